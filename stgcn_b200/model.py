"""Drop-in 10-block ST-GCN models: L_STGCN (src/lightning_model.py:30-112) and
STGCN (src/network/stgcn.py:7-80).  Same forward API -- x (N,T,V,C_in) ->
(N, nr_classes) -- same parameter names/shapes; pytorch-lightning is not
required (plain nn.Module; `training_step`-style helpers live in bench.py).
"""
import torch
import torch.nn as nn

from . import functional as Fn
from . import graph
from .modules import SpatialTemporalConv, prepare_training_operands

LAYERS = [(None, 64, 1), (64, 64, 1), (64, 64, 1), (64, 64, 1), (64, 128, 2), (128, 128, 1), (128, 128, 1),
          (128, 256, 2), (256, 256, 1), (256, 256, 1)]          # lightning_model.py:65-86


def _prepare(conv, x):
    """Before a forward that will be differentiated: the weight-only operands of all ten blocks in one launch."""
    if torch.is_grad_enabled() and x.is_cuda and any(p.requires_grad for p in conv[0].parameters()):
        prepare_training_operands(list(conv), x.shape[0], x.shape[2], x.device)


def _stack(c_in, A, masks, gamma, pad, compute_dtype, **kw):
    blocks = []
    for i, (ci, co, s) in enumerate(LAYERS):
        blocks.append(SpatialTemporalConv(c_in if ci is None else ci, co, A * masks[i], gamma, s, pad,
                                          compute_dtype=compute_dtype, **kw))
    return nn.Sequential(*blocks)


class L_STGCN(nn.Module):
    def __init__(self, hparams, dataset_filters=None, compute_dtype=torch.float32, A=None):
        super().__init__()
        self.hparams = hparams
        self.dataset_filters = dataset_filters
        temporal_padding = (hparams.gamma - 1) // 2
        if A is None:
            A = graph.get_normalized_adjacency_matrices(hparams.partitioning, hparams.d,
                                                        distance_file=getattr(hparams, 'distance_file', None) or None,
                                                        distances=getattr(hparams, 'distances', None),
                                                        skeleton=getattr(hparams, 'skeleton', 'body25'))
        self.K, self.V = A.shape[0], A.shape[1]
        self.nr_classes = hparams.nr_classes
        self.n_layers = 10
        if hparams.use_edge_importance:
            jit = [2 * (torch.randn_like(A) - 0.5) * hparams.max_mask_jitter for _ in range(self.n_layers)]
            self.Masks = nn.ParameterList([nn.Parameter(jit[i] + torch.ones(A.shape)) for i in range(self.n_layers)])
        else:
            self.Masks = [torch.ones(A.shape) for _ in range(self.n_layers)]
        self.conv = _stack(hparams.C_in, A, self.Masks, hparams.gamma, temporal_padding, compute_dtype,
                           dropout_rate=hparams.dropout_rate, residual=hparams.residual).float()
        self.fc_layer = nn.Linear(256, hparams.nr_classes).float()
        self.softmax = nn.Softmax(dim=1)
        self.compute_dtype = compute_dtype
        # Beyond the reference (its own TODO at lightning_model.py:52 / st_graphconv.py:129), opt-in with
        # hparams.edge_importance_mode = 'learnable': the masks are what is learned -- every forward uses A_base * Masks[i] as the
        # block's adjacency, Masks[i] receive dA (.) A_base, and spatialConv.A is a frozen base.  Default 'reference': A * Mask is
        # baked into a detached trainable spatialConv.A at construction and the masks never get a gradient, as in the reference.
        self.edge_importance_mode = getattr(hparams, 'edge_importance_mode', 'reference')
        if self.edge_importance_mode == 'learnable':
            if not hparams.use_edge_importance:
                raise ValueError("edge_importance_mode='learnable' needs use_edge_importance=True")
            with torch.no_grad():
                for b in self.conv:
                    b.spatialConv.A.copy_(A.float())          # the base graph, without the mask baked in
                    b.spatialConv.A.requires_grad_(False)
        elif self.edge_importance_mode != 'reference':
            raise ValueError('edge_importance_mode must be "reference" or "learnable"')

    def forward(self, x):
        x = x.permute(0, 3, 1, 2)                 # (N,T,V,C) -> logical (N,C,T,V): already channels-last in memory
        if self.edge_importance_mode == 'learnable':
            for b, mk in zip(self.conv, self.Masks):
                b._A_override = b.spatialConv.A.detach() * mk        # (K,V,V) element-wise: the only torch op, outside the chain
        _prepare(self.conv, x)
        x = self.conv(x)
        return Fn.head_forward(x, self.fc_layer.weight, self.fc_layer.bias)

    def compute_accuracy(self, pred, gt):
        pred = torch.argmax(self.softmax(pred), axis=1)
        return (pred == gt).sum().item() / len(pred)


class STGCN(nn.Module):
    def __init__(self, C_in, gamma, nr_classes, strat=graph.Strategy.UNI_LABELING, d=1, edge_importance=True,
                 compute_dtype=torch.float32, A=None):
        super().__init__()
        self.nr_classes = nr_classes
        temporal_padding = (gamma - 1) // 2
        if A is None:
            A = graph.get_normalized_adjacency_matrices(strat, d)
        self.K, self.V = A.shape[0], A.shape[1]
        self.C_in, self.C_out = C_in, 256
        if edge_importance:
            self.Masks = nn.ParameterList([nn.Parameter(torch.ones(A.shape)) for _ in range(10)])
        else:
            self.Masks = [torch.ones(A.shape) for _ in range(10)]
        self.conv = _stack(C_in, A, self.Masks, gamma, temporal_padding, compute_dtype).float()
        self.fc_layer = nn.Linear(256, nr_classes).float()
        self.softmax = nn.Softmax(dim=1)

    def forward(self, x):
        x = x.permute(0, 3, 1, 2)
        _prepare(self.conv, x)
        x = self.conv(x)
        x = Fn.head_forward(x, self.fc_layer.weight, self.fc_layer.bias)
        return self.softmax(x)
