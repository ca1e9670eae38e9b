"""Stock-PyTorch comparator of the ST-GCN block stack  --  TEST / BENCH INFRASTRUCTURE ONLY.

The op sequence of the reference, restated with the same stock torch modules the reference instantiates, so that
`bench.py --torch-gpu-baseline` can time "what the reference does" (cuDNN / cuBLAS / ATen kernels, eager) on the same B200
as the product:

    SpatialTemporalConv.forward  /root/reference/src/network/st_graphconv.py:85-109
        nn.BatchNorm2d (:34) -> SpatialConv (:139-152: nn.Conv2d 1x1 (:136), view (:149), einsum 'kvw,nkctw->nctv' (:150))
        -> nn.Conv2d (9,1) stride (s,1) pad (4,0) (:41-43) -> nn.BatchNorm2d (:46) -> ReLU on a clone (:105)
        [residual variant :60-82]
    L_STGCN.forward              /root/reference/src/lightning_model.py:91-112  (permute, 10 blocks, avg_pool2d, Linear)

The reference itself cannot travel to the GPU box (/root/reference does not exist there); tests/test_oracle_golden.py pins
THIS module's forward against the fixtures generated from the unmodified reference.  Nothing in stgcn_b200/ imports it; its users are tests/test_oracle_golden.py (the pin), bench.py (`--impl reference`,
`cpu_baseline`, `torch_gpu_baseline`) and tools/bench_block.py (the same comparator per block, BASELINE cfg2).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

LAYERS = [(None, 64, 1), (64, 64, 1), (64, 64, 1), (64, 64, 1), (64, 128, 2), (128, 128, 1), (128, 128, 1),
          (128, 256, 2), (256, 256, 1), (256, 256, 1)]


class StockSpatialConv(nn.Module):
    def __init__(self, C_in, C_out, A):
        super().__init__()
        self.A = nn.Parameter(A.float())
        self.K, self.V, self.C_out = A.shape[0], A.shape[1], C_out
        self.W = nn.Conv2d(C_in, self.K * C_out, (1, 1))

    def forward(self, f_in):
        N, _, T, _ = f_in.shape
        f = self.W(f_in).view(N, self.K, self.C_out, T, self.V)
        return torch.einsum('kvw,nkctw->nctv', self.A.to(f.dtype), f)


class StockBlock(nn.Module):
    def __init__(self, C_in, C_out, A, gamma, stride, pad, dropout_rate=0.0, residual=False):
        super().__init__()
        self.residual = residual
        self.res_conv = None
        if residual and not (C_in == C_out and stride == 1):
            self.res_conv = nn.Conv2d(C_in, C_out, kernel_size=1, stride=(stride, 1))
        self.batch_n = nn.BatchNorm2d(C_in)
        self.spatialConv = StockSpatialConv(C_in, C_out, A)
        self.temporalConv = nn.Conv2d(C_out, C_out, kernel_size=(gamma, 1), stride=(stride, 1), padding=(pad, 0))
        self.batch_n_2 = nn.BatchNorm2d(C_out)
        self.dropout = nn.Dropout(dropout_rate, inplace=True) if dropout_rate != 0 else None

    def forward(self, f_in):
        if self.residual:
            res = f_in.clone()
            f = F.relu(self.batch_n(f_in).clone())
            f = self.spatialConv(f)
            f = F.relu(self.batch_n_2(f).clone())
            f = self.temporalConv(f)
            f = f + (self.res_conv(res) if self.res_conv is not None else res)
        else:
            f = self.batch_n_2(self.temporalConv(self.spatialConv(self.batch_n(f_in))))
        f = F.relu(f.clone())
        if self.dropout is not None:
            f = self.dropout(f.clone())
        return f


class StockSTGCN(nn.Module):
    def __init__(self, C_in, A, classes, gamma=9, dropout_rate=0.0, residual=False):
        super().__init__()
        pad = (gamma - 1) // 2
        self.conv = nn.Sequential(*[StockBlock(C_in if ci is None else ci, co, A, gamma, s, pad, dropout_rate, residual)
                                    for ci, co, s in LAYERS])
        self.fc_layer = nn.Linear(256, classes)

    def forward(self, x):                       # (N,T,V,C) like lightning_model.py:96
        x = x.permute(0, 3, 1, 2)
        x = self.conv(x)
        x = F.avg_pool2d(x, (x.shape[2], x.shape[3])).view(x.shape[0], -1)
        return self.fc_layer(x)
