#!/usr/bin/env python
"""bench.py -- headline benchmark of the ST-GCN block stack on B200 (contract in the task brief).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg4|cfg5|cfg3]
                    [--dtype bf16|f32] [--seqs-per-gpu S]

Workloads (BASELINE.json `configs`):
  cfg4 (default, the configuration the metric is quoted on): 10-layer ST-GCN (64x4,128x3,256x3), V=25, K=3, 60 classes,
        bf16 TRAINING step = forward + cross-entropy + backward + Adam, 32 sequences x 3x300x25x2 per GPU (persons folded into
        the batch: N_eff = 64; global batch 256 at 8 GPUs).  Weak scaling; gradients all-reduced with NCCL in buckets
        overlapped with backward.
  cfg5: Kinetics-skeleton shape, V=18 (OpenPose-18 graph), 400 classes, 128 sequences per GPU (N=1024 at 8 GPUs), training;
        its inference number rides along in `infer`.
  cfg3: INFERENCE, N=64 sequences (N_eff=128), V=25, 60 classes, 1 GPU, bf16 or fp32 (--dtype).
`value` = sequences/s of the whole job with inputs resident in HBM, L2 flushed between timed steps; `e2e` = the same through
the public module API with pinned-host inputs copied in and the result read back every step.  `roofline` = the dominant
kernel on ALGORITHMIC FLOPs; `roofline_step` = the whole step; `torch_gpu_baseline` = the reference's stock-module op
sequence (oracle/torch_stock.py) timed on the same GPU; `cpu_baseline` = the same stock-module sequence on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

T_FRAMES, C_IN, M_PERSONS = 300, 3, 2
LAYERS = [(3, 64, 1), (64, 64, 1), (64, 64, 1), (64, 64, 1), (64, 128, 2), (128, 128, 1), (128, 128, 1),
          (128, 256, 2), (256, 256, 1), (256, 256, 1)]
WORKLOADS = {
    'cfg4': dict(v=25, k=3, classes=60, seqs=32, mode='train',
                 name='BASELINE cfg4: 10-layer ST-GCN (64x4,128x3,256x3) training, V=25 K=3 60 classes',
                 metric='ST-GCN training sequences/s (3x300x25x2, 10-layer, fwd+CE+bwd+Adam)'),
    'cfg5': dict(v=18, k=3, classes=400, seqs=128, mode='train',
                 name='BASELINE cfg5: Kinetics-skeleton 10-layer ST-GCN training, V=18 (OpenPose-18 graph) K=3 400 classes',
                 metric='ST-GCN training sequences/s (3x300x18x2, 10-layer, fwd+CE+bwd+Adam)'),
    'cfg3': dict(v=25, k=3, classes=60, seqs=64, mode='infer',
                 name='BASELINE cfg3: 10-layer ST-GCN (64x4,128x3,256x3) inference, N=64 V=25 K=3 60 classes',
                 metric='ST-GCN inference sequences/s (3x300x25x2, 10-layer, eval forward)'),
}


def flops_forward_per_neff(v, k, classes, t=T_FRAMES):
    """Algorithmic forward FLOPs per N_eff unit; (total, dense GEMM part)."""
    total = gemm = 0.0
    for ci, co, s in LAYERS:
        to = (t - 1) // s + 1
        g1 = 2.0 * t * v * ci * k * co
        agg = 2.0 * k * co * t * v * v
        g2 = 2.0 * to * v * 9 * co * co
        total += g1 + agg + g2
        gemm += g1 + g2
        t = to
    return total + 2.0 * 256 * classes, gemm


def fused_min_bytes_per_neff(v, train, t=T_FRAMES, esize=2):
    """Fused-minimum activation traffic per N_eff (SURVEY.md section 8(d) 'Algorithmic bytes'): forward reads each block input and
    writes each block output once; training adds one re-read of every block output (batch statistics) and, in backward, the
    reads of x and dy and the write of dx per block."""
    fwd = bwd = 0.0
    for ci, co, s in LAYERS:
        to = (t - 1) // s + 1
        xin, yout = t * v * ci * esize, to * v * co * esize
        fwd += xin + yout
        bwd += yout + xin + yout + xin          # stats re-read of y; backward: read x, read dy, write dx
        t = to
    return fwd + (bwd if train else 0.0)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, index):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q,
                                       '--format=csv,noheader,nounits', '-lms', '20'], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(',')])

    def stop(self):
        if self.p is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.p.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == 'Active' for r in self.rows)]
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None, 'reasons': reasons,
                'samples': len(sm)}


def synthetic_batch(wl, seqs, seed, device=None, pin=False):
    import torch
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(seqs * M_PERSONS, T_FRAMES, wl['v'], C_IN, generator=g)        # persons folded into the batch
    y = torch.randint(0, wl['classes'], (seqs * M_PERSONS,), generator=g)
    if pin:
        return x.pin_memory(), y.pin_memory()
    return (x.to(device), y.to(device)) if device is not None else (x, y)


def hparams(wl):
    return argparse.Namespace(C_in=C_IN, nr_classes=wl['classes'], gamma=9, dropout_rate=0.0, residual=False,
                              use_edge_importance=False, max_mask_jitter=0.001, partitioning=2, d=1, distance_file=None,
                              distances=None)


def adjacency(wl):
    """V=25: dense O(1) learned-adjacency stand-in (SURVEY 8(d) input (ii); the block treats A as an arbitrary dense learned
    operand).  V=18: the OpenPose-18 graph of stgcn_b200.graph, spatial-configuration partitions, D^-1/2 A D^-1/2."""
    import numpy as np
    import torch
    if wl['v'] == 25:
        return torch.randn(wl['k'], 25, 25, generator=torch.Generator().manual_seed(0)) * 0.2
    from stgcn_b200 import graph
    A = graph.get_normalized_adjacency_matrices(graph.Strategy.SPATIAL_CONFIGURATION, 1, distances=np.random.RandomState(0).rand(wl['v']),
                                                skeleton='openpose18', dense_alpha=False)
    return A / A.abs().max()


def stock_model(wl, device, A=None):
    """The reference's op sequence as stock torch modules (oracle/torch_stock.py), default init under seed 0."""
    import torch
    from oracle.torch_stock import StockSTGCN
    torch.manual_seed(0)
    return StockSTGCN(C_IN, adjacency(wl) if A is None else A, wl['classes']).to(device)


def run_reference(args, wl):
    """CPU arm: the reference's path = its stock torch modules (nn.BatchNorm2d / nn.Conv2d / einsum, mkldnn kernels) as restated in
    oracle/torch_stock.py (the reference itself does not travel to the GPU box), all host threads, a bounded sample of the same
    workload: BASELINE cfg1's batch, 8 sequences (N_eff = 16) of the workload's shape per step."""
    import torch
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    train = wl['mode'] == 'train'
    model = stock_model(wl, 'cpu').train(train)
    opt = torch.optim.Adam(model.parameters(), lr=0.01)
    seqs = 8
    x, y = synthetic_batch(wl, seqs, 0)

    def step():
        if not train:
            with torch.no_grad():
                return float(model(x).sum())
        opt.zero_grad(set_to_none=True)
        loss = torch.nn.functional.cross_entropy(model(x), y)
        loss.backward()
        opt.step()
        return float(loss.detach())
    if not train:                         # running statistics off their initial values first
        model.train()
        with torch.no_grad():
            model(x)
        model.eval()
    warm = max(1, min(args.warmup, 5))         # ~0.7 s per step on 16 cores: the whole arm stays well under a minute
    for _ in range(warm):
        step()
    steps = max(1, min(args.steps, 20))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    val = seqs / dt
    what = 'fwd+CE+bwd+Adam' if train else 'eval forward'
    sample = '%d sequences (N_eff=%d) x 3x300x%d, %s fp32, %d steps' % (seqs, seqs * 2, wl['v'], what, steps)
    print(json.dumps({'impl': 'reference', 'metric': wl['metric'], 'value': val, 'unit': 'seq/s',
                      'n_gpus': args.gpus, 'steps': steps, 'warmup': warm, 'ms_per_step': dt * 1e3,
                      'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
                      'config': {'workload': wl['name'], 'seqs_per_step': seqs, 'T': T_FRAMES,
                                 'note': 'reference arm: the reference\'s stock torch modules (oracle/torch_stock.py, pinned to the '
                                         'reference by tests/test_oracle_golden.py) on the host cores; bounded sample = BASELINE '
                                         'cfg1 batch (N_eff=16) of the same workload'},
                      'cpu_baseline': {'value': val, 'unit': 'seq/s', 'cores': cores, 'kind': 'port', 'sample': sample},
                      'e2e': {'value': val, 'unit': 'seq/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}))


def torch_gpu_baseline(wl, seqs, dev, steps, flush, A):
    """GPU-vs-GPU comparator: the reference's stock-module op sequence (cuDNN / cuBLAS / ATen, eager) on this same B200, same
    shape and step count, in fp32 (torch defaults) and under bf16 autocast; L2 flushed between steps like the product."""
    import torch
    train = wl['mode'] == 'train'
    out = {}
    x, y = synthetic_batch(wl, seqs, 100, dev)
    for tag, autocast in (('eager_fp32', False), ('autocast_bf16', True)):
        try:
            model = stock_model(wl, dev, A).train()
            opt = torch.optim.Adam(model.parameters(), lr=0.01)

            def step():
                with torch.autocast('cuda', dtype=torch.bfloat16, enabled=autocast):
                    if not train:
                        with torch.no_grad():
                            return model(x)
                    opt.zero_grad(set_to_none=True)
                    loss = torch.nn.functional.cross_entropy(model(x).float(), y)
                loss.backward()
                opt.step()
                return loss
            for _ in range(3):
                step()
            if not train:
                model.eval()
                step()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                flush.zero_(); step()
            e1.record(); torch.cuda.synchronize()
            ms_all = e0.elapsed_time(e1)
            e0.record()
            for _ in range(steps):
                flush.zero_()
            e1.record(); torch.cuda.synchronize()
            ms = (ms_all - e0.elapsed_time(e1)) / steps
            out[tag] = {'value': seqs / (ms * 1e-3), 'unit': 'seq/s', 'ms_per_step': ms}
            del model, opt
            torch.cuda.empty_cache()
        except Exception as e:          # e.g. out of memory at a large shape: report, do not fail the bench
            out[tag] = {'error': str(e)[:200]}
    out['what'] = 'oracle/torch_stock.py: nn.BatchNorm2d -> nn.Conv2d(1x1) -> einsum -> nn.Conv2d(9x1) -> BN -> ReLU x10, ' \
                  'avg_pool2d, Linear; torch %s eager; %s' % (torch.__version__, 'fwd+CE+bwd+torch.optim.Adam' if train else 'eval forward')
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default='cfg4', choices=sorted(WORKLOADS))
    ap.add_argument('--seqs-per-gpu', type=int, default=0, help='override the workload\'s sequences per GPU')
    ap.add_argument('--dtype', default='bf16', choices=['bf16', 'f32'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-torch-baseline', action='store_true')
    ap.add_argument('--eager', action='store_true', help='A/B aid: launch the step kernel by kernel instead of replaying its CUDA graph')
    ap.add_argument('--only-train', action='store_true', help='profiling aid: skip the e2e, inference and baseline legs')
    ap.add_argument('--only-infer', action='store_true', help='profiling aid: time the eval-mode forward only')
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == 'reference':
        return run_reference(args, wl)

    import torch
    import torch.distributed as dist
    from stgcn_b200 import CapturedForward, CapturedTrainStep, L_STGCN, _lib, cross_entropy
    from stgcn_b200.optim import Adam
    from stgcn_b200.parallel import GradBucketReducer
    from stgcn_b200.pipeline import HostBatchPrefetcher
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    W = max(3, args.warmup)
    S = args.seqs_per_gpu or wl['seqs']
    infer_only = wl['mode'] == 'infer' or args.only_infer
    dt = torch.bfloat16 if args.dtype == 'bf16' else torch.float32
    torch.manual_seed(0)
    A = adjacency(wl)
    model = L_STGCN(hparams(wl), compute_dtype=dt, A=A).to(dev).train()
    opt = Adam(model.parameters(), lr=0.01)                 # one launch per 64 parameter tensors (stgcn_adam_step)
    reducer = GradBucketReducer(model.parameters(), 3) if world > 1 else None
    n_batches = 4 if S <= 64 else 2                         # distinct resident batches
    dev_batches = [synthetic_batch(wl, S, 100 + rank * 10 + i, dev) for i in range(n_batches)]
    flush = torch.empty(160 * 1024 * 1024, dtype=torch.uint8, device=dev)      # > 126 MB L2
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    peak_tf = peaks.get('bf16_tflops_sustained', 1400.0)
    peak_src = 'MEASURED_PEAKS.json bf16_tflops_sustained' if peaks else 'fallback 1.4 PFLOP/s (sustained)'
    fwd_flops, gemm_flops = flops_forward_per_neff(wl['v'], wl['k'], wl['classes'])
    n_eff = S * M_PERSONS

    def train_step(x, y):
        opt.zero_grad(set_to_none=True)
        loss = cross_entropy(model(x), y)
        loss.backward()
        if reducer is not None:
            reducer.finish()
        opt.step()
        return loss

    cap = [None]

    def leave():
        """End of a rank: destroy the captured graph first (it references NCCL work and the scratch arena), then the process
        group.  A watchdog ends the process if NCCL teardown does not return (every result has been printed by then)."""
        sys.stdout.flush(); sys.stderr.flush()
        if world > 1:
            torch.cuda.synchronize()
            if cap[0] is not None:
                cap[0].close()
            t = threading.Timer(30.0, lambda: os._exit(0))
            t.daemon = True
            t.start()
            dist.destroy_process_group()
            t.cancel()

    def step_fn(x, y):
        """The public training step: the captured CUDA graph of train_step (stgcn_b200.CapturedTrainStep), or train_step itself."""
        return cap[0](x, y) if cap[0] is not None else train_step(x, y)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_ms = [0.0]

    def timed(fn, steps):
        """K steps bracketed by barrier+sync, device-timed with CUDA events; max over ranks.  host_ms[0] = wall time the host
        needed to ENQUEUE the K steps (when it is close to the device time, the step is launch-bound)."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        for i in range(steps):
            fn(i)
        host_ms[0] = (time.perf_counter() - t0) * 1e3
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    def roofline_of(prof, cls, kernel):
        um = prof.get(cls)
        if not um or um[0] <= 0:
            return None
        ach = um[1] / (um[0] * 1e-3) / 1e12
        traffic, traffic_src = None, None
        try:    # DRAM bytes per launch of the same kernel from the committed ncu capture (dram__bytes_read.sum + dram__bytes_write.sum)
            tj = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json')))[kernel]
            traffic, traffic_src = tj['avg_dram_bytes_per_launch'], tj['source']
        except Exception:
            pass
        return {'bound': 'tensor', 'kernel': kernel, 'achieved': ach, 'peak': peak_tf, 'unit': 'TFLOP/s', 'frac': ach / peak_tf,
                'traffic': traffic, 'traffic_unit': 'DRAM bytes per launch (ncu)', 'traffic_source': traffic_src,
                'flop_per_launch': um[1] / max(um[2], 1), 'launches_timed': um[2], 'ms_total': um[0],
                'flops': 'algorithmic (2 x rows x taps x C x N of the forward contraction; padded channels and unused dgrad taps not counted)',
                'peak_source': peak_src}

    def measure_inference():
        """Eval-mode forward at this workload's batch: L2 flushed between steps (flush time subtracted), device-timed; also the
        same through pinned-host inputs with the logits read back (e2e)."""
        model.eval()
        res = {}
        with torch.no_grad():
            for i in range(W):
                model(dev_batches[i % n_batches][0])
            if args.eager:
                fn = model
            else:
                fn = CapturedForward(model, dev_batches[0][0])          # the eval forward as one CUDA graph
                for i in range(W):
                    fn(dev_batches[i % n_batches][0])
            ms_t = timed(lambda i: (flush.zero_(), fn(dev_batches[i % n_batches][0])), args.steps)
            res['host_enqueue_ms_per_step'] = host_ms[0] / args.steps
            ms_f = timed(lambda i: flush.zero_(), args.steps)
            ms = (ms_t - ms_f) / args.steps
            # e2e: pinned host batch -> device, forward, logits -> pinned host, every step
            hx = [b[0].cpu().pin_memory() for b in dev_batches]
            out_host = torch.empty((S * M_PERSONS, wl['classes']), dtype=torch.float32).pin_memory()
            xin = torch.empty_like(dev_batches[0][0])

            def e2e(i):
                xin.copy_(hx[i % n_batches], non_blocking=True)
                out_host.copy_(fn(xin), non_blocking=True)
            for i in range(2):
                e2e(i)
            ms_e_seq = timed(e2e, args.steps) / args.steps
            ms_e = ms_e_seq
            if not args.eager:
                # the serving loop of the public API (CapturedForward.stream): the copy of batch i+1 overlaps the forward of batch i;
                # every step still copies its own input from pinned host memory and reads its logits back, inside the timed region
                def piped(total):
                    gen = fn.stream(hx[i % n_batches] for i in range(total))
                    return lambda i: next(gen, None)
                timed(piped(4), 4)
                ms_e = timed(piped(args.steps), args.steps) / args.steps
            _lib.timing_reset(); _lib.timing_enable(True)
            timed(lambda i: (flush.zero_(), model(dev_batches[i % n_batches][0])), args.steps)
            _lib.timing_enable(False)
        prof = _lib.timing_read()
        model.train()
        tf = fwd_flops * n_eff * world / (ms * 1e-3) / 1e12
        res.update({'value': S * world / (ms * 1e-3), 'unit': 'seq/s', 'ms_per_step': ms, 'seqs_per_gpu': S,
                    'mode': 'eager launches' if args.eager else 'one CUDA graph per forward (CapturedForward)',
                    'l2': 'flushed (160 MB write) between timed steps, flush time subtracted',
                    'model_tflops': tf,
                    'roofline_step': {'bound': 'tensor', 'achieved': tf / world, 'peak': peak_tf, 'unit': 'TFLOP/s per GPU',
                                      'frac': tf / world / peak_tf, 'peak_source': peak_src},
                    'roofline': roofline_of(prof, 'umma_conv_gemm', 'conv_gemm_umma_kernel') or roofline_of(prof, 'simt_conv_gemm', 'conv_gemm_simt_kernel'),
                    'e2e': {'value': S * world / (ms_e * 1e-3), 'unit': 'seq/s', 'ms_per_step': ms_e,
                            'h2d_bytes_per_step': hx[0].numel() * hx[0].element_size(), 'd2h_bytes_per_step': out_host.numel() * 4,
                            'mode': 'CapturedForward.stream (input copy of the next batch overlaps the forward)' if not args.eager else 'copy, forward, read back on one stream',
                            'sequential_ms_per_step': ms_e_seq},
                    'kernel_ms_per_step': {k: round(v[0] / args.steps, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}})
        return res

    def common_fields(ms_step, value, metric):
        return {'metric': metric, 'value': value, 'unit': 'seq/s', 'n_gpus': world, 'steps': args.steps, 'warmup': W,
                'ms_per_step': ms_step, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': args.dtype,
                'data': 'synthetic',
                'config': {'workload': wl['name'], 'seqs_per_gpu': S, 'n_eff_per_gpu': n_eff, 'global_batch_seqs': S * world,
                           'T': T_FRAMES, 'V': wl['v'], 'K': wl['k'], 'classes': wl['classes'], 'parallelism': 'dp%d' % world,
                           'l2': 'flushed (160 MB write) between timed steps, flush time subtracted',
                           'bn': 'per-replica statistics'}}

    def add_baselines(out):
        if world != 1:
            return
        if not args.no_torch_baseline:
            out['torch_gpu_baseline'] = torch_gpu_baseline(wl, S, dev, args.steps, flush, A)
            for tag in ('eager_fp32', 'autocast_bf16'):
                b = out['torch_gpu_baseline'].get(tag, {})
                if 'value' in b:
                    b['speedup_of_this_repo'] = out['value'] / b['value']
        if not args.no_cpu_baseline:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), '--impl', 'reference', '--workload', args.workload,
                                '--steps', '6', '--warmup', '1'], capture_output=True, text=True)
            try:
                out['cpu_baseline'] = json.loads(r.stdout.strip().splitlines()[-1])['cpu_baseline']
            except Exception:
                out['cpu_baseline'] = {'error': (r.stderr or r.stdout)[-300:]}

    # ---- inference workloads ---------------------------------------------------------------------------------------------
    if infer_only:
        for i in range(2):
            train_step(*dev_batches[i % n_batches])          # move the BatchNorm running statistics off their initial values
        sampler = ClockSampler(local) if rank == 0 else None
        l0 = _lib.launch_count()
        inf = measure_inference()
        launches_eager_pass = _lib.launch_count() - l0
        clocks = sampler.stop() if sampler else None
        if rank != 0:
            return leave()
        out = common_fields(inf['ms_per_step'], inf['value'], wl['metric'] if wl['mode'] == 'infer' else WORKLOADS['cfg3']['metric'])
        out.update({'model_tflops': inf['model_tflops'], 'e2e': inf['e2e'], 'roofline': inf['roofline'],
                    'roofline_step': inf['roofline_step'], 'gpu_launches': launches_eager_pass,
                    'gpu_launches_note': 'library launches counted over the warm-up, capture and instrumented eager passes of this leg',
                    'step_mode': inf['mode'], 'host_enqueue_ms_per_step': inf['host_enqueue_ms_per_step'], 'clocks': clocks,
                    'kernel_ms_per_step': inf['kernel_ms_per_step']})
        add_baselines(out)
        print(json.dumps(out))
        return leave()

    # ---- device-resident training ------------------------------------------------------------------------------------------
    for i in range(W):
        train_step(*dev_batches[i % n_batches])
    l0 = _lib.launch_count()
    if not args.eager:
        cap[0] = CapturedTrainStep(model, opt, *dev_batches[0], reducer=reducer)      # warm-up steps inside are undone
    launches_per_step = (_lib.launch_count() - l0) // 3 if cap[0] is not None else None  # 2 warm-up steps + the captured one
    for i in range(W):
        step_fn(*dev_batches[i % n_batches])
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = _lib.launch_count()
    ms_train = timed(lambda i: (flush.zero_(), step_fn(*dev_batches[i % n_batches])), args.steps)
    host_enqueue_ms = host_ms[0] / args.steps
    # a replayed graph launches the captured kernels without passing through the library's host counter
    launches = launches_per_step * args.steps if cap[0] is not None else _lib.launch_count() - l0
    # the L2 flush is inside the timed region: subtract its own measured cost
    ms_flush = timed(lambda i: flush.zero_(), args.steps)
    ms_step = (ms_train - ms_flush) / args.steps
    value = S * world / (ms_step * 1e-3)
    # second pass over the same K steps with a CUDA-event pair around every launch of the library (on the launching
    # stream): per-kernel-class durations for the roofline line.  The events cost a little, so `value` is the pass above.
    _lib.timing_reset(); _lib.timing_enable(True)
    ms_instr = (timed(lambda i: (flush.zero_(), train_step(*dev_batches[i % n_batches])), args.steps) - ms_flush) / args.steps
    _lib.timing_enable(False)
    prof = _lib.timing_read()
    clocks = sampler.stop() if sampler else None      # sampled every 20 ms across the timed pass and the instrumented pass
    kernel_ms = {k: round(v[0] / args.steps, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}
    if args.only_train:
        if rank == 0:
            print(json.dumps({'value': value, 'ms_per_step': ms_step, 'ms_per_step_instrumented': ms_instr, 'gpu_launches': launches,
                              'host_enqueue_ms_per_step': host_enqueue_ms, 'step_mode': 'eager' if args.eager else 'cuda_graph',
                              'kernel_ms_per_step': kernel_ms}))
        return leave()

    # ---- end to end: pinned host batch -> H2D -> step -> loss D2H ------------------------
    # public input pipeline (stgcn_b200.pipeline): the pinned batch of step i+1 is copied on a side stream while step i runs;
    # every step still moves its own inputs host -> device inside the timed region and reads its loss back
    host_batches = [synthetic_batch(wl, S, 100 + rank * 10 + i, pin=True) for i in range(n_batches)]
    pf = HostBatchPrefetcher(dev)
    e2e_k = [0]
    loss_host = [torch.zeros(1).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]

    def e2e_step(i):
        k = e2e_k[0]
        if pf.count == 0:
            pf.push(host_batches[k % n_batches])
        pf.push(host_batches[(k + 1) % n_batches])
        xin, yin = pf.pop()
        loss = step_fn(xin, yin)
        pf.done()
        e2e_k[0] = k + 1
        # device -> host read of every step's loss through pinned memory; the host blocks on the PREVIOUS step's copy (deferred
        # logging), so it can queue step i+1 while step i runs, and on the last step also on its own
        slot = k & 1
        loss_host[slot].copy_(loss.detach().reshape(1), non_blocking=True)
        loss_ev[slot].record()
        val = None
        if k > 0:
            loss_ev[slot ^ 1].synchronize()
            val = float(loss_host[slot ^ 1])
        if i == args.steps - 1:
            loss_ev[slot].synchronize()
            val = float(loss_host[slot])
        return val
    for i in range(2):
        e2e_step(i)
    ms_e2e = timed(e2e_step, args.steps) / args.steps
    h2d = sum(t.numel() * t.element_size() for t in host_batches[0])

    # ---- inference at this workload's shape rides along (cfg3 proper: --workload cfg3) ------------------------------------
    inf = measure_inference()
    if rank != 0:
        return leave()
    train_flops = 3 * fwd_flops * n_eff * world
    model_tf = train_flops / (ms_step * 1e-3) / 1e12
    step_bytes, step_bytes_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json')))['training_step_' + args.workload]
        step_bytes, step_bytes_src = tj['dram_bytes_per_step'], tj['source']
    except Exception:
        pass
    min_bytes = fused_min_bytes_per_neff(wl['v'], True) * n_eff
    out = common_fields(ms_step, value, wl['metric'])
    out.update({'model_tflops': model_tf,
                'e2e': {'value': S * world / (ms_e2e * 1e-3), 'unit': 'seq/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 4,
                        'ms_per_step': ms_e2e},
                'roofline': roofline_of(prof, 'umma_conv_gemm', 'conv_gemm_umma_kernel'),
                'roofline_step': {'bound': 'tensor', 'achieved': model_tf / world, 'peak': peak_tf, 'unit': 'TFLOP/s per GPU',
                                  'frac': model_tf / world / peak_tf, 'algorithmic_flops_per_step_per_gpu': train_flops / world,
                                  'hbm_bytes_per_step': step_bytes, 'hbm_bytes_source': step_bytes_src,
                                  'hbm_bytes_fused_minimum': min_bytes,
                                  'hbm_over_minimum': (step_bytes / min_bytes) if step_bytes else None, 'peak_source': peak_src},
                'infer': {k: inf[k] for k in ('value', 'unit', 'ms_per_step', 'seqs_per_gpu', 'mode', 'l2', 'model_tflops',
                                              'roofline_step', 'host_enqueue_ms_per_step')},
                'gpu_launches': launches, 'step_mode': 'eager launches' if args.eager else 'one CUDA graph per step (CapturedTrainStep)',
                'host_enqueue_ms_per_step': host_enqueue_ms, 'clocks': clocks, 'ms_per_step_instrumented': ms_instr,
                'kernel_ms_per_step': kernel_ms})
    out['infer']['e2e'] = inf['e2e']
    add_baselines(out)
    print(json.dumps(out))
    leave()


if __name__ == '__main__':
    main()
