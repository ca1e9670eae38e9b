#!/usr/bin/env python
"""bench.py -- headline benchmark of the ST-GCN block stack on B200 (contract in the task brief).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--seqs-per-gpu S]

Workload (BASELINE.json configs[3], cfg4): the 10-layer ST-GCN (64x4,128x3,256x3), V=25, K=3 (spatial
partitioning), 60 classes, bf16 TRAINING step = forward + cross-entropy + backward + Adam, on synthetic
NTU-shaped skeletons (S sequences x 3 x 300 x 25 x 2 persons per GPU; persons folded into the batch, so
N_eff = 2*S).  Weak scaling: S per GPU is fixed (default 32 -> global batch 256 at 8 GPUs as cfg4 says);
gradients are all-reduced with NCCL in buckets overlapped with backward.  `value` = sequences/s of the
whole job with inputs resident in HBM; `e2e` = the same through the public module API with pinned-host
inputs copied in and the loss read back every step.  An inference line (cfg3) rides along in `infer`.
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

T_FRAMES, V_JOINTS, C_IN, M_PERSONS, CLASSES, K_PART = 300, 25, 3, 2, 60, 3
METRIC = 'ST-GCN training sequences/s (3x300x25x2, 10-layer, fwd+CE+bwd+Adam)'
WORKLOAD = 'BASELINE cfg4: 10-layer ST-GCN (64x4,128x3,256x3) training, V=25 K=3 60 classes'
LAYERS = [(3, 64, 1), (64, 64, 1), (64, 64, 1), (64, 64, 1), (64, 128, 2), (128, 128, 1), (128, 128, 1),
          (128, 256, 2), (256, 256, 1), (256, 256, 1)]


def flops_forward_per_neff(v=V_JOINTS, k=K_PART, t=T_FRAMES, classes=CLASSES):
    """Algorithmic forward FLOPs per N_eff unit; tconv part returned separately."""
    total = tconv = 0.0
    for ci, co, s in LAYERS:
        to = (t - 1) // s + 1
        g1 = 2.0 * t * v * ci * k * co
        agg = 2.0 * k * co * t * v * v
        g2 = 2.0 * to * v * 9 * co * co
        total += g1 + agg + g2
        tconv += g2
        t = to
    return total + 2.0 * 256 * classes, tconv


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


def synthetic_batch(seqs, seed, device=None, pin=False):
    import torch
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(seqs * M_PERSONS, T_FRAMES, V_JOINTS, C_IN, generator=g)        # persons folded into the batch
    y = torch.randint(0, CLASSES, (seqs * M_PERSONS,), generator=g)
    if pin:
        return x.pin_memory(), y.pin_memory()
    return (x.to(device), y.to(device)) if device is not None else (x, y)


def hparams():
    import numpy as np
    return argparse.Namespace(C_in=C_IN, nr_classes=CLASSES, gamma=9, dropout_rate=0.0, residual=False,
                              use_edge_importance=False, max_mask_jitter=0.001, partitioning=2, d=1, distance_file=None,
                              distances=np.random.RandomState(0).rand(V_JOINTS))


def sane_adjacency():
    # dense O(1) learned-adjacency stand-in (SURVEY 8(d) input (ii)); the reference's own normalisation gives
    # entries ~1e5 that overflow an un-calibrated network
    import torch
    return torch.randn(K_PART, V_JOINTS, V_JOINTS, generator=torch.Generator().manual_seed(0)) * 0.2


def run_reference(args):
    """CPU arm: the oracle port of the reference's path (the reference is Python+torch and does not travel to
    the GPU box), all host threads, a bounded sample of the same workload: N_eff = 4 rows of the same shape."""
    import torch
    from oracle import stgcn_oracle as O
    from stgcn_b200 import L_STGCN
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = L_STGCN(hparams(), A=sane_adjacency())
    state = {k: v.detach().clone() for k, v in model.state_dict().items()}
    params = [v.requires_grad_(True) for k, v in state.items() if v.dtype.is_floating_point and 'running' not in k]
    opt = torch.optim.Adam(params, lr=0.01)
    cfgs = O.stack_cfgs(C_IN, 9, 0.0, False)
    seqs = 2
    x, y = synthetic_batch(seqs, 0)

    def step():
        opt.zero_grad(set_to_none=True)
        loss = O.cross_entropy(O.model_forward(x, state, cfgs, True), y)
        loss.backward()
        opt.step()
        return float(loss.detach())
    for _ in range(max(1, min(args.warmup, 1))):
        step()
    steps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    val = seqs / dt
    sample = '%d sequences (N_eff=%d) x 3x300x25, fwd+CE+bwd+Adam fp32, %d steps' % (seqs, seqs * 2, steps)
    print(json.dumps({'impl': 'reference', 'metric': METRIC, 'value': val, 'unit': 'seq/s',
                      'n_gpus': args.gpus, 'steps': steps, 'warmup': 1, 'ms_per_step': dt * 1e3, 'higher_is_better': True,
                      'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
                      'config': {'workload': WORKLOAD, 'seqs_per_step': seqs, 'T': T_FRAMES,
                                 'note': 'reference arm: CPU port of the path (oracle), bounded sample of the same workload'},
                      'cpu_baseline': {'value': val, 'unit': 'seq/s', 'cores': cores, 'kind': 'port', 'sample': sample},
                      'e2e': {'value': val, 'unit': 'seq/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--seqs-per-gpu', type=int, default=32)
    ap.add_argument('--dtype', default='bf16', choices=['bf16', 'f32'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--eager', action='store_true', help='A/B aid: launch the step kernel by kernel instead of replaying its CUDA graph')
    ap.add_argument('--only-train', action='store_true', help='profiling aid: skip the e2e and inference legs')
    ap.add_argument('--only-infer', action='store_true', help='profiling aid: time the eval-mode forward only')
    args = ap.parse_args()
    if args.impl == 'reference':
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from stgcn_b200 import L_STGCN, _lib
    from stgcn_b200.parallel import GradBucketReducer
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    W = max(3, args.warmup)
    S = args.seqs_per_gpu
    dt = torch.bfloat16 if args.dtype == 'bf16' else torch.float32
    torch.manual_seed(0)
    model = L_STGCN(hparams(), compute_dtype=dt, A=sane_adjacency()).to(dev).train()
    from stgcn_b200 import cross_entropy
    from stgcn_b200.optim import Adam
    opt = Adam(model.parameters(), lr=0.01)                 # one launch per 64 parameter tensors (stgcn_adam_step)
    reducer = GradBucketReducer(model.parameters(), 3) if world > 1 else None
    n_batches = 4                                           # distinct resident batches: 4 x S x 2 x 300x25x3 fp32
    dev_batches = [synthetic_batch(S, 100 + rank * 10 + i, dev) for i in range(n_batches)]
    host_batches = [synthetic_batch(S, 100 + rank * 10 + i, pin=True) for i in range(n_batches)]
    flush = torch.empty(160 * 1024 * 1024, dtype=torch.uint8, device=dev)      # > 126 MB L2

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
        """End of a rank.  Every collective of the run is behind us (the last one is the MAX all-reduce of the last timed region), so
        ranks leave independently.  With a captured step the NCCL communicator is still referenced by a live CUDA graph and tearing
        it down (destroy_process_group / interpreter shutdown) was observed to hang: release the graph, drain the device and exit
        the process directly."""
        sys.stdout.flush(); sys.stderr.flush()
        if world > 1:
            torch.cuda.synchronize()
            if cap[0] is None:
                dist.destroy_process_group()
            else:
                os._exit(0)

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

    if args.only_infer:
        for i in range(2):
            train_step(*dev_batches[i % n_batches])          # calibrate the BatchNorm running statistics
        model.eval()
        with torch.no_grad():
            for i in range(W):
                model(dev_batches[i % n_batches][0])
            l0 = _lib.launch_count()
            ms_t = timed(lambda i: (flush.zero_(), model(dev_batches[i % n_batches][0])), args.steps)
            launches = _lib.launch_count() - l0
            ms_f = timed(lambda i: flush.zero_(), args.steps)
            ms_inf = (ms_t - ms_f) / args.steps
            _lib.timing_reset(); _lib.timing_enable(True)
            timed(lambda i: (flush.zero_(), model(dev_batches[i % n_batches][0])), args.steps)
            _lib.timing_enable(False)
        prof = _lib.timing_read()
        if rank == 0:
            print(json.dumps({'infer_seq_s': S * world / (ms_inf * 1e-3), 'ms_per_step': ms_inf, 'gpu_launches': launches,
                              'kernel_ms_per_step': {k: round(v[0] / args.steps, 4) for k, v in prof.items()}}))
        return leave()

    # ---- device-resident training -------------------------------------------------------
    for i in range(W):
        train_step(*dev_batches[i % n_batches])
    l0 = _lib.launch_count()
    if not args.eager:
        from stgcn_b200 import CapturedTrainStep
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

    # ---- end to end: pinned host batch -> H2D -> step -> loss D2H ------------------------
    # public input pipeline (stgcn_b200.pipeline): the pinned batch of step i+1 is copied on a side stream while step i runs;
    # every step still moves its own inputs host -> device inside the timed region and reads its loss back
    from stgcn_b200.pipeline import HostBatchPrefetcher
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
    if args.only_train:
        if rank == 0:
            print(json.dumps({'value': value, 'ms_per_step': ms_step, 'ms_per_step_instrumented': ms_instr, 'gpu_launches': launches,
                              'host_enqueue_ms_per_step': host_enqueue_ms, 'step_mode': 'eager' if args.eager else 'cuda_graph',
                              'kernel_ms_per_step': {k: round(v[0] / args.steps, 4) for k, v in prof.items()}}))
        return leave()
    for i in range(2):
        e2e_step(i)
    ms_e2e = timed(e2e_step, args.steps) / args.steps
    h2d = sum(t.numel() * t.element_size() for t in host_batches[0])

    # ---- inference (cfg3) -------------------------------------------------------------------
    model.eval()
    with torch.no_grad():
        for i in range(W):
            model(dev_batches[i % n_batches][0])
        if args.eager:
            infer_fn = model
        else:
            from stgcn_b200 import CapturedForward
            infer_fn = CapturedForward(model, dev_batches[0][0])          # the eval forward as one CUDA graph
            for i in range(W):
                infer_fn(dev_batches[i % n_batches][0])
        ms_inf = timed(lambda i: infer_fn(dev_batches[i % n_batches][0]), args.steps) / args.steps
        host_enqueue_inf = host_ms[0] / args.steps
    model.train()

    if rank != 0:
        return leave()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    peak_tf = peaks.get('bf16_tflops_sustained', 1400.0)
    fwd_flops, _ = flops_forward_per_neff()
    um = prof.get('umma_conv_gemm')
    roof = None
    traffic, traffic_src = None, None
    try:    # DRAM bytes per launch of the same kernel from the committed ncu capture (dram__bytes_read.sum + dram__bytes_write.sum)
        tj = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json')))['conv_gemm_umma_kernel']
        traffic, traffic_src = tj['avg_dram_bytes_per_launch'], tj['source']
    except Exception:
        pass
    if um:
        ach = um[1] / (um[0] * 1e-3) / 1e12
        roof = {'bound': 'tensor', 'kernel': 'conv_gemm_umma_kernel', 'achieved': ach, 'peak': peak_tf, 'unit': 'TFLOP/s',
                'frac': ach / peak_tf, 'traffic': traffic, 'traffic_unit': 'DRAM bytes per launch (ncu)', 'traffic_source': traffic_src,
                'flop_per_launch': um[1] / max(um[2], 1), 'launches': um[2], 'ms_total': um[0],
                'peak_source': 'MEASURED_PEAKS.json bf16_tflops_sustained' if peaks else 'fallback 1.4 PFLOP/s'}
    out = {'metric': METRIC, 'value': value, 'unit': 'seq/s',
           'n_gpus': world, 'steps': args.steps, 'warmup': W, 'ms_per_step': ms_step, 'higher_is_better': True,
           'scaling': 'weak', 'vs_baseline': None, 'dtype': args.dtype, 'data': 'synthetic',
           'config': {'workload': WORKLOAD,
                      'seqs_per_gpu': S, 'n_eff_per_gpu': S * M_PERSONS, 'global_batch_seqs': S * world, 'T': T_FRAMES,
                      'parallelism': 'dp%d' % world, 'l2': 'flushed (160 MB write) between timed steps, flush time subtracted',
                      'bn': 'per-replica statistics'},
           'model_tflops': 3 * fwd_flops * S * M_PERSONS * world / (ms_step * 1e-3) / 1e12,
           'e2e': {'value': S * world / (ms_e2e * 1e-3), 'unit': 'seq/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 4,
                   'ms_per_step': ms_e2e},
           'infer': {'value': S * world / (ms_inf * 1e-3), 'unit': 'seq/s', 'ms_per_step': ms_inf, 'host_enqueue_ms_per_step': host_enqueue_inf,
                     'mode': 'eager launches' if args.eager else 'one CUDA graph per forward (CapturedForward)',
                     'model_tflops': fwd_flops * S * M_PERSONS * world / (ms_inf * 1e-3) / 1e12},
           'gpu_launches': launches, 'step_mode': 'eager launches' if args.eager else 'one CUDA graph per step (CapturedTrainStep)',
           'host_enqueue_ms_per_step': host_enqueue_ms, 'clocks': clocks, 'roofline': roof, 'ms_per_step_instrumented': ms_instr,
           'kernel_ms_per_step': {k: round(v[0] / args.steps, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}}
    if not args.no_cpu_baseline and world == 1:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), '--impl', 'reference', '--steps', '2', '--warmup', '1'],
                           capture_output=True, text=True)
        try:
            out['cpu_baseline'] = json.loads(r.stdout.strip().splitlines()[-1])['cpu_baseline']
        except Exception:
            out['cpu_baseline'] = {'error': (r.stderr or r.stdout)[-300:]}
    print(json.dumps(out))
    leave()


if __name__ == '__main__':
    main()
