"""The alternative kernel paths behind the A/B environment switches must stay parity-green.

The switches are read once per process, so each one is exercised by re-running the bf16 block / model parity tests in a
child interpreter with the variable set.
"""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize('var', ['STGCN_FUSED_SPATIAL=1', 'STGCN_FUSED_SPATIAL=0', 'STGCN_NO_HALO=1', 'STGCN_SPLIT_DA=1',
                                 'STGCN_NO_FUSED_STATS=1', 'STGCN_FUSED_VSUM=1', 'STGCN_NO_TAP_STACK=1'])
def test_switched_path_is_parity_green(var):
    if os.environ.get('STGCN_SWITCH_CHILD'):
        pytest.skip('child run')
    env = dict(os.environ, STGCN_SWITCH_CHILD='1')
    name, value = var.split('=')
    env[name] = value
    r = subprocess.run([sys.executable, '-m', 'pytest', os.path.join(ROOT, 'tests', 'test_gpu_block.py'),
                        os.path.join(ROOT, 'tests', 'test_gpu_umma.py'), '-q', '-x', '-m', 'gpu', '-k', 'bfloat16 or umma_conv_gemm'],
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-1000:]


def test_host_batch_prefetcher_round_trip():
    from stgcn_b200.pipeline import HostBatchPrefetcher
    dev = torch.device('cuda', 0)
    host = [(torch.randn(8, 20, 25, 3).pin_memory(), torch.randint(0, 60, (8,)).pin_memory()) for _ in range(5)]
    pf = HostBatchPrefetcher(dev)
    pf.push(host[0])
    acc = []
    for i in range(5):
        if i + 1 < 5:
            pf.push(host[i + 1])
        x, y = pf.pop()
        acc.append((x.double().sum() + y.double().sum()))       # queued on the current stream after the copy
        z = x * 2                                                # keep the slot busy a little
        pf.done()
        del z
    torch.cuda.synchronize()
    for i in range(5):
        ref = host[i][0].double().sum() + host[i][1].double().sum()
        assert abs(float(acc[i]) - float(ref)) < 1e-6 * max(1.0, abs(float(ref)))
    with pytest.raises(RuntimeError):
        pf.pop()
