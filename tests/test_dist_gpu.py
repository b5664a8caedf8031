"""Data parallelism on real GPUs: tools/check_dp.py under torchrun on 2 GPUs -- gradient identity
with the single-process global batch, three Adam steps eager and with the whole step (closure + NCCL all-reduce +
Adam) replayed as one CUDA graph.  Skipped on a single-GPU box; the CPU suite covers the same host logic on gloo."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')
def test_dp2_gradient_and_whole_step_graph_match_single_process():
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', '29533', os.path.join(ROOT, 'tools', 'check_dp.py')]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    sys.stdout.write(res.stdout[-6000:])
    assert res.returncode == 0, res.stdout[-4000:] + res.stderr[-4000:]
    assert 'NCCL captured' in res.stdout and 'bit-identical across ranks' in res.stdout
