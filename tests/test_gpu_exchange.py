"""The winner exchange fused into the search kernel (tp_exchange_*, tp_launch_exchange), one process per GPU over
CUDA IPC peer mappings: needs at least two GPUs (ranks sharing a GPU would spin on each other), so it is skipped on
a one-GPU box; bench.py --gpus N drives the same path.  The N > 1 host protocol is covered on CPU by
tests/test_multi_rank_cpu.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_fused_exchange_on_all_gpus():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU visible: the fused exchange needs one GPU per rank")
    n = min(n, 8)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", "29613", os.path.join(HERE, "exchange_ranks.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "exchange ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
