"""Hardware parity of the slab path: SlabDamBreak over torch.distributed/NCCL on 2 (and 4, if present) GPUs of the box
must reproduce the single-GPU run - the LocalComm test in test_slab.py proves the decomposition, this one the real
transport (send/recv of migrants and halo records, the all-reduce of the time-step maximum).  Skipped below 2 GPUs."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("world,dp,steps", [(2, 0.05, 10), (2, 1.0 / 48.0, 4), (4, 0.05, 6)])
def test_nccl_slabs_equal_single_gpu(world, dp, steps):
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs on the box")
    port = str(29650 + os.getpid() % 300 + world)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
                        "--master-port", port, os.path.join(ROOT, "tests", "slab_nccl_worker.py"), repr(dp), str(steps)],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "slab nccl parity ok" in r.stdout, (r.stdout[-3000:], r.stderr[-3000:])
