"""torchrun worker: N-slab run over NCCL (DistComm) against the single-GPU run on rank 0.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 scratch/dist_check.py [dp] [steps]
"""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pk = importlib.import_module("sphinxsys-learning_b200")
slab = importlib.import_module("sphinxsys-learning_b200.slab")

dp = float(sys.argv[1]) if len(sys.argv) > 1 else 0.05
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
comm = slab.DistComm(torch.device("cuda", lr))
many = slab.SlabDamBreak(pk.host, comm, dp=dp, real_bytes=8, device=lr)
one = pk.cases.DamBreak(pk.host, ndim=3, dp=dp, real_bytes=8, with_observer=False, device=lr) if comm.rank == 0 else None
seq = []
for s in range(steps):
    rn = many.advection_step()
    if one is not None:
        r1 = one.advection_step()
        assert r1 == rn, (s, r1, rn)
    seq.append(rn)
en = many.mechanical_energy()
nf = many.n_fluid()
if one is not None:
    e1 = one.mechanical_energy()
    assert nf == one.water.n, (nf, one.water.n)
    assert abs(e1 - en) <= 1e-9 * abs(e1), (e1, en)
    print(f"dist slab check ok: world={comm.world} dp={dp} steps={steps} n_fluid={nf} energy={en!r} (single {e1!r}) "
          f"dt sequence identical, cuts={many.cuts}")
dist.barrier()
dist.destroy_process_group()
