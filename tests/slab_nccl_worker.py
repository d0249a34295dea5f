"""torchrun worker of tests/test_slab_nccl.py: the N-slab dam break over REAL NCCL (slab.DistComm, one process per GPU)
against the single-GPU run on rank 0 - (Dt, sub-steps) sequence bit-exact, per-particle fields keyed by the global
lattice id < 1e-9, total mechanical energy, particle count.  Prints one line 'slab nccl parity ok ...' on rank 0.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port P tests/slab_nccl_worker.py [dp] [steps]
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
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
comm = slab.DistComm(torch.device("cuda", lr))
PROBES = [[x, y, 0.25] for x in (0.31, 0.93, 1.02, 1.49, 1.97) for y in (0.12, 0.55)]
many = slab.SlabDamBreak(pk.host, comm, dp=dp, real_bytes=8, device=lr, probes=PROBES)
one = pk.cases.DamBreak(pk.host, ndim=3, dp=dp, real_bytes=8, with_observer=True, probes=PROBES, device=lr) if comm.rank == 0 else None
if one is not None:
    pos0 = one.water.get("Position").astype(np.float64)
    oid0, _ = one.water.ids()
    idx = np.rint((pos0 - many.sys_lo) / dp - 0.5).astype(np.int64)
    gid_sorted = (idx[:, 0] * int(many.n_lat[1]) + idx[:, 1]) * int(many.n_lat[2]) + idx[:, 2]
    gid_by_oid = np.empty_like(gid_sorted)
    gid_by_oid[oid0] = gid_sorted
for s in range(steps):
    rn = many.advection_step()
    pn = many.probe_pressure()                       # owner-rank evaluation + all-reduce(sum): every rank takes part
    if one is not None:
        r1 = one.advection_step()
        assert r1 == rn, ("(Dt, sub-steps) differ at advection step", s, r1, rn)
        assert np.allclose(np.asarray(one.probe_pressure()), pn, rtol=1e-9, atol=1e-12), ("probe pressures differ at step", s)
en = many.mechanical_energy()
nf = many.n_fluid()
worst = {}
for name in ("Position", "Velocity", "Density", "Pressure", "DensityChangeRate", "Force"):
    g, v = many.owned(name)
    parts = [None] * comm.world
    dist.all_gather_object(parts, (g, v))
    if one is not None:
        g = np.concatenate([p[0] for p in parts])
        v = np.concatenate([p[1] for p in parts], axis=0)
        assert len(np.unique(g)) == len(g) == one.water.n, "every particle owned exactly once"
        oid, _ = one.water.ids()
        ref = one.water.get(name)
        order, o2 = np.argsort(gid_by_oid[oid]), np.argsort(g)
        assert np.array_equal(gid_by_oid[oid][order], g[o2])
        a, b = v[o2].astype(np.float64).reshape(len(g), -1), ref[order].astype(np.float64).reshape(len(g), -1)
        worst[name] = float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))
        assert worst[name] < 1e-9, (name, worst[name])
if one is not None:
    e1 = one.mechanical_energy()
    assert nf == one.water.n and abs(e1 - en) <= 1e-12 * abs(e1), (nf, e1, en)
    print(f"slab nccl parity ok: world={comm.world} dp={dp} steps={steps} n_fluid={nf} sub-steps={many.n_acoustic} "
          f"max rel-L2 {max(worst.values()):.2e} energy {en!r} cuts={many.cuts}", flush=True)
many.close()            # IPC mappings closed before the owners free their landing buffers
dist.barrier()
dist.destroy_process_group()
