"""torchrun worker: where a slab advection step spends its time, per rank (CUDA events around each phase).
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29612 scratch/slab_profile.py [dp1] [steps]"""
import importlib, os, sys, time, json
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pk = importlib.import_module("sphinxsys-learning_b200")
slab = importlib.import_module("sphinxsys-learning_b200.slab")
lr = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
dp1 = float(sys.argv[1]) if len(sys.argv) > 1 else 1 / 272
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dp = 1.0 / round((1.0 / dp1) * world ** (1.0 / 3.0))
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
comm = slab.DistComm(torch.device("cuda", lr))
case = slab.SlabDamBreak(pk.host, comm, dp=dp, real_bytes=8, device=lr, stream=stream.cuda_stream)
for _ in range(3):
    case.advection_step()
acc, host = {}, {}
def timed(name, fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(stream); r = fn(); e1.record(stream)
    host[name] = host.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
    acc.setdefault(name, []).append((e0, e1)); return r
P = case.parts[0]; S = P.case.sys
import math
def step():
    Dt = timed("adv_dt", lambda: case._min_over_slabs(lambda Q: Q.case.sys.advection_dt(Q.case.water, case.U_ref)))
    timed("density", lambda: S.density_summation(P.case.inner, P.case.wall_contact, free_surface=True))
    relax = 0.0
    while relax < Dt:
        timed("K1", lambda: S.integration_1st_half_init(P.case.inner, case.dt))
        timed("halo_a", lambda: case._halo_records(0))
        timed("K2", lambda: S.integration_1st_half_interact(P.case.inner, P.case.wall_contact, True, case.dt))
        timed("halo_b", lambda: case._halo_records(1))
        timed("K3", lambda: S.integration_2nd_half(P.case.inner, P.case.wall_contact, True, case.dt))
        case.n_acoustic += 1
        case.dt = timed("dt_allreduce", lambda: case._min_over_slabs(lambda Q: Q.case.sys.acoustic_dt(Q.case.water)))
        relax += case.dt
    timed("reconfigure", case.reconfigure)
dist.barrier(); torch.cuda.synchronize()
w0 = time.perf_counter()
for _ in range(steps):
    step()
torch.cuda.synchronize(); wall = (time.perf_counter() - w0) * 1e3
res = {k: round(sum(a.elapsed_time(b) for a, b in v) / steps, 2) for k, v in acc.items()}
res["host_ms"] = {k: round(v / steps, 2) for k, v in host.items()}
res["wall_ms_per_step"] = round(wall / steps, 2); res["rank"] = comm.rank; res["owned"] = P.oe - P.ob; res["resident"] = P.case.water.n
out = [None] * world
dist.all_gather_object(out, res)
if comm.rank == 0:
    for r in out: print(json.dumps(r))
dist.destroy_process_group()
