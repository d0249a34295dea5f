"""torchrun worker: where a slab advection step spends its time, per rank.
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29612 scratch/slab_profile.py [dp1] [steps]
Phase A: the production step (device-resident sub-step loop), CUDA events around whole steps.
Phase B: the same step with a synchronisation after every phase (host clock), to attribute the time."""
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
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
dp = 1.0 / round((1.0 / dp1) * world ** (1.0 / 3.0))
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
comm = slab.DistComm(torch.device("cuda", lr))
case = slab.SlabDamBreak(pk.host, comm, dp=dp, real_bytes=8, device=lr, stream=stream.cuda_stream)
for _ in range(3):
    case.advection_step()
P = case.parts[0]; S = P.case.sys; w = P.case.water
# ---- phase A
dist.barrier(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
sub0 = case.n_acoustic
e0.record(stream)
for _ in range(steps):
    case.advection_step()
e1.record(stream); torch.cuda.synchronize()
res = {"rank": comm.rank, "owned": P.oe - P.ob, "resident": w.n, "wall": P.case.wall.n, "production_ms_per_step": round(e0.elapsed_time(e1) / steps, 2),
       "substeps_per_step": (case.n_acoustic - sub0) / steps}
# ---- phase B
acc = {}
def timed(name, fn):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = fn(); torch.cuda.synchronize()
    acc[name] = acc.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
    return r
def reconfigure():
    n_keep, n_l, n_r = timed("rc_select", lambda: w.slab_select(P.has_left, P.has_right))
    width = sum(w.width(nm) for nm in slab.MIGRATE_VARS)
    bl = torch.empty((n_l, width), dtype=torch.float64, device=case.device); br = torch.empty((n_r, width), dtype=torch.float64, device=case.device)
    def pack():
        if n_l: w.slab_pack(1, slab.MIGRATE_VARS, bl.data_ptr())
        if n_r: w.slab_pack(2, slab.MIGRATE_VARS, br.data_ptr())
    timed("rc_pack", pack)
    (fl, fr), = timed("rc_exchange", lambda: comm.exchange([(bl if P.has_left else None, br if P.has_right else None)]))
    nl = fl.shape[0] if fl is not None else 0; nr = fr.shape[0] if fr is not None else 0
    timed("rc_rebuild", lambda: w.slab_rebuild(slab.MIGRATE_VARS, fl.data_ptr() if nl else 0, nl, fr.data_ptr() if nr else 0, nr))
    timed("rc_cell_list", w.update_cell_list)
    timed("rc_inner", P.case.inner.update)
    timed("rc_wall", P.case.wall_contact.update)
    timed("rc_halo_maps", case._halo_maps)
    res["migrate_rows"] = [n_l, n_r, nl, nr]
def step():
    Dt = timed("adv_dt", lambda: case._min_over_slabs(lambda Q: Q.case.sys.advection_dt(Q.case.water, case.U_ref)))
    timed("density+union", lambda: S.density_summation(P.case.inner, P.case.wall_contact, free_surface=True))
    S.relaxation_begin(Dt, case.dt)
    while True:
        timed("K1", lambda: S.integration_1st_half_init(P.case.inner, -1.0))
        timed("halo_a", lambda: case._halo_records(0))
        timed("K2", lambda: S.integration_1st_half_interact(P.case.inner, P.case.wall_contact, True, -1.0))
        timed("halo_b", lambda: case._halo_records(1))
        timed("K3", lambda: S.integration_2nd_half(P.case.inner, P.case.wall_contact, True, -1.0))
        timed("dt_allreduce", lambda: (comm.allreduce_max_device([P.slot]), S.substep_ctl(w, 0.6, 1, 1)))
        st = S.relaxation_state()
        if st["done"]:
            break
    case.dt = st["dt"]; case.n_acoustic += st["substeps"]
    reconfigure()
dist.barrier()
for _ in range(steps):
    step()
res["phases_ms_per_step"] = {k: round(v / steps, 2) for k, v in acc.items()}
res["phases_sum"] = round(sum(acc.values()) / steps, 2)
out = [None] * world
dist.all_gather_object(out, res)
if comm.rank == 0:
    for r in out: print(json.dumps(r))
dist.destroy_process_group()
