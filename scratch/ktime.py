"""per-kernel timing of the dam-break path (development tool; not part of tests/bench)"""
import importlib, sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
pk = importlib.import_module("sphinxsys-learning_b200")
if os.environ.get("SPHB200_LIB"):
    pk.capi.LIB_PATH = os.path.abspath(os.environ["SPHB200_LIB"])
QUICK = os.environ.get("QUICK") == "1"
dp = float(sys.argv[1]) if len(sys.argv) > 1 else 1 / 272
reals = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [8, 4]
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
for rb in reals:
    c = pk.cases.DamBreak(pk.host, ndim=3, dp=dp, real_bytes=rb, with_observer=False, device=0, stream=stream.cuda_stream)
    S, w = c.sys, c.water
    for _ in range(3):
        c.advection_step()
    def timeit(fn, reps=5):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps): fn()
        e1.record(stream); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    dt = c.dt
    res = {"real": rb, "n_f": w.n, "n_w": c.wall.n}
    res["first_half"] = timeit(lambda: S.integration_1st_half(c.inner, c.wall_contact, True, 0.0))
    res["second_half"] = timeit(lambda: S.integration_2nd_half(c.inner, c.wall_contact, True, 0.0))
    if QUICK:
        print(os.environ.get("SPHB200_LIB", "default"), json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in res.items()}), flush=True)
        del c, S, w
        torch.cuda.empty_cache()
        continue
    res["density"] = timeit(lambda: S.density_summation(c.inner, c.wall_contact, True))
    res["cell_list"] = timeit(lambda: w.update_cell_list(), 3)
    res["inner_update"] = timeit(lambda: c.inner.update(), 3)
    res["wall_update"] = timeit(lambda: c.wall_contact.update(), 3)
    res["adv_step"] = timeit(lambda: c.advection_step(), 3)
    n0 = c.n_acoustic; c.advection_step(); res["substeps_per_adv"] = c.n_acoustic - n0
    res["psteps_per_s"] = w.n * res["substeps_per_adv"] / (res["adv_step"] * 1e-3)
    print(json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in res.items()}), flush=True)
    del c, S, w
    torch.cuda.empty_cache()
