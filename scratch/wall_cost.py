"""K2/K3 with and without the wall contact (development tool)"""
import importlib, sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
pk = importlib.import_module("sphinxsys-learning_b200")
dp = float(sys.argv[1]) if len(sys.argv) > 1 else 1 / 272
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
c = pk.cases.DamBreak(pk.host, ndim=3, dp=dp, real_bytes=8, with_observer=False, device=0, stream=stream.cuda_stream)
S = c.sys
for _ in range(3):
    c.advection_step()
def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps): fn()
    e1.record(stream); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
res = {}
res["k1"] = timeit(lambda: S.integration_1st_half_init(c.inner, 0.0))
res["k2_wall"] = timeit(lambda: S.integration_1st_half_interact(c.inner, c.wall_contact, True, 0.0))
res["k2_nowall"] = timeit(lambda: S.integration_1st_half_interact(c.inner, None, True, 0.0))
res["k3_wall"] = timeit(lambda: S.integration_2nd_half(c.inner, c.wall_contact, True, 0.0))
res["k3_nowall"] = timeit(lambda: S.integration_2nd_half(c.inner, None, True, 0.0))
res["density_wall"] = timeit(lambda: S.density_summation(c.inner, c.wall_contact, True))
res["density_nowall"] = timeit(lambda: S.density_summation(c.inner, None, True))
print(json.dumps({k: round(v, 3) for k, v in res.items()}))
