"""Developed-flow measurement: the 20 M dam break is run from the lattice to t >= T (default 1.0: the
front has left the column and runs along the floor), then the same quantities as bench.py's resident leg are measured on that
state: ms per advection step, sub-steps, particle-steps/s, K2 / K3 kernel times, union-list slots, masked-search flag.
    python scratch/developed_state.py [T] [dp] > gpurun_out/developed.json"""
import importlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
pk = importlib.import_module("sphinxsys-learning_b200")
T = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
dp = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0 / 272.0
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
c = pk.cases.DamBreak(pk.host, ndim=3, dp=dp, real_bytes=8, with_observer=True, device=0, stream=stream.cuda_stream)
S, w = c.sys, c.water
def measure(tag):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sub0 = c.n_acoustic
    e0.record(stream)
    for _ in range(6):
        c.advection_step()
    e1.record(stream); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 6
    sub = (c.n_acoustic - sub0) / 6
    def kms(fn):
        fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(3): fn()
        b.record(stream); torch.cuda.synchronize()
        return a.elapsed_time(b) / 3
    keep = {nm: w.get(nm) for nm in ("DensityChangeRate", "Force", "Density", "Position", "Velocity", "Pressure")}
    k3 = kms(lambda: S.integration_2nd_half(c.inner, c.wall_contact, True, 0.0))
    k2 = kms(lambda: S.integration_1st_half(c.inner, c.wall_contact, True, 0.0))
    for nm, v in keep.items():
        w.set(nm, v)
    st, sw = c.inner.stats(), c.wall_contact.stats()
    counts = c.inner.neighbor_counts()[0]
    return {"state": tag, "time": c.time, "advection_steps_done": c.n_advection, "ms_per_advection_step": ms, "substeps_per_advection_step": sub,
            "particle_steps_per_s": w.n * sub / (ms * 1e-3), "K2_ms": k2, "K3_ms": k3, "pairs_per_particle": (st["pairs"] + sw["pairs"]) / w.n,
            "neighbours_min_max": [int(counts.min()), int(counts.max())], "union_slots_per_group": st["union_entries"] / max(st["union_groups"], 1),
            "masked_search_inner": st["masked_search"], "masked_search_wall": sw["masked_search"],
            "front_x": float(S.position_bound(w)[0]), "max_speed": S.maximum_speed(w), "energy": c.mechanical_energy()}
for _ in range(3):
    c.advection_step()
out = [measure("lattice (start)")]
t0 = time.time()
while c.time < T:
    c.advection_step()
out.append(measure(f"developed (t >= {T})"))
out[-1]["wall_seconds_to_reach"] = time.time() - t0
print(json.dumps({"n_fluid": w.n, "n_wall": c.wall.n, "dp": dp, "states": out}, indent=1))
