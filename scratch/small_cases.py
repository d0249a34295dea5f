"""Throughput of the small reference configs (launch-bound on a B200): sub-steps/s with the host-driven loop
(one sphb200_acoustic_dt round trip per sub-step) and with the device-resident loop (sphb200_acoustic_relaxation).
    python scratch/small_cases.py > gpurun_out/small_cases.json"""
import importlib
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pk = importlib.import_module("sphinxsys-learning_b200")
pk.capi.load()
out = {}
for name, make in (("C1 2-D dam break, 3200 fluid", lambda f: pk.cases.DamBreak(pk.host, ndim=2, dp=0.025, fused_loop=f)),
                   ("C2 Taylor-Green, 10000", lambda f: pk.cases.TaylorGreen(pk.host, dp=0.01, fused_loop=f)),
                   ("C3 3-D dam break default, 8000 fluid", lambda f: pk.cases.DamBreak(pk.host, ndim=3, dp=0.05, fused_loop=f))):
    for fused in (False, True):
        c = make(fused)
        for _ in range(20):
            c.advection_step()
        c.sys.synchronize()
        n0, l0, t0 = c.n_acoustic, c.sys.launches, time.perf_counter()
        for _ in range(200):
            c.advection_step()
        c.sys.synchronize()
        dt = time.perf_counter() - t0
        sub = c.n_acoustic - n0
        out[f"{name}, {'device-resident loop' if fused else 'host-driven loop'}"] = {
            "substeps_per_s": sub / dt, "particle_steps_per_s": c.water.n * sub / dt, "advection_steps_per_s": 200 / dt,
            "launches_per_substep": (c.sys.launches - l0) / sub}
c = pk.cases.TwistingColumn(pk.host, dp=0.1)
for _ in range(50):
    c.step()
c.sys.synchronize()
t0 = time.perf_counter()
for _ in range(1000):
    c.step()
c.sys.synchronize()
dt = time.perf_counter() - t0
out["C5 twisting column, 6100"] = {"steps_per_s": 1000 / dt, "particle_steps_per_s": c.column.n * 1000 / dt}
print(json.dumps(out, indent=1))
