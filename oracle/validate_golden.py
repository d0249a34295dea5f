"""ORACLE - TEST INFRASTRUCTURE ONLY.  Pins the CPU restatement against the reference's own regression
fixtures (tests/golden/*.json, extracted by tests/golden/make_golden.py): runs a case on the oracle and
reports the deviation from the stored runs.  `python oracle/validate_golden.py [case ...] [--full]`.
The short variants are what tests/test_oracle_golden.py asserts; --full runs the whole stored series.
"""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pk = importlib.import_module("sphinxsys-learning_b200")
from oracle import binding as ob  # noqa: E402


def golden(case):
    with open(os.path.join(ROOT, "tests", "golden", case + ".json")) as fh:
        return json.load(fh)["quantities"]


def envelope(q):
    """per-snapshot min/max over the stored runs (full-resolution prefix common to all strides)"""
    runs = [np.asarray(r["series"]) for r in q["runs"].values() if r["stride"] == 1]
    n = min(r.shape[1] for r in runs)
    stack = np.stack([r[:, :n] for r in runs])
    return stack.min(0), stack.max(0)


def dambreak_2d(n_samples=4, backend=ob, real_bytes=8):
    """TotalMechanicalEnergy every 200 advection steps (Dambreak.cpp:143-144,200-204)"""
    g = golden("dambreak_2d")["WaterBody_TotalMechanicalEnergy"]
    lo, hi = envelope(g)
    c = pk.cases.DamBreak(backend, ndim=2, dp=0.025, real_bytes=real_bytes)
    out = [c.mechanical_energy()]
    for s in range(1, n_samples):
        for _ in range(200):
            c.advection_step()
        out.append(c.mechanical_energy())
    out = np.asarray(out)
    ref = 0.5 * (lo[0, :n_samples] + hi[0, :n_samples])
    return {"case": "dambreak_2d", "ours": out.tolist(), "golden_lo": lo[0, :n_samples].tolist(),
            "golden_hi": hi[0, :n_samples].tolist(), "max_rel_dev": float(np.max(np.abs(out - ref) / np.abs(ref)))}


def taylor_green(n_samples=3, backend=ob, real_bytes=8):
    """TotalKineticEnergy at t = 0, 0.1, 0.2, ... (taylor_green.cpp:155-192)"""
    g = golden("taylor_green_2d")["WaterBody_TotalKineticEnergy"]
    lo, hi = envelope(g)
    c = pk.cases.TaylorGreen(backend, dp=0.01, real_bytes=real_bytes)
    out = [c.sys.total_kinetic_energy(c.water)]
    for s in range(1, n_samples):
        t_end = c.time
        integ = 0.0
        while integ < 0.1:
            t0 = c.time
            c.advection_step()
            integ += c.time - t0
        out.append(c.sys.total_kinetic_energy(c.water))
    out = np.asarray(out)
    ref = 0.5 * (lo[0, :n_samples] + hi[0, :n_samples])
    return {"case": "taylor_green_2d", "ours": out.tolist(), "golden_lo": lo[0, :n_samples].tolist(),
            "golden_hi": hi[0, :n_samples].tolist(), "max_rel_dev": float(np.max(np.abs(out - ref) / np.abs(ref)))}


def twisting_column(n_samples=200, backend=ob, real_bytes=8):
    """observer Position/Velocity after every step (twisting_column.cpp:154-175)"""
    gq = golden("twisting_column_3d")
    pos_runs = [np.asarray(r["series"]) for r in gq["MyObserver_Position"]["runs"].values() if r["stride"] == 1]
    ref = pos_runs[0][0]  # (snapshots, 3)
    c = pk.cases.TwistingColumn(backend, dp=0.1, real_bytes=real_bytes)
    out = [c.observed()[0]]
    for s in range(1, n_samples):
        c.step()
        out.append(c.observed()[0])
    out = np.asarray(out)
    dev = np.abs(out - ref[:n_samples])
    scale = np.abs(ref[:n_samples]).max()
    return {"case": "twisting_column_3d", "ours_last": out[-1].tolist(), "golden_last": ref[n_samples - 1].tolist(),
            "max_rel_dev": float(dev.max() / scale)}


def dambreak_3d(n_samples=2, backend=ob, real_bytes=8):
    """TotalMechanicalEnergy at t = 0, 1, 2 ... (dambreak.cpp:144,197)"""
    g = golden("dambreak_3d")["WaterBody_TotalMechanicalEnergy"]
    lo, hi = envelope(g)
    c = pk.cases.DamBreak(backend, ndim=3, dp=0.05, real_bytes=real_bytes)
    out = [c.mechanical_energy()]
    for s in range(1, n_samples):
        integ = 0.0
        while integ < 1.0:
            t0 = c.time
            c.advection_step()
            integ += c.time - t0
        out.append(c.mechanical_energy())
    out = np.asarray(out)
    ref = 0.5 * (lo[0, :n_samples] + hi[0, :n_samples])
    return {"case": "dambreak_3d", "ours": out.tolist(), "golden_lo": lo[0, :n_samples].tolist(),
            "golden_hi": hi[0, :n_samples].tolist(), "max_rel_dev": float(np.max(np.abs(out - ref) / np.abs(ref)))}


ALL = {"dambreak_2d": dambreak_2d, "taylor_green": taylor_green, "twisting_column": twisting_column, "dambreak_3d": dambreak_3d}
FULL = {"dambreak_2d": 23, "taylor_green": 51, "twisting_column": 3127, "dambreak_3d": 21}

if __name__ == "__main__":
    full = "--full" in sys.argv
    names = [a for a in sys.argv[1:] if not a.startswith("--")] or list(ALL)
    for nm in names:
        t = time.time()
        res = ALL[nm](FULL[nm]) if full else ALL[nm]()
        res["seconds"] = round(time.time() - t, 1)
        print(json.dumps(res), flush=True)
