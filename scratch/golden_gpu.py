"""run-level observables of the CUDA path against the reference's regression fixtures (development probe)"""
import importlib, json, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pk = importlib.import_module("sphinxsys-learning_b200")
from oracle import validate_golden as vg
for name, n in (("dambreak_2d", 12), ("taylor_green", 16), ("twisting_column", 1200), ("dambreak_3d", 5)):
    t = time.time()
    r = vg.ALL[name](n, backend=pk.host)
    print(name, n, "max_rel_dev", r["max_rel_dev"], "seconds", round(time.time() - t, 1), flush=True)
