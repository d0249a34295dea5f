"""two advection steps of the 20 M dam break (development tool: run under ncu for a launch list)"""
import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
pk = importlib.import_module("sphinxsys-learning_b200")
dp = float(sys.argv[1]) if len(sys.argv) > 1 else 1 / 272
rb = int(sys.argv[2]) if len(sys.argv) > 2 else 8
c = pk.cases.DamBreak(pk.host, ndim=3, dp=dp, real_bytes=rb, with_observer=False, device=0)
for _ in range(2):
    c.advection_step()
torch.cuda.synchronize()
print("ok", c.n_acoustic)
