import importlib, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pk = importlib.import_module("sphinxsys-learning_b200")
from oracle import binding as ob
o = pk.cases.TaylorGreen(ob, dp=0.02)
for s in range(10):
    print(s, o.advection_step(), flush=True)
for name in ("Density", "Pressure", "Velocity", "Position", "ViscousForce", "KernelGradientIntegral"):
    v = o.water.get(name); oid,_ = o.water.ids()
    print(name, v.shape, np.abs(v).max())
