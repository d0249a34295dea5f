import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pk = importlib.import_module("sphinxsys-learning_b200")
pk.capi.load()
g = pk.cases.TaylorGreen(pk.host, dp=0.02)
for s in range(12):
    r = g.advection_step()
    g.sys.synchronize() if hasattr(g.sys, "synchronize") else None
    print(s, r, flush=True)
    o, sid = g.water.ids()
    print(" ids ok", o[:4], flush=True)
    v = g.water.get("Velocity")
    print(" vel", v[:2].ravel(), flush=True)
