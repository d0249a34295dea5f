"""hit ratio of union lists for groups of 2 / 4 / 8 consecutive particles (development tool)"""
import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
pk = importlib.import_module("sphinxsys-learning_b200")
dp = float(sys.argv[1]) if len(sys.argv) > 1 else 1 / 100
c = pk.cases.DamBreak(pk.host, ndim=3, dp=dp, real_bytes=8, with_observer=False, device=0)
for _ in range(3):
    c.advection_step()
counts, j = c.inner.neighbors()
n = len(counts)
i = np.repeat(np.arange(n, dtype=np.int64), counts)
for G in (2, 4, 8):
    key = (i // G) * n + j
    u = np.unique(key).size
    print(G, "entries/group", u / (n / G), "hit", len(j) / (u * G))
