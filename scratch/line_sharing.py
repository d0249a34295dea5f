"""How many distinct 128-byte lines does one warp-wide neighbour gather touch? (CPU experiment on oracle lists)"""
import importlib, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pk = importlib.import_module("sphinxsys-learning_b200")
from oracle import binding as ob
dp = float(sys.argv[1]) if len(sys.argv) > 1 else 0.025
jit = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
c = pk.cases.DamBreak(ob, ndim=3, dp=dp, with_observer=False)
w = c.water
if jit:
    rng = np.random.default_rng(1); pos = w.get("Position"); w.set("Position", pos + rng.uniform(-jit, jit, pos.shape) * dp)
    w.update_cell_list(); c.update_configuration()
lin, _ = w.cell_keys()
order = np.argsort(lin, kind="stable")          # GPU order: stable counting sort by linear cell index
rank = np.empty_like(order); rank[order] = np.arange(len(order))
counts, j = c.inner.neighbors()
off = np.concatenate([[0], np.cumsum(counts)])
n = w.n
lists = [np.sort(rank[j[off[i]:off[i + 1]]]) for i in order]   # list of particle (new index) in new index space
cell_sorted = lin[order]
REC_PER_LINE = int(sys.argv[3]) if len(sys.argv) > 3 else 4
def cost(lists_of_32, width=None):
    m = max(len(l) for l in lists_of_32)
    wf = 0
    for k in range(m):
        js = [l[k] for l in lists_of_32 if k < len(l) and l[k] >= 0]
        wf += len(set(x // REC_PER_LINE for x in js))
    return m, wf
tot_it = tot_wf = tot_pairs = 0
tot_it2 = tot_wf2 = 0
nz = c.water.info()["cells"][2]
GROUP = int(os.environ.get("GROUP", "1"))
for s in range(0, n - 31, 32 * 7):   # sample every 7th slice
    L = lists[s:s + 32]
    it, wf = cost(L)
    tot_it += it; tot_wf += wf; tot_pairs += sum(len(l) for l in L)
    # run-aligned: split each list by (x,y) column of j's cell, pad each run to the slice max
    runs = {}
    for li, l in enumerate(L):
        cols = cell_sorted[l] // (nz * GROUP)
        for col in np.unique(cols):
            runs.setdefault(col, [None] * 32)[li] = l[cols == col]
    it2 = wf2 = 0
    for col, per_lane in sorted(runs.items()):
        pl = [x if x is not None else np.zeros(0, int) for x in per_lane]
        a, b = cost(pl)
        it2 += a; wf2 += b
    tot_it2 += it2; tot_wf2 += wf2
print(f"dp={dp} jitter={jit} rec/line={REC_PER_LINE}: pairs/slice-iter  current: iters {tot_it} wavefronts {tot_wf} ({tot_wf/tot_it:.1f}/iter, {tot_wf/tot_pairs:.2f}/pair)"
      f" | run-aligned: iters {tot_it2} (+{100*(tot_it2/tot_it-1):.0f}%) wavefronts {tot_wf2} ({tot_wf2/tot_it2:.1f}/iter, {tot_wf2/tot_pairs:.2f}/pair)")
# ---- G-grouping: thread owns G consecutive particles, gathers the union of their lists once
for G in (2, 3, 4):
    tot_union = tot_pairs = tot_it = tot_wf = 0
    for s in range(0, n - 32 * G + 1, 32 * G * 7):
        groups = [lists[s + t * G: s + (t + 1) * G] for t in range(32)]
        U = [np.unique(np.concatenate(g)) for g in groups]
        it, wf = cost(U)
        tot_it += it; tot_wf += wf
        tot_union += sum(len(u) for u in U); tot_pairs += sum(len(l) for g in groups for l in g)
    print(f"G={G}: union entries/pair {tot_union/tot_pairs:.3f}  slice iters per particle-slice {tot_it/ (tot_pairs/ (sum(len(l) for l in lists)/n) /32):.1f}  wavefronts/pair {tot_wf/tot_pairs:.3f}  body slots/useful pair {G*tot_union/tot_pairs:.2f}  padded slots/useful {G*32*tot_it/tot_pairs:.2f}")
# ---- S lanes per particle: lane (p, s) walks entries S*k+s of particle p's ascending list
for S in (2, 4, 8):
    P = 32 // S
    tot_it = tot_wf = tot_pairs = 0
    for s0 in range(0, n - P + 1, P * 29):
        L = lists[s0:s0 + P]
        m = max((len(l) + S - 1) // S for l in L)
        for k in range(m):
            js = [l[S * k + s] for l in L for s in range(S) if S * k + s < len(l)]
            tot_wf += len(set(x // REC_PER_LINE for x in js))
        tot_it += m; tot_pairs += sum(len(l) for l in L)
    print(f"S={S} lanes/particle: wavefronts/iter {tot_wf/tot_it:.1f}  wavefronts/pair {tot_wf/tot_pairs:.3f}  padded slots/useful {32*tot_it/tot_pairs:.2f}")
