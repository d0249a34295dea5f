"""Shared helpers of the parity tests: identical seeded inputs for both backends and order-independent
comparisons (the reference's own particle order is non-deterministic, SURVEY §0.3, so everything is
keyed by OriginalID)."""
import numpy as np

FP_TOL = {8: 1e-6, 4: 1e-4}  # north-star tolerances: rel-L2 per-step fields, fp64 / fp32


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    den = np.linalg.norm(b)
    if den == 0.0:
        return float(np.linalg.norm(a))
    return float(np.linalg.norm(a - b) / den)


def field(body, name, width=None):
    """variable `name` in original-id order"""
    v = body.get(name, width)
    oid, _ = body.ids()
    out = np.empty_like(v)
    out[oid] = v
    return out


def pair_set(relation, k=0, target=None):
    """sorted (original_i, original_j) pair array of a relation"""
    counts, j = relation.neighbors(k)
    src_oid, _ = relation.body.ids()
    tgt = target if target is not None else relation.body
    tgt_oid, _ = tgt.ids()
    i = np.repeat(np.arange(relation.body.n, dtype=np.int64), counts)
    pairs = np.stack([src_oid[i], tgt_oid[j]], axis=1)
    order = np.lexsort((pairs[:, 1], pairs[:, 0]))
    return pairs[order]


def jitter(body, amplitude, seed=12345):
    """positions + U(-a, a)*dp with a fixed seed, same for every backend"""
    pos = field(body, "Position")
    rng = np.random.default_rng(seed)
    noise = rng.uniform(-amplitude, amplitude, size=pos.shape) * body.dp
    oid, _ = body.ids()
    body.set("Position", (pos + noise.astype(pos.dtype))[oid])


def set_field(body, name, values_by_original_id):
    oid, _ = body.ids()
    body.set(name, np.ascontiguousarray(values_by_original_id)[oid])


def seeded_state(case, seed=7):
    """non-trivial velocity / density / force state so every term of the pair sums is exercised"""
    w = case.water
    rng = np.random.default_rng(seed)
    d = case.ndim if hasattr(case, "ndim") else 2
    dtype = case.sys.dtype
    n = w.n
    vel = rng.normal(0.0, 0.3, size=(n, d)).astype(dtype)
    rho = (1.0 + rng.uniform(0.0, 0.02, size=n)).astype(dtype)
    drho = rng.normal(0.0, 0.5, size=n).astype(dtype)
    force = (rng.normal(0.0, 1.0, size=(n, d)) * float(case.dp) ** d).astype(dtype)
    return {"Velocity": vel, "Density": rho, "DensityChangeRate": drho, "Force": force}
