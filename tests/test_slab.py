"""Slab decomposition.  The reference has no multi-device semantics to match, so the
parity statement is: an N-slab run equals the single-GPU run on identical inputs (per-particle fields
keyed by the global lattice id, counts, dt sequence), and the single-GPU run is pinned to the oracle by
test_gpu_parity.py.  Host logic (page cuts, id matching, the send/recv protocol over gloo with
world_size 2) runs without a GPU."""
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def slab():
    importlib.import_module("sphinxsys-learning_b200")
    return importlib.import_module("sphinxsys-learning_b200.slab")


def test_page_cuts_balance_and_cover(slab):
    per_page = [0, 0, 7, 7, 7, 7, 7, 7, 7, 7, 0, 0, 0]
    for n in (1, 2, 4, 8):
        cuts = slab.page_cuts(per_page, n)
        assert cuts[0] == 0 and cuts[-1] == len(per_page) and len(cuts) == n + 1
        assert all(b > a for a, b in zip(cuts, cuts[1:]))
        owned = [sum(per_page[a:b]) for a, b in zip(cuts, cuts[1:])]
        assert sum(owned) == sum(per_page)
        assert max(owned) - min(owned) <= 7
    # more slabs than fluid pages still gives a valid (if unbalanced) cover
    cuts = slab.page_cuts([0, 5, 5, 0], 4)
    assert cuts == sorted(cuts) and len(set(cuts)) == 5


def test_recut_moves_cuts_towards_balance_within_the_neighbour_bound(slab):
    # a column of fluid that has slumped to the right: the old cuts leave slab 0 over-full
    per_page = np.array([0, 0] + [9] * 12 + [5] * 12 + [2] * 12 + [0] * 4)
    old = slab.page_cuts([0, 0] + [9] * 12 + [0] * 28, 3)
    assert slab.imbalance(per_page, old) > 0.5
    cuts = list(old)
    for _ in range(12):
        new = slab.recut(per_page, cuts)
        assert new[0] == 0 and new[-1] == len(per_page) and all(b > a for a, b in zip(new, new[1:]))
        for r in range(1, 3):
            # a cut never moves past the middle of a neighbouring slab: one neighbour exchange suffices
            room = max(min(cuts[r] - cuts[r - 1], cuts[r + 1] - cuts[r]) // 2 - 1, 0)
            assert abs(new[r] - cuts[r]) <= room
            assert new[r] >= cuts[r - 1] + 2 or new[r] == cuts[r]
            assert new[r] <= cuts[r + 1] - 2 or new[r] == cuts[r]
        if new == cuts:
            break
        assert slab.imbalance(per_page, new) <= slab.imbalance(per_page, cuts) + 1e-12
        cuts = new
    assert cuts == slab.page_cuts(per_page, 3)       # bounded moves converge to the balanced cuts
    assert slab.imbalance(per_page, cuts) < 0.06


def test_match_by_gid(slab):
    rng = np.random.default_rng(3)
    gid = rng.permutation(1000)[:200]
    send = rng.permutation(gid)
    recv = rng.permutation(gid)
    si, ri = slab.match_by_gid(send, recv)
    assert np.array_equal(send[si], recv[ri])


WORKER = r'''
import importlib, os, sys
sys.path.insert(0, sys.argv[1])
import torch, torch.distributed as dist
importlib.import_module("sphinxsys-learning_b200")
slab = importlib.import_module("sphinxsys-learning_b200.slab")
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:" + sys.argv[2], rank=int(sys.argv[3]), world_size=2)
comm = slab.DistComm(torch.device("cpu"))
r = comm.rank
assert comm.local_slabs() == [r]
# variable-size exchange (migration + halo): rank 0 sends 3 rows right, rank 1 sends 5 rows left
to_left = torch.full((5 if r == 1 else 0, 4), float(10 + r))
to_right = torch.full((3 if r == 0 else 0, 4), float(20 + r))
(fl, fr), = comm.exchange([(to_left, to_right)])
if r == 0:
    assert fl is None and fr.shape == (5, 4) and float(fr[0, 0]) == 11.0
else:
    assert fr is None and fl.shape == (3, 4) and float(fl[0, 0]) == 20.0
# fixed-size exchange (halo records)
(fl, fr), = comm.exchange_fixed([(torch.full((2, 8), float(r)), torch.full((2, 8), float(r)))], [(2, 2)])
other = fr if r == 0 else fl
assert other.shape == (2, 8) and float(other[0, 0]) == float(1 - r)
assert comm.allreduce([float(r + 1)], "min") == 1.0 and comm.allreduce([float(r + 1)], "sum") == 3.0
import numpy as np
tot = comm.allreduce_sum_array([np.arange(5, dtype=np.int64) * (r + 1)])     # per-page fluid counts of a re-cut
assert tot.dtype == np.int64 and tot.tolist() == [0, 3, 6, 9, 12]
dist.destroy_process_group()
print("ok", r)
'''


def test_dist_comm_protocol_over_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
             for r in range(2)]
    outs = [p.communicate(timeout=300)[0].decode() for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f"ok {r}" in o, o


@pytest.mark.gpu
@pytest.mark.parametrize("n_slabs,real_bytes,pair_kernel,round1_path",
                         [(2, 8, -1, False), (3, 8, 0, False), (2, 4, -1, False), (3, 4, 1, False), (4, 8, -1, False), (3, 8, -1, True)])
def test_slabs_equal_single_gpu(pk, gpu, slab, n_slabs, real_bytes, pair_kernel, round1_path):
    """N slabs emulated in one process (LocalComm) against the undecomposed run; round1_path: host-driven sub-step
    loop, the two-sort reconfigure and send/recv halo records instead of the device-resident loop,
    sphb200_slab_select/pack/rebuild and the halo push into peer memory (sphb200_rows_push / rows_wait_scatter)"""
    dp, steps = 0.05, 8
    one = pk.cases.DamBreak(gpu, ndim=3, dp=dp, real_bytes=real_bytes, with_observer=False)
    many = slab.SlabDamBreak(gpu, slab.LocalComm(n_slabs), dp=dp, real_bytes=real_bytes, device_loop=not round1_path,
                             fast_reconfigure=not round1_path, peer_halo=not round1_path)
    one.sys.set_pair_kernel(pair_kernel)
    for P in many.parts:
        P.case.sys.set_pair_kernel(pair_kernel)
    assert many.n_fluid() == one.water.n
    assert sum(1 for P in many.parts if P.oe > P.ob) == n_slabs  # every slab owns fluid
    # global ids of the single run, from its initial lattice
    pos0 = one.water.get("Position").astype(np.float64)
    oid0, _ = one.water.ids()
    idx = np.rint((pos0 - many.sys_lo) / dp - 0.5).astype(np.int64)
    gid_sorted = (idx[:, 0] * int(many.n_lat[1]) + idx[:, 1]) * int(many.n_lat[2]) + idx[:, 2]
    gid_by_oid = np.empty_like(gid_sorted)
    gid_by_oid[oid0] = gid_sorted
    for s in range(steps):
        r1, rn = one.advection_step(), many.advection_step()
        assert r1[1] == rn[1]
        assert r1[0] == rn[0]  # max-type reductions are order independent: bit-exact dt
    assert many.n_fluid() == one.water.n
    tol = 1e-9 if real_bytes == 8 else 1e-4
    oid, _ = one.water.ids()
    for name in ("Position", "Velocity", "Density", "Pressure", "DensityChangeRate", "Force"):
        ref = one.water.get(name)
        ref_by_gid = dict()
        order = np.argsort(gid_by_oid[oid])
        g, v = many.owned(name)
        o2 = np.argsort(g)
        assert np.array_equal(gid_by_oid[oid][order], g[o2])
        a, b = v[o2].astype(np.float64).reshape(len(g), -1), ref[order].astype(np.float64).reshape(len(g), -1)
        err = np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)
        assert err < tol, (name, err)
    e1, en = one.mechanical_energy(), many.mechanical_energy()
    assert abs(e1 - en) <= 1e-9 * abs(e1) if real_bytes == 8 else abs(e1 - en) <= 1e-4 * abs(e1)


def _compare_owned_with_single(one, many, dp, tol):
    pos_names = ("Position", "Velocity", "Density", "Pressure", "DensityChangeRate", "Force")
    oid, _ = one.water.ids()
    worst = 0.0
    for name in pos_names:
        ref = one.water.get(name)
        g, v = many.owned(name)
        assert len(g) == one.water.n and len(np.unique(g)) == len(g)     # every particle owned exactly once
        order, o2 = np.argsort(one._gid_by_oid[oid]), np.argsort(g)
        assert np.array_equal(one._gid_by_oid[oid][order], g[o2])
        a, b = v[o2].astype(np.float64).reshape(len(g), -1), ref[order].astype(np.float64).reshape(len(g), -1)
        err = np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)
        worst = max(worst, err)
        assert err < tol, (name, err)
    return worst


@pytest.mark.gpu
@pytest.mark.parametrize("n_slabs,recut_interval,in_place,peer_halo", [(3, 0, True, True), (3, 5, True, True), (4, 3, True, False), (3, 5, False, True)])
def test_migration_and_recut_over_a_collapsing_column(pk, gpu, slab, monkeypatch, n_slabs, recut_interval, in_place, peer_halo):
    """long enough for particles to cross the cuts (migrants must stay behind as the sender's halo copies) and, with
    recut_interval > 0, for the cuts themselves to follow the front (whole pages of fluid and wall change slabs):
    the decomposed run still equals the undecomposed one, halo sets included (checked through the sums)"""
    dp, steps = 0.05, 140
    if not in_place:
        monkeypatch.setenv("SPHB200_SLAB_INPLACE", "0")       # the gather path of sphb200_slab_rebuild
    one = pk.cases.DamBreak(gpu, ndim=3, dp=dp, real_bytes=8, with_observer=False)
    many = slab.SlabDamBreak(gpu, slab.LocalComm(n_slabs), dp=dp, real_bytes=8, recut_interval=recut_interval, recut_threshold=0.0,
                             peer_halo=peer_halo)
    pos0 = one.water.get("Position").astype(np.float64)
    oid0, _ = one.water.ids()
    idx = np.rint((pos0 - many.sys_lo) / dp - 0.5).astype(np.int64)
    gid_sorted = (idx[:, 0] * int(many.n_lat[1]) + idx[:, 1]) * int(many.n_lat[2]) + idx[:, 2]
    one._gid_by_oid = np.empty_like(gid_sorted)
    one._gid_by_oid[oid0] = gid_sorted
    cuts0 = list(many.cuts)
    owner0 = {}
    for P in many.parts:
        g = many._gid(P)[P.ob:P.oe].cpu().numpy()
        owner0.update({int(k): P.rank for k in g})
    for s in range(steps):
        r1, rn = one.advection_step(), many.advection_step()
        assert r1[1] == rn[1] and abs(r1[0] - rn[0]) <= 1e-12 * r1[0], (s, r1, rn)
        assert sum(P.oe - P.ob for P in many.parts) == one.water.n
    moved = 0
    for P in many.parts:
        g = many._gid(P)[P.ob:P.oe].cpu().numpy()
        moved += sum(1 for k in g if owner0[int(k)] != P.rank)
    assert moved > 50, moved                                 # particles did change slabs
    if recut_interval:
        assert many.n_recuts >= 1 and list(many.cuts) != cuts0, (many.n_recuts, many.cuts)
        assert slab.imbalance(many.fluid_per_page(), many.cuts) < slab.imbalance(many.fluid_per_page(), cuts0)
    else:
        assert list(many.cuts) == cuts0
    worst = _compare_owned_with_single(one, many, dp, 1e-7)
    e1, en = one.mechanical_energy(), many.mechanical_energy()
    assert abs(e1 - en) <= 1e-9 * abs(e1), (e1, en, worst)


@pytest.mark.gpu
@pytest.mark.parametrize("n_slabs", [2, 4])
def test_probes_are_evaluated_by_the_owning_slab(pk, gpu, slab, n_slabs):
    """observables of a decomposed run - the pressure probes (ObservingAQuantity) come from the slab that owns
    the probe's page, summed over the slabs; equal to the undecomposed run's, also for probes next to a cut"""
    dp = 0.05
    probes = [[x, y, 0.25] for x in (0.31, 0.93, 1.02, 1.49, 1.97) for y in (0.12, 0.55)]
    one = pk.cases.DamBreak(gpu, ndim=3, dp=dp, real_bytes=8, with_observer=True, probes=probes)
    many = slab.SlabDamBreak(gpu, slab.LocalComm(n_slabs), dp=dp, real_bytes=8, probes=probes)
    cut_x = [many.mesh_lo[0] + c * many.spacing for c in many.cuts[1:-1]]
    assert any(min(abs(p[0] - cx) for cx in cut_x) < 2 * many.spacing for p in probes)     # some probes sit near a cut
    for s in range(6):
        r1, rn = one.advection_step(), many.advection_step()
        assert r1 == rn
        a, b = np.asarray(one.probe_pressure()), np.asarray(many.probe_pressure())
        if s == 5:
            many.close()          # releases the landing buffers and contexts; a second close is a no-op
            many.close()
        assert np.allclose(a, b, rtol=1e-9, atol=1e-12), (s, a, b)
    assert np.abs(a).min() > 1e-4 and np.abs(a).max() > 0.02          # pressure has built up at every probe


@pytest.mark.gpu
def test_refused_peer_mapping_falls_back_to_send_recv(pk, gpu, slab, monkeypatch, capsys):
    """a box that refuses the mapping of a neighbour's memory: all slabs agree on the send/recv transport (another
    transport for the same records, not another compute path) and the run is unchanged"""
    def refuse(self):
        raise RuntimeError("peer access denied (test)")
    monkeypatch.setattr(slab.SlabDamBreak, "_peer_setup", refuse)
    one = pk.cases.DamBreak(gpu, ndim=3, dp=0.05, real_bytes=8, with_observer=False)
    many = slab.SlabDamBreak(gpu, slab.LocalComm(3), dp=0.05, real_bytes=8, peer_halo=True)
    assert many.peer_halo is False and "unavailable" in capsys.readouterr().err
    for _ in range(3):
        assert one.advection_step() == many.advection_step()
    assert abs(one.mechanical_energy() - many.mechanical_energy()) <= 1e-9 * abs(one.mechanical_energy())
