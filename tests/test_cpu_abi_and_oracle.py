"""CPU-side tests (no GPU): the C-ABI library loads and exports every declared symbol, the oracle
reproduces the reference's closed-form known answers and stays inside the reference's own regression
envelopes (tests/golden, extracted from the reference's regression_test_tool XML files)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(pk):
    header = open(os.path.join(ROOT, "include", "sphb200.h")).read()
    declared = set(re.findall(r"\b(sphb200_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 45
    lib = ctypes.CDLL(pk.capi.LIB_PATH)
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert declared == set(pk.capi.SIGNATURES), declared ^ set(pk.capi.SIGNATURES)


def test_no_device_is_a_loud_error(pk):
    """the product path must not fall back to the CPU"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pk.capi.SphB200Error) as e:
        pk.host.SPHSystem(3, [0, 0, 0], [1, 1, 1], 0.1)
    assert e.value.status == pk.capi.ERR_NO_DEVICE


def test_cpp_adapter_main_builds_and_fails_loudly_without_a_gpu(pk):
    """include/sphb200_adapter.hpp + examples/dambreak_3d.cpp (the reference's main on the C++ host side)"""
    import subprocess
    import torch
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "examples")])
    exe = os.path.join(ROOT, "examples", "dambreak_3d")
    assert os.path.exists(exe)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the run itself is covered by the gpu tests")
    r = subprocess.run([exe, "0.1", "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "no CUDA device" in r.stderr


def test_reference_main_compiles_unmodified_on_the_cuda_policy(tmp_path):
    """integration/: the reference's tests/3d_examples/test_3d_dambreak/dambreak.cpp and the CK main
    tests/tests_sycl/3d_examples/test_3d_dambreak_sycl/dambreak.cpp, read where they lie and not a byte changed, compile against the class skeleton + the CUDA execution policy (src_cuda) + libsphb200; without a GPU the
    binary fails loudly (the reference's convention: message + exit 1)"""
    import subprocess
    import torch
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "integration")])
    mk = open(os.path.join(ROOT, "integration", "Makefile")).read()
    # compiled from the reference tree, not from copies: the classic main and its CK twin of the SYCL test tree
    assert "$(REF)/tests/3d_examples/test_3d_dambreak/dambreak.cpp" in mk
    assert "$(REF)/tests/tests_sycl/3d_examples/test_3d_dambreak_sycl/dambreak.cpp" in mk
    assert "$(REF)/tests/2d_examples/test_2d_dambreak/Dambreak.cpp" in mk
    assert "$(REF)/tests/3d_examples/test_3d_twisting_column/twisting_column.cpp" in mk
    assert "$(REF)/tests/2d_examples/test_2d_taylor_green/taylor_green.cpp" in mk
    assert "$(REF)/tests/tests_sycl/2d_examples/test_2d_dambreak_sycl/dambreak_sycl.cpp" in mk
    for name in ("test_3d_dambreak_cuda", "test_3d_dambreak_ck_cuda", "test_2d_dambreak_cuda", "test_3d_twisting_column_cuda",
                 "test_2d_taylor_green_cuda", "test_2d_dambreak_ck_cuda"):
        exe = os.path.join(ROOT, "integration", "_build", name)
        if not os.path.exists("/root/reference") and not os.path.exists(exe):
            pytest.skip("reference tree absent and no prebuilt binary")
        assert os.path.exists(exe), name
        if torch.cuda.is_available():
            continue  # a GPU is present: the run itself is covered by the gpu tests
        r = subprocess.run([exe, "--state_recording=0"], capture_output=True, text=True, cwd=tmp_path)
        assert r.returncode == 1 and "no CUDA device" in r.stdout + r.stderr, name


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "sphinxsys-learning_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in src and "import oracle" not in src and "from oracle" not in src, f


def test_lattice_sizes_and_kernel_kats(pk, oracle):
    """N_f = 3200 (2-D dp 0.025), 8000 (3-D dp 0.05); sigma0 / W0 closed forms"""
    c2 = pk.cases.DamBreak(oracle, ndim=2, dp=0.025)
    assert c2.water.n == 3200 and c2.water.info()["cells"] == [93, 93]
    i2 = c2.water.info()
    assert abs(i2["sigma0"] - 1616.7570433713395) < 1e-9 and abs(i2["W0"] - 527.377326221665) < 1e-9
    c3 = pk.cases.DamBreak(oracle, ndim=3, dp=0.05)
    assert c3.water.n == 8000 and c3.water.info()["cells"] == [52, 26, 14]
    i3 = c3.water.info()
    assert abs(i3["sigma0"] - 8076.006270278393) < 1e-8 and abs(i3["W0"] - 1521.2807487163413) < 1e-9
    counts, _ = c3.inner.neighbors()
    assert counts.max() == 80
    counts, _ = c2.inner.neighbors()
    assert counts.max() == 20
    # golden sample 0 of the reference: total mechanical energy of the initial 2-D column is exactly 1
    assert abs(c2.mechanical_energy() - 1.0000000000000016) < 1e-13
    # contact search depth formula of the reference (base_body_relation.cpp:10-16)
    assert oracle.load().orc_search_depth(c3.sys.h, c3.wall_contact.id, 0) in (1, 2)
    tc = pk.cases.TwistingColumn(oracle, dp=0.1)
    assert tc.column.n == 6100
    tg = pk.cases.TaylorGreen(oracle, dp=0.01)
    assert tg.water.n == 10000
    assert abs(tg.sys.total_kinetic_energy(tg.water) - 0.25) < 1e-12  # golden sample 0 = 0.25000000000000011


# ------------------------------------------------------------------------------------------------
# Run-level pin of the oracle: the FULL stored series of every fixture the reference's tests hold for the
# path (tests/golden), under (a) the north-star strict criterion - every sample within 1 % of the stored
# runs - where the fixture demonstrably comes from the current classic sources, and (b) the reference's
# own acceptance test (testResult: DTW distance <= 1.01 x stored, restated in oracle/validate_golden.py)
# everywhere.  DESIGN.md section 4 holds the per-case findings.
# ------------------------------------------------------------------------------------------------
def test_twisting_column_reproduces_the_stored_run_digit_for_digit():
    """all 3127 stored samples of the free-end observer (6 significant digits in the XML), position and
    velocity: equal to the fixture's own precision.  (SolidBody smoothing length 1.15 dp,
    bodies/predefined_bodies.h:44-54 - with 1.3 dp the series is 11 % off after 1200 steps.)"""
    from oracle import validate_golden as vg
    r = vg.twisting_column(3127)
    assert r["steps"] == 3126  # the reference's loop also ends after 3126 steps (t >= 0.5)
    assert r["MyObserver_Position"]["max_rel_dev"] < 2e-6 and r["MyObserver_Position"]["dtw_ok"], r
    assert r["MyObserver_Velocity"]["max_rel_dev"] < 1e-5 and r["MyObserver_Velocity"]["dtw_ok"], r


def test_dambreak_3d_full_series_within_one_percent():
    """the headline case: all 21 stored energy samples (t = 0..20) within 1 %, the six probe-pressure
    series (one sample per advection step) under the reference's DTW criterion"""
    from oracle import validate_golden as vg
    r = vg.dambreak_3d(21)
    e = r["WaterBody_TotalMechanicalEnergy"]
    assert e["snapshots"] == [21, 21] and e["max_rel_dev"] < 0.01 and e["dtw_ok"], r
    p = r["FluidObserver_Pressure"]
    assert p["dtw_ok"] and 0.8 * 2180 <= r["advection_steps"] <= 1.2 * 2180, r


def test_dambreak_2d_and_taylor_green_pass_the_reference_acceptance_test():
    """full series (23 / 51 samples) under RegressionTestDynamicTimeWarping::testResult; the strict 1 %
    window is asserted where it holds (dam break: samples 0-4, Taylor-Green: 0-1)"""
    from oracle import validate_golden as vg
    r = vg.dambreak_2d(23)
    assert r["WaterBody_TotalMechanicalEnergy"]["dtw_ok"] and r["FluidObserver_Pressure"]["dtw_ok"], r
    assert r["WaterBody_TotalMechanicalEnergy"]["first_sample_over_1pct"] >= 5, r
    assert abs(r["ours_energy"][0] - 1.0000000000000016) < 1e-13
    r = vg.taylor_green(51)
    assert r["WaterBody_TotalKineticEnergy"]["dtw_ok"] and r["WaterBody_MaximumSpeed"]["dtw_ok"], r
    assert r["WaterBody_TotalKineticEnergy"]["first_sample_over_1pct"] >= 2, r


def test_ck_dambreak_2d_passes_the_reference_acceptance_test():
    """the CK oracle over the full stored series of tests_sycl/2d_examples/test_2d_dambreak_sycl (23 samples, 4400 advection
    steps): DTW within 1.01 x stored for energy and probe.  (The 3-D twin takes 150 s on the CPU and runs on the GPU:
    tests/test_gpu_parity.py::test_ck_run_level_observables_against_the_sycl_fixtures; oracle result in DESIGN.md section 4.)"""
    from oracle import validate_golden as vg
    r = vg.dambreak_2d_ck(23)
    assert r["WaterBody_TotalMechanicalEnergy"]["dtw_ok"] and r["FluidObserver_Pressure"]["dtw_ok"], r
    assert r["WaterBody_TotalMechanicalEnergy"]["snapshots"] == [23, 23]
    assert abs(r["ours_energy"][0] - 1.0) < 1e-13


@pytest.mark.xfail(strict=True, reason="fixtures predate the current classic scheme (DESIGN.md section 4): the 2-D dam-break "
                   "energy file is byte-identical to the CK/SYCL twin's, the Taylor-Green runs were recorded with an "
                   "un-limited transport-velocity correction; kept visible, not hidden")
def test_strict_one_percent_on_the_stale_fixtures():
    from oracle import validate_golden as vg
    r2 = vg.dambreak_2d(23)
    rt = vg.taylor_green(51)
    assert r2["WaterBody_TotalMechanicalEnergy"]["max_rel_dev"] < 0.01 and rt["WaterBody_TotalKineticEnergy"]["max_rel_dev"] < 0.01


def test_taylor_green_fixture_vintage_is_reproduced():
    """oracle/experiments/tg_variants.py: with every switch at the current sources' setting the numpy
    restatement reproduces liboracle (so the restatement is not the variable); switching ONLY the
    transport-velocity limiter off (NoLimiter, the pre-TruncatedLinear scheme) moves the t = 0.1 sample
    from 2.1e-3 to < 1e-4 of the stored 0.20646315402108778 and the maximum speed from 7.8e-3 to < 5e-4"""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle", "experiments"))
    import tg_variants as tv
    from oracle import validate_golden as vg
    ke_gold, vmax_gold = tv.GOLD[0.1]
    (_, _, ke, vmax), = tv.sample({}, (0.1,))
    lib = vg.taylor_green(2)["ours_ke"][1]
    assert abs(ke - lib) < 1e-12 * lib
    assert abs(ke - ke_gold) / ke_gold > 1e-3 and abs(vmax - vmax_gold) / vmax_gold > 5e-3
    (_, _, ke, vmax), = tv.sample({"tvc": "plain"}, (0.1,))
    assert abs(ke - ke_gold) / ke_gold < 1e-4 and abs(vmax - vmax_gold) / vmax_gold < 5e-4


def test_reference_dtw_restatement_known_answers():
    """dtw_reference against hand-evaluated cases of dynamic_time_warping_method.hpp:16-55"""
    from oracle import validate_golden as vg
    a = np.arange(12.0)
    assert vg.dtw_reference(a, a) == 0.0
    assert vg.dtw_reference(a, a + 1.0) == 2.0     # band edge restarts the accumulation: far below the true 11 + ...
    assert vg.dtw_reference(a, np.arange(20.0)) == float("inf")  # length changed by > 20 %: the reference exits
    v = np.stack([a, a], 1)
    assert vg.dtw_reference(v, v) == 0.0


def test_ck_acoustic_oracle_agrees_with_an_independent_restatement(pk, oracle):
    """row f3: the C++ oracle of the CK dam break (oracle/oracle_ck.hpp: corrected acoustic steps on Compression /
    Displacement, tabulated kernel, LinearCorrectionMatrix(0.5), interval-triggered advection steps) against the numpy
    restatement that was written first and independently (oracle/experiments/dambreak2d_ck.py: k-d tree neighbours,
    vectorised pair sums): 200 acoustic steps = 40 advection steps of the 2-D main, every dt equal, fields to 1e-10"""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle", "experiments"))
    import dambreak2d_ck as nk
    dp = 0.05
    o = pk.cases.DamBreakCK(oracle, ndim=2, dp=dp, real_bytes=8, with_observer=False)
    m = nk.DamBreak2DCK(dp=dp)
    interval, present, adv = m.advection_dt(), 0.0, 0
    for it in range(200):
        dt = m.acoustic_dt()
        m.time += dt
        present += dt
        m.first_half(dt)
        m.second_half(dt)
        did = o.acoustic_step_2d()
        assert abs(o.dt - dt) <= 1e-12 * dt, (it, o.dt, dt)
        assert did == (present > interval)
        if did:
            present, interval = 0.0, m.advection_dt()
            adv += 1
            m.x = m.x + m.dpos
            m.relations()
            m.density_regularization()
            m.advection_setup()
            m.correction_matrix()
    assert adv >= 30
    n = len(m.x)
    for name, ref in (("Position", m.x), ("Velocity", m.v), ("Density", m.rho), ("Compression", m.comp), ("CompressionRate", m.crate),
                      ("Force", m.F), ("LinearCorrectionMatrix", m.B.reshape(n, -1))):
        got = pk.cases.by_original_id(o.water, o.water.get(name)).reshape(n, -1)
        ref = np.asarray(ref).reshape(n, -1)
        err = np.linalg.norm(got - ref) / max(np.linalg.norm(ref), 1e-300)
        assert err < 1e-10, (name, err)
    assert abs(o.mechanical_energy() - m.energy()) < 1e-12


@pytest.mark.parametrize("ndim,dp,weight,side", [(2, 0.025, 2.0, 0.5), (3, 0.05, 1.0, 0.25)])
def test_force_on_structure_oracle_hydrostatic_known_answer(pk, oracle, ndim, dp, weight, side):
    """row f4, a physical known-answer for the restatement of FSI::PressureForceFromFluid (no reference fixture pins it):
    the water column at rest with hydrostatic pressure rho g (H - y) pushes the tank down with its weight rho g V and the
    wall it leans on sideways with rho g H^2 / 2 (x width); nothing along z; no viscous force without motion.  The SPH sum
    reaches both within the kernel-truncation error of the boundary (4-7 % at these resolutions), with the right signs."""
    c = pk.cases.DamBreakCK(oracle, ndim=ndim, dp=dp, real_bytes=8, with_observer=False)
    w, s, S = c.water, c.wall, c.sys
    w.set("Pressure", (1.0 - w.get("Position")[:, 1]).astype(np.float64))
    rel = S.contact(s, [w])
    rel.update()
    S.ck_pressure_force_from_fluid(rel, False)
    total = s.get("PressureForceFromFluid").sum(0)
    assert abs(total[1] + weight) < 0.08 * weight and abs(total[0] + side) < 0.08 * side, total
    if ndim == 3:
        assert abs(total[2]) < 1e-12
    assert np.allclose(s.get("ForcePrior"), s.get("PressureForceFromFluid"))       # ForcePriorCK: prior started at zero
    S.ck_viscous_force_from_fluid(rel, False)
    assert np.abs(s.get("ViscousForceFromFluid")).max() == 0.0


def test_ck_correction_matrix_oracle_lattice_known_answer(pk, oracle):
    """LinearCorrectionMatrix on the undisturbed lattice: in the interior the moment matrix -sum r_ij (x) grad W_ij V_j is
    isotropic and within the kernel's discretisation error of the identity (h = 1.3 dp: 0.979), so its inverse is
    1.0214 I - off-diagonals at rounding level, the same on every interior particle"""
    c = pk.cases.DamBreakCK(oracle, ndim=3, dp=0.05, real_bytes=8, with_observer=False)
    c.sys.ck_linear_correction_matrix(c.inner, c.wall_contact, 0.0)
    B = c.water.get("LinearCorrectionMatrix").reshape(-1, 3, 3)
    pos = c.water.get("Position")
    interior = np.all((pos > 0.2) & (pos < np.array([1.8, 0.8, 0.3])), axis=1)
    assert interior.sum() > 500
    diag = B[interior][:, [0, 1, 2], [0, 1, 2]]
    assert np.abs(diag - diag[0, 0]).max() < 1e-12 and abs(diag[0, 0] - 1.0) < 0.03
    off = B[interior] - diag[:, :, None] * np.eye(3)
    assert np.abs(off).max() < 1e-12
