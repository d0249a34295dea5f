"""The reference's benchmark cases as data + driver loops, backend-agnostic.

Geometry and loop order restate the reference mains (host-side setup, not part of the GPU path):
  tests/2d_examples/test_2d_dambreak/Dambreak.cpp:8-75,166-220
  tests/3d_examples/test_3d_dambreak/dambreak.cpp:10-136,159-195
  tests/2d_examples/test_2d_taylor_green/taylor_green.cpp:12-57,155-192
  tests/3d_examples/test_3d_twisting_column/twisting_column.cpp:13-67,128-175
Lattice generator: for_3D_build/particle_generator/particle_generator_lattice_3d.cpp:12-25 (x-major,
last axis fastest, cell centres of a buffer-less mesh anchored at the expanded system lower bound).
Wall normals: geometries/base_geometry.cpp:45-59 over BinaryShapes::findClosestPoint (:118-140) and
GeometricBox::findClosestPoint (geometric_element.cpp:19-60).

A "backend" is any module exposing SPHSystem with the surface of host.py (the CUDA library) - the
tests pass oracle/binding.py as the checker backend.
"""
import math

import numpy as np


def lattice(lo, hi, dp, dtype=np.float64):
    """all lattice positions of the system box (expanded bounds lo/hi), reference order"""
    lo = np.asarray(lo, dtype=dtype)
    hi = np.asarray(hi, dtype=dtype)
    dp = dtype(dp)
    n = np.ceil((hi - lo) / dp).astype(np.int64)
    axes = [(lo[d] + np.arange(n[d]).astype(dtype) * dp) + dtype(0.5) * dp for d in range(len(lo))]
    grid = np.meshgrid(*axes, indexing="ij")
    return np.stack([g.ravel() for g in grid], axis=1)


def lattice_select(lo, hi, dp, dtype, keep, max_points=4_000_000, x_range=None):
    """lattice(lo, hi, dp)[keep(points)] evaluated in chunks of x-planes (same order, bounded memory);
    x_range = (x_lo, x_hi) keeps only the planes with x_lo <= x < x_hi (slab decomposition)"""
    lo = np.asarray(lo, dtype=dtype)
    hi = np.asarray(hi, dtype=dtype)
    dpt = dtype(dp)
    n = np.ceil((hi - lo) / dpt).astype(np.int64)
    axes = [(lo[d] + np.arange(n[d]).astype(dtype) * dpt) + dtype(0.5) * dpt for d in range(len(lo))]
    plane = int(np.prod(n[1:]))
    step = max(1, max_points // max(plane, 1))
    out = []
    planes = np.arange(int(n[0]))
    if x_range is not None:
        planes = planes[(axes[0] >= dtype(x_range[0])) & (axes[0] < dtype(x_range[1]))]
    for k in range(0, len(planes), step):
        grid = np.meshgrid(axes[0][planes[k:k + step]], *axes[1:], indexing="ij")
        pts = np.stack([g.ravel() for g in grid], axis=1)
        out.append(pts[keep(pts)])
    return np.concatenate(out, axis=0) if out else np.zeros((0, len(lo)), dtype)


def in_box(p, center, half):
    return np.all(np.abs(p - np.asarray(center, p.dtype)) <= np.asarray(half, p.dtype), axis=1)


def _box_closest(p, center, half):
    """GeometricBox::findClosestPoint in the box frame"""
    c = p - center
    clamped = np.clip(c, -half, half)
    inside = np.all(clamped == c, axis=1)
    d_side = half - np.abs(c)
    which = np.argmin(d_side, axis=1)  # first minimum, as the reference's strict '<' scan
    proj = c.copy()
    rows = np.arange(len(c))
    proj[rows, which] = np.where(c[rows, which] < 0, -half[which], half[which])
    out = np.where(inside[:, None], proj, clamped)
    return out + center


def shell_normals(p, center, half_outer, half_inner):
    """NormalDirectionFromBodyShape for 'outer box minus inner box' (all p are inside the shell)"""
    center = np.asarray(center, p.dtype)
    ho, hi_ = np.asarray(half_outer, p.dtype), np.asarray(half_inner, p.dtype)
    co = _box_closest(p, center, ho)
    ci = _box_closest(p, center, hi_)
    do = np.linalg.norm(p - co, axis=1)
    di = np.linalg.norm(p - ci, axis=1)
    closest = np.where((di <= do)[:, None], ci, co)  # later sub-shape wins ties (dist <= dist_min)
    disp = closest - p
    return disp / np.linalg.norm(disp, axis=1)[:, None]


# ------------------------------------------------------------------------------------------------
# dam break (2-D and 3-D)
# ------------------------------------------------------------------------------------------------
class DamBreak:
    def __init__(self, backend, ndim=3, dp=0.05, real_bytes=8, with_observer=True, fluid_box=None, tank=None, x_range=None,
                 defer_configuration=False, fused_loop=True, wall_h_ratio=1.15, widen_wall_cells=True, probes=None, **sys_kw):
        """x_range: keep only the lattice planes with x in [x_lo, x_hi) (one slab of a decomposed run);
        defer_configuration: stop before the first cell list / configuration (the slab driver finishes)"""
        dtype = np.float64 if real_bytes == 8 else np.float32
        self.ndim, self.dp = ndim, dp
        BW = dp * 4
        if ndim == 3:
            DL, DH, DW = tank or (5.366, 2.0, 0.5)
            LL, LH, LW = fluid_box or (2.0, 1.0, 0.5)
            tank_sz, water_sz = [DL, DH, DW], [LL, LH, LW]
            probes = probes or [[DL, y, 0.5 * DW] for y in (0.01, 0.1, 0.2, 0.24, 0.252, 0.266)]
        else:
            DL, DH = tank or (5.366, 5.366)
            LL, LH = fluid_box or (2.0, 1.0)
            tank_sz, water_sz = [DL, DH], [LL, LH]
            probes = probes or [[DL, 0.2]]
        if probes is not None:   # other observation points than the reference main's (tests)
            probes = [list(p) for p in probes]
        self.probes = probes
        self.rho0, self.g = 1.0, 1.0
        self.U_ref = 2.0 * math.sqrt(self.g * water_sz[1])
        self.c0 = 10.0 * self.U_ref
        self.gravity = [0.0, -self.g] + ([0.0] if ndim == 3 else [])
        lo_user = [-BW] * ndim
        hi_user = [t + BW for t in tank_sz]
        self.sys = backend.SPHSystem(ndim, lo_user, hi_user, dp, real_bytes=real_bytes, **sys_kw)
        half_w = np.asarray(water_sz) * 0.5
        half_in = np.asarray(tank_sz) * 0.5
        half_out = half_in + BW
        water = lattice_select(self.sys.lo, self.sys.hi, dp, dtype, lambda p: in_box(p, half_w, half_w), x_range=x_range)
        shell = lattice_select(self.sys.lo, self.sys.hi, dp, dtype,
                               lambda p: in_box(p, half_in, half_out) & ~in_box(p, half_in, half_in), x_range=x_range)
        vol = dtype(dp) ** ndim
        # ---- fluid body (materials/base_material.cpp:35-41: rho = rho0, m = rho0 * Vol)
        self.water = w = self.sys.add_body("WaterBody", len(water), dp)
        w.set_fluid(self.rho0, self.c0, 0.0)
        w.set("Position", water)
        w.set("VolumetricMeasure", np.full(len(water), vol, dtype))
        w.set("Mass", np.full(len(water), dtype(self.rho0) * vol, dtype))
        w.set("Density", np.full(len(water), self.rho0, dtype))
        for name, width in (("Pressure", 1), ("DensityChangeRate", 1), ("Velocity", ndim), ("Force", ndim),
                            ("ForcePrior", ndim), ("PreviousGravityForce", ndim)):
            w.register(name, width, True)
        # ---- wall body (Solid: rho0 = 1, static)
        # The reference's wall is a SolidBody, smoothing length 1.15 dp (predefined_bodies.h:44-54).  The fluid-wall contact
        # uses the LARGER of the two kernels (neighborhood.cpp:75-80), i.e. the fluid's, so the wall's own h only sets the
        # cell size of the wall's cell list, which neighbour sets do not depend on.  On the CUDA library the wall's cells
        # are widened to the fluid's cut-off (sphb200_body_set_cell_spacing): the contact search then takes the
        # single-pass masked path instead of the two-pass generic one.  The oracle keeps the reference's cells.
        self.wall = s = self.sys.add_body("WallBoundary", len(shell), dp, h_ratio=wall_h_ratio)
        if widen_wall_cells and hasattr(s, "set_cell_spacing") and wall_h_ratio < w.h_ratio:
            s.set_cell_spacing(w.info()["rc"])
        s.set_fluid(1.0, 1.0, 0.0)
        s.set("Position", shell)
        s.set("VolumetricMeasure", np.full(len(shell), vol, dtype))
        s.set("Mass", np.full(len(shell), vol, dtype))
        s.set("NormalDirection", shell_normals(shell, half_in, half_out, half_in).astype(dtype))
        s.register("Velocity", ndim, True)
        s.register("Acceleration", ndim, True)
        self.inner = self.sys.inner(w)
        self.wall_contact = self.sys.contact(w, [s])
        self.observer = self.obs_contact = None
        if with_observer:
            self.observer = o = self.sys.add_body("FluidObserver", len(probes), dp)
            o.set("Position", np.asarray(probes, dtype))
            self.obs_contact = self.sys.contact(o, [w])
        self.time = 0.0
        self.n_advection = 0
        self.n_acoustic = 0
        self.dt = 0.0
        self.fused_loop = fused_loop
        self._substeps_guess = 6
        if defer_configuration:
            return
        # initializeSystemCellLinkedLists / Configurations; gravity prior force (dambreak.cpp:131-136)
        w.update_cell_list()
        s.update_cell_list()
        if self.observer:
            self.observer.update_cell_list()
        self.update_configuration(observer=True)
        self.sys.gravity_force(w, self.gravity)
        self.time = 0.0
        self.n_advection = 0
        self.n_acoustic = 0
        self.dt = 0.0

    def update_configuration(self, observer=True):
        self.inner.update()
        self.wall_contact.update()
        if observer and self.obs_contact:
            self.obs_contact.update()

    def acoustic_substep(self, dt):
        S = self.sys
        S.integration_1st_half(self.inner, self.wall_contact, True, dt)
        S.integration_2nd_half(self.inner, self.wall_contact, True, dt)
        self.n_acoustic += 1

    def advection_step(self, sort_interval=100, observer=True, before_configuration=None, reconfigure=True):
        """one outer iteration of dambreak.cpp:159-195; returns (Dt, number of acoustic sub-steps).
        reconfigure=False leaves out the closing cell-list / relation update (a caller that replaces the state from the
        host before the next step rebuilds the configuration then anyway).
        before_configuration(): called between the sub-step loop and the configuration update, where
        the 2-D main samples its observables (Dambreak.cpp:196-207, i.e. with the previous step's
        observer configuration)"""
        S, w = self.sys, self.water
        Dt = S.advection_dt(w, self.U_ref)
        S.density_summation(self.inner, self.wall_contact, free_surface=True)
        relaxation, k = 0.0, 0
        if self.fused_loop and hasattr(S, "acoustic_relaxation"):
            # the same loop, resident on the device (sphb200_acoustic_relaxation): no host round trip per sub-step
            k, self.dt, _, relaxation = S.acoustic_relaxation(self.inner, self.wall_contact, True, True, Dt, 0.6,
                                                              order=1 if self.ndim == 3 else 0, dt_carry=self.dt,
                                                              max_substeps=self._substeps_guess)
            self._substeps_guess = max(k, 1)   # steady state: exactly the launches that are needed, none that returns at once
            self.time += relaxation
            self.n_acoustic += k
        while relaxation < Dt:
            if self.ndim == 3:  # dambreak.cpp:168-177: step with the previous dt (0 at start), then reduce
                self.acoustic_substep(self.dt)
                self.dt = S.acoustic_dt(w)
            else:  # Dambreak.cpp:183-192: reduce first, then step
                self.dt = S.acoustic_dt(w)
                self.acoustic_substep(self.dt)
            relaxation += self.dt
            self.time += self.dt
            k += 1
        if before_configuration is not None:
            before_configuration()
        self.n_advection += 1
        if not reconfigure:
            return Dt, k
        if sort_interval and self.n_advection % sort_interval == 0 and self.n_advection != 1:
            w.sort()
        w.update_cell_list()
        self.update_configuration(observer)
        return Dt, k

    def mechanical_energy(self):
        return self.sys.total_mechanical_energy(self.water, self.gravity)

    def probe_pressure(self):
        return self.sys.observe(self.obs_contact, "Pressure", 1)


class DamBreakCK(DamBreak):
    """The CK-flavour dam break: tests/tests_sycl/3d_examples/test_3d_dambreak_sycl/dambreak.cpp:183-222 (ndim = 3) and
    tests/tests_sycl/2d_examples/test_2d_dambreak_sycl/dambreak_sycl.cpp:170-240 (ndim = 2).  Same bodies as the classic
    case; the dynamics are CompressionSummation + DensityRegularization, AdvectionStepSetup, the spatial-temporal free
    surface indication (3-D main only), LinearCorrectionMatrix(0.5) and the corrected acoustic steps on
    (Compression, CompressionRate, Displacement), all with the tabulated kernel."""

    def __init__(self, backend, ndim=3, dp=0.05, real_bytes=8, with_observer=True, **kw):
        super().__init__(backend, ndim=ndim, dp=dp, real_bytes=real_bytes, with_observer=with_observer, **kw)
        w = self.water
        for name, width in (("Displacement", ndim), ("Compression", 1), ("CompressionRate", 1)):
            w.register(name, width, True)
        w.set("Compression", np.ones(w.n, np.float64 if real_bytes == 8 else np.float32))
        self.advection_steps = 1
        if ndim == 2:   # dambreak_sycl.cpp:150-156: the first advection step's preparation happens before the loop
            self._prepare_2d()
            self._interval, self._present = self.sys.ck_advection_dt(w, self.U_ref), 0.0

    def _prepare_2d(self):
        S = self.sys
        S.ck_density_regularization(self.inner, self.wall_contact, True)
        S.ck_advection_step_setup(self.water)
        S.ck_linear_correction_matrix(self.inner, self.wall_contact, 0.5)

    def advection_step(self, sort_interval=100, observer=True, before_configuration=None):
        """3-D main: one outer iteration (dambreak.cpp:187-222); returns (Dt, sub-steps)"""
        assert self.ndim == 3
        S, w = self.sys, self.water
        S.ck_density_regularization(self.inner, self.wall_contact, True)
        S.ck_advection_step_setup(w)
        Dt = S.ck_advection_dt(w, self.U_ref)
        S.ck_free_surface_indication(self.inner, self.wall_contact)
        S.ck_linear_correction_matrix(self.inner, self.wall_contact, 0.5)
        relaxation, k = 0.0, 0
        while relaxation < Dt:
            self.dt = S.ck_acoustic_dt(w)
            S.ck_acoustic_1st_half(self.inner, self.wall_contact, self.dt)
            S.ck_acoustic_2nd_half(self.inner, self.wall_contact, self.dt)
            relaxation += self.dt
            self.time += self.dt
            k += 1
        self.n_acoustic += k
        S.ck_update_particle_position(w)
        if before_configuration is not None:
            before_configuration()
        self.n_advection += 1
        if sort_interval and self.n_advection % sort_interval == 0 and self.n_advection != 1:
            w.sort()
        w.update_cell_list()
        self.update_configuration(observer)
        return Dt, k

    def acoustic_step_2d(self, on_advection=None):
        """2-D main: one iteration of its single loop (dambreak_sycl.cpp:176-228): an acoustic step, and the advection-step
        work whenever the accumulated time passes the interval; returns True when an advection step was completed"""
        assert self.ndim == 2
        S, w = self.sys, self.water
        self.dt = S.ck_acoustic_dt(w)
        self.time += self.dt
        self._present += self.dt
        S.ck_acoustic_1st_half(self.inner, self.wall_contact, self.dt)
        S.ck_acoustic_2nd_half(self.inner, self.wall_contact, self.dt)
        self.n_acoustic += 1
        if self._present <= self._interval:
            return False
        self._present, self._interval = 0.0, S.ck_advection_dt(w, self.U_ref)
        self.advection_steps += 1
        self.n_advection += 1
        S.ck_update_particle_position(w)
        if on_advection is not None:
            on_advection()
        if self.advection_steps % 100:   # sic (dambreak_sycl.cpp:214): order only, the results do not depend on it
            w.sort()
        w.update_cell_list()
        self.update_configuration(self.observer is not None)
        self._prepare_2d()
        return True


# ------------------------------------------------------------------------------------------------
# 2-D Taylor-Green vortex
# ------------------------------------------------------------------------------------------------
class TaylorGreen:
    def __init__(self, backend, dp=0.01, real_bytes=8, fused_loop=True, **sys_kw):
        dtype = np.float64 if real_bytes == 8 else np.float32
        self.fused_loop, self._substeps_guess = fused_loop, 6
        self.rho0, self.U_f, self.c0 = 1.0, 1.0, 10.0
        self.mu = self.rho0 * self.U_f * 1.0 / 100.0
        self.sys = backend.SPHSystem(2, [0.0, 0.0], [1.0, 1.0], dp, real_bytes=real_bytes, **sys_kw)
        pts = lattice(self.sys.lo, self.sys.hi, dp, dtype)
        pts = pts[in_box(pts, [0.5, 0.5], [0.5, 0.5])]
        n = len(pts)
        vol = dtype(dp) ** 2
        self.water = w = self.sys.add_body("WaterBody", n, dp)
        w.set_fluid(self.rho0, self.c0, self.mu)
        w.set_periodic(0, 0.0, 1.0)
        w.set_periodic(1, 0.0, 1.0)
        w.set("Position", pts)
        w.set("VolumetricMeasure", np.full(n, vol, dtype))
        w.set("Mass", np.full(n, dtype(self.rho0) * vol, dtype))
        w.set("Density", np.full(n, self.rho0, dtype))
        x, y = pts[:, 0].astype(np.float64), pts[:, 1].astype(np.float64)
        vel = np.stack([-np.cos(2 * np.pi * x) * np.sin(2 * np.pi * y), np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y)], 1)
        w.set("Velocity", vel.astype(dtype))
        for name, width in (("Pressure", 1), ("DensityChangeRate", 1), ("Force", 2), ("ForcePrior", 2),
                            ("PreviousViscousForce", 2)):
            w.register(name, width, True)
        self.inner = self.sys.inner(w)
        self.configure()
        self.time, self.n_advection, self.n_acoustic = 0.0, 0, 0

    def configure(self):
        w = self.water
        w.update_cell_list()
        w.periodic_cell_list(0)
        w.periodic_cell_list(1)
        self.inner.update()

    def advection_step(self):
        S, w = self.sys, self.water
        Dt = S.advection_dt(w, self.U_f, viscous=True)
        S.density_summation(self.inner, None, free_surface=False)
        S.viscous_force(self.inner)
        S.transport_velocity_correction(self.inner, 0.2)
        relaxation, k = 0.0, 0
        if self.fused_loop and hasattr(S, "acoustic_relaxation"):
            k, _, _, relaxation = S.acoustic_relaxation(self.inner, None, True, False, Dt, 0.6, order=2,
                                                        max_substeps=self._substeps_guess)
            self._substeps_guess = max(k, 1)   # steady state: exactly the launches that are needed, none that returns at once
            self.time += relaxation
        while relaxation < Dt:
            dt = min(S.acoustic_dt(w), Dt)
            relaxation += dt
            S.integration_1st_half(self.inner, None, True, dt)
            S.integration_2nd_half(self.inner, None, False, dt)
            self.time += dt
            k += 1
        self.n_acoustic += k
        self.n_advection += 1
        w.periodic_bounding(0)
        w.periodic_bounding(1)
        self.configure()
        return Dt, k


# ------------------------------------------------------------------------------------------------
# 3-D twisting column (total-Lagrangian Neo-Hookean solid)
# ------------------------------------------------------------------------------------------------
class TwistingColumn:
    def __init__(self, backend, dp=0.1, real_bytes=8, decomposed=True, **sys_kw):
        dtype = np.float64 if real_bytes == 8 else np.float32
        PL, PH, PW = 6.0, 1.0, 1.0
        SL = dp * 1.0
        self.decomposed = decomposed
        self.sys = backend.SPHSystem(3, [-SL, -0.5 * PH, -0.5 * PW], [PL, 0.5 * PH, 0.5 * PW], dp, real_bytes=real_bytes, **sys_kw)
        pts = lattice(self.sys.lo, self.sys.hi, dp, dtype)
        col = in_box(pts, [0.5 * (PL - SL), 0, 0], [0.5 * (PL + SL), 0.5 * PH, 0.5 * PW])
        hold = in_box(pts, [-0.5 * SL, 0, 0], [0.5 * SL, 0.5 * PH, 0.5 * PW])
        pts = pts[col | hold]
        n = len(pts)
        self.holder_ids = np.nonzero(in_box(pts, [-0.5 * SL, 0, 0], [0.5 * SL, 0.5 * PH, 0.5 * PW]))[0]
        rho0 = 1100.0
        vol = dtype(dp) ** 3
        self.column = c = self.sys.add_body("Column", n, dp, h_ratio=1.15)  # SolidBody: bodies/predefined_bodies.h:44-54
        c.set_elastic(rho0, 1.7e7, 0.45)
        c.set("Position", pts)
        c.set("InitialPosition", pts)
        c.set("VolumetricMeasure", np.full(n, vol, dtype))
        c.set("Mass", np.full(n, dtype(rho0) * vol, dtype))
        c.set("Density", np.full(n, rho0, dtype))
        x, y, z = (pts[:, k].astype(np.float64) for k in range(3))
        omega = -400.0 * np.sin(np.pi * x / (2.0 * PL))
        radius = np.sqrt(y ** 2 + z ** 2)
        ang = np.arctan2(y, z)
        vel = np.zeros((n, 3))
        on = x > 0.0
        vel[on, 1] = (omega * radius * np.cos(ang))[on]
        vel[on, 2] = (-omega * radius * np.sin(ang))[on]
        c.set("Velocity", vel.astype(dtype))
        c.set("DeformationGradient", np.tile(np.eye(3).ravel(), (n, 1)).astype(dtype))
        for name, width in (("DeformationRate", 9), ("Force", 3), ("ForcePrior", 3)):
            c.register(name, width, True)
        for name, width in (("LinearGradientCorrectionMatrix", 9), ("StressOnParticle", 9), ("StressPK1OnParticle", 9),
                            ("InverseTransposedDeformation", 9), ("DeterminantTerm", 1)):
            c.register(name, width, True)  # sorted along with the rest (the solid is sorted once, before use)
        self.observer = o = self.sys.add_body("MyObserver", 1, dp)
        o.set("Position", np.asarray([[PL, 0.0, 0.0]], dtype))
        self.inner = self.sys.inner(c)
        self.obs_contact = self.sys.contact(o, [c])
        c.update_cell_list()
        o.update_cell_list()
        self.inner.update()
        self.obs_contact.update()
        self.sys.correction_matrix(self.inner)
        self.time, self.dt, self.n_steps = 0.0, 0.0, 0

    def step(self):
        S = self.sys
        S.solid_1st_half(self.inner, self.decomposed, self.dt)
        S.solid_constrain(self.column, self.holder_ids)
        S.solid_2nd_half(self.inner, self.dt)
        self.n_steps += 1
        self.dt = S.solid_dt(self.column, 0.5)
        self.time += self.dt
        return self.dt

    def observed(self):
        return (self.sys.observe(self.obs_contact, "Position", 3)[0], self.sys.observe(self.obs_contact, "Velocity", 3)[0])


def by_original_id(body, values):
    """reorder a per-particle array from the body's current order to original-id order"""
    oid, _ = body.ids()
    out = np.empty_like(values)
    out[oid] = values
    return out
