"""Host-side mirror (Python) of the reference's interface for the hot path, over the C ABI.

The production host binding is the C++ adapter in include/sphb200_adapter.hpp (the reference is C++);
this module is the same surface for the test-suite and bench harness.  Method names follow the
reference's operators so the parity tests read like the reference's own mains:

  SPHSystem            sphinxsys_system/sph_system.h:50-154
  Body                 bodies/base_body.h:62-201 + particles/base_particles.h:80-263 (named variables)
  InnerRelation / ContactRelation        body_relations/*.h  (updateConfiguration)
  density_summation / integration_1st_half / ...   particle_dynamics/fluid_dynamics/* exec(dt)

Every call goes to libsphb200.so; nothing is computed in Python.
"""
import ctypes as C

import numpy as np

from . import capi


def _np_real(real_bytes):
    return np.float64 if real_bytes == 8 else np.float32


class SPHSystem:
    """One context per GPU. `lo`/`hi` are the user's system_domain_bounds; like the reference's ctor
    (sph_system.cpp:39) they are expanded by 4*dp_ref."""

    backend = "sphb200"

    def __init__(self, ndim, lo, hi, dp_ref, real_bytes=8, device=0, expand=True, stream=None):
        self.lib = capi.load()
        self.ndim, self.real_bytes, self.dtype = int(ndim), int(real_bytes), _np_real(real_bytes)
        ex = 4.0 * dp_ref if expand else 0.0
        self.lo = np.asarray(lo, dtype=np.float64) - ex
        self.hi = np.asarray(hi, dtype=np.float64) + ex
        lo3 = (C.c_double * 3)(*list(self.lo) + [0.0] * (3 - ndim))
        hi3 = (C.c_double * 3)(*list(self.hi) + [0.0] * (3 - ndim))
        h = C.c_void_p()
        st = self.lib.sphb200_ctx_create(device, ndim, real_bytes, lo3, hi3, C.byref(h))
        if st != capi.OK:
            why = {capi.ERR_NO_DEVICE: "no CUDA device (there is no CPU fallback)",
                   capi.ERR_INVALID: "invalid arguments", capi.ERR_CUDA: "CUDA error"}.get(st, "?")
            raise capi.SphB200Error(st, f"ctx_create failed: {why}")
        self.h = h
        self.bodies, self.relations = [], []
        if stream is not None:
            self.check(self.lib.sphb200_ctx_set_stream(self.h, C.c_void_p(int(stream))))

    def check(self, st):
        if st != capi.OK:
            raise capi.SphB200Error(st, self.lib.sphb200_last_error(self.h).decode())

    def close(self):
        if self.h:
            self.lib.sphb200_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_pair_kernel(self, mode):
        """-1 automatic, 0 list-per-particle, 1 group-tiled (include/sphb200.h)"""
        self.check(self.lib.sphb200_ctx_set_pair_kernel(self.h, int(mode)))

    def synchronize(self):
        self.check(self.lib.sphb200_ctx_synchronize(self.h))

    def stage_wait(self):
        self.check(self.lib.sphb200_ctx_stage_wait(self.h))

    def fetch_wait(self):
        self.check(self.lib.sphb200_ctx_fetch_wait(self.h))

    @property
    def stream(self):
        return self.lib.sphb200_ctx_stream(self.h)

    @property
    def launches(self):
        return int(self.lib.sphb200_ctx_launch_count(self.h))

    def add_body(self, name, n, dp, h_ratio=1.3):
        b = Body(self, name, n, dp, h_ratio)
        self.bodies.append(b)
        return b

    def inner(self, body):
        r = Relation(self, body, None)
        self.relations.append(r)
        return r

    def contact(self, body, targets):
        r = Relation(self, body, list(targets))
        self.relations.append(r)
        return r

    def self_contact(self, body, part_ids=None):
        """SelfSurfaceContactRelation over a body part by particle (original ids), or the whole body"""
        r = Relation(self, body, None, builder="self", part_ids=part_ids)
        self.relations.append(r)
        return r

    def surface_contact(self, body, targets, part_ids=None):
        """SurfaceContactRelation: contact builder with the kernel of the averaged smoothing length"""
        r = Relation(self, body, list(targets), builder="surface", part_ids=part_ids)
        self.relations.append(r)
        return r

    def exclusive_scan(self, values):
        v = np.ascontiguousarray(values, dtype=np.uint32)
        out = np.zeros_like(v)
        total = C.c_uint64()
        self.check(self.lib.sphb200_exclusive_scan_u32(
            self.h, v.ctypes.data_as(C.POINTER(C.c_uint32)), out.ctypes.data_as(C.POINTER(C.c_uint32)),
            v.size, C.byref(total)))
        return out, int(total.value)

    def exclusive_scan_static(self, d_in_ptr, d_out_ptr, n):
        """sphb200_exclusive_scan_u32_static: device pointers (n in, n + 1 out), scratch in a static device array, no allocation"""
        self.check(self.lib.sphb200_exclusive_scan_u32_static(self.h, C.c_void_p(d_in_ptr), C.c_void_p(d_out_ptr), int(n)))

    # ---- dynamics (exec) ---------------------------------------------------------------------
    def density_summation(self, inner, contact=None, free_surface=False):
        self.check(self.lib.sphb200_density_summation(inner.h, contact.h if contact else None, int(free_surface)))

    def integration_1st_half(self, inner, wall=None, riemann=True, dt=0.0):
        self.check(self.lib.sphb200_integration_1st_half(inner.h, wall.h if wall else None, int(riemann), dt))

    def integration_1st_half_init(self, inner, dt=0.0):
        self.check(self.lib.sphb200_integration_1st_half_init(inner.h, dt))

    def integration_1st_half_interact(self, inner, wall=None, riemann=True, dt=0.0):
        self.check(self.lib.sphb200_integration_1st_half_interact(inner.h, wall.h if wall else None, int(riemann), dt))

    def integration_2nd_half(self, inner, wall=None, riemann=True, dt=0.0):
        self.check(self.lib.sphb200_integration_2nd_half(inner.h, wall.h if wall else None, int(riemann), dt))

    def acoustic_dt(self, body, cfl=0.6):
        out = C.c_double()
        self.check(self.lib.sphb200_acoustic_dt(body.h, cfl, C.byref(out)))
        return out.value

    def advection_dt(self, body, U_ref, cfl=0.25, viscous=False):
        out = C.c_double()
        self.check(self.lib.sphb200_advection_dt(body.h, U_ref, cfl, int(viscous), C.byref(out)))
        return out.value

    def advection_dt_device(self, body, U_ref, cfl=0.25, viscous=False):
        """Dt stays on the device (no host synchronisation); acoustic_relaxation(Dt=None) uses it"""
        self.check(self.lib.sphb200_advection_dt(body.h, U_ref, cfl, int(viscous), None))

    def acoustic_relaxation(self, inner, wall=None, riemann_1st=True, riemann_2nd=True, Dt=None, cfl=0.6, order=1, dt_carry=-1.0,
                            max_substeps=8, wait=True, resume=False):
        """the sub-step loop of one advection step on the device; wait=True returns (sub-steps, last dt, time, relaxation_time)"""
        out = (C.c_double * 5)() if wait else None
        self.check(self.lib.sphb200_acoustic_relaxation(inner.h, wall.h if wall else None, int(riemann_1st), int(riemann_2nd),
                                                        -1.0 if Dt is None else Dt, cfl, int(order), dt_carry, int(max_substeps),
                                                        int(resume), out))
        if wait:
            return int(out[0]), out[1], out[2], out[3]

    def relaxation_begin(self, Dt, dt_carry=-1.0):
        self.check(self.lib.sphb200_relaxation_begin(self.h, Dt, dt_carry))

    def substep_ctl(self, body, cfl=0.6, order=1, phase=1):
        self.check(self.lib.sphb200_substep_ctl(body.h, cfl, int(order), int(phase)))

    def reduce_slot_ptr(self, slot=0):
        p = C.c_void_p()
        self.check(self.lib.sphb200_reduce_slot_ptr(self.h, int(slot), C.byref(p)))
        return p.value

    def rows_gather(self, src_ptr, row_bytes, idx_ptr, n, dst_ptr):
        self.check(self.lib.sphb200_rows_gather(self.h, C.c_void_p(src_ptr), int(row_bytes), C.c_void_p(idx_ptr), int(n), C.c_void_p(dst_ptr)))

    def rows_scatter(self, dst_ptr, row_bytes, idx_ptr, n, src_ptr):
        self.check(self.lib.sphb200_rows_scatter(self.h, C.c_void_p(dst_ptr), int(row_bytes), C.c_void_p(idx_ptr), int(n), C.c_void_p(src_ptr)))

    # ---- halo records over peer memory (csrc/peer.cu)
    def peer_alloc(self, nbytes):
        p = C.c_void_p()
        self.check(self.lib.sphb200_peer_alloc(self.h, int(nbytes), C.byref(p)))
        return p.value

    def peer_free(self, ptr):
        self.check(self.lib.sphb200_peer_free(self.h, C.c_void_p(ptr)))

    def peer_export(self, ptr):
        buf = C.create_string_buffer(64)
        self.check(self.lib.sphb200_peer_export(self.h, C.c_void_p(ptr), buf))
        return buf.raw

    def peer_open(self, handle):
        p = C.c_void_p()
        self.check(self.lib.sphb200_peer_open(self.h, C.create_string_buffer(bytes(handle), 64), C.byref(p)))
        return p.value

    def peer_close(self, ptr):
        self.check(self.lib.sphb200_peer_close(self.h, C.c_void_p(ptr)))

    def rows_push(self, src_ptr, row_bytes, idx_ptr, n, peer_landing, peer_flag, epoch):
        self.check(self.lib.sphb200_rows_push(self.h, C.c_void_p(src_ptr), int(row_bytes), C.c_void_p(idx_ptr), int(n),
                                              C.c_void_p(peer_landing), C.c_void_p(peer_flag), int(epoch)))

    def rows_wait_scatter(self, dst_ptr, row_bytes, idx_ptr, n, landing, flag, epoch):
        self.check(self.lib.sphb200_rows_wait_scatter(self.h, C.c_void_p(dst_ptr), int(row_bytes), C.c_void_p(idx_ptr), int(n),
                                                      C.c_void_p(landing), C.c_void_p(flag), int(epoch)))

    def peer_timeouts(self):
        out = C.c_int64()
        self.check(self.lib.sphb200_peer_timeouts(self.h, C.byref(out)))
        return out.value

    def relaxation_state(self):
        out = (C.c_double * 5)()
        self.check(self.lib.sphb200_relaxation_state(self.h, out))
        return {"substeps": int(out[0]), "dt": out[1], "time": out[2], "relaxation_time": out[3], "done": bool(out[4])}

    def gravity_force(self, body, g):
        g3 = (C.c_double * 3)(*list(g) + [0.0] * (3 - len(g)))
        self.check(self.lib.sphb200_gravity_force(body.h, g3))

    def viscous_force(self, inner, wall=None):
        self.check(self.lib.sphb200_viscous_force(inner.h, wall.h if wall else None))

    def transport_velocity_correction(self, inner, coefficient=0.2):
        self.check(self.lib.sphb200_transport_velocity_correction(inner.h, coefficient))

    def ck_density_regularization(self, inner, contact=None, free_surface=True):
        self.check(self.lib.sphb200_ck_density_regularization(inner.h, contact.h if contact else None, int(free_surface)))

    def ck_linear_correction_matrix(self, inner, contact=None, alpha=0.0):
        self.check(self.lib.sphb200_ck_linear_correction_matrix(inner.h, contact.h if contact else None, alpha))

    def ck_free_surface_indication(self, inner, contact=None):
        self.check(self.lib.sphb200_ck_free_surface_indication(inner.h, contact.h if contact else None))

    def ck_compression_summation(self, inner, contact=None):
        self.check(self.lib.sphb200_ck_compression_summation(inner.h, contact.h if contact else None))

    def ck_density_regularize(self, body, free_surface=True):
        self.check(self.lib.sphb200_ck_density_regularize(body.h, int(free_surface)))

    def ck_pressure_force_from_fluid(self, contact, corrected=True):
        self.check(self.lib.sphb200_ck_pressure_force_from_fluid(contact.h, int(corrected)))

    def ck_viscous_force_from_fluid(self, contact, corrected=True):
        self.check(self.lib.sphb200_ck_viscous_force_from_fluid(contact.h, int(corrected)))

    def ck_advection_step_setup(self, body):
        self.check(self.lib.sphb200_ck_advection_step_setup(body.h))

    def ck_update_particle_position(self, body):
        self.check(self.lib.sphb200_ck_update_particle_position(body.h))

    def ck_acoustic_dt(self, body, cfl=0.6):
        out = C.c_double()
        self.check(self.lib.sphb200_ck_acoustic_dt(body.h, cfl, C.byref(out)))
        return out.value

    def ck_advection_dt(self, body, U_ref, cfl=0.25):
        out = C.c_double()
        self.check(self.lib.sphb200_ck_advection_dt(body.h, U_ref, cfl, C.byref(out)))
        return out.value

    def ck_acoustic_1st_half(self, inner, wall=None, dt=0.0):
        self.check(self.lib.sphb200_ck_acoustic_step_1st_half(inner.h, wall.h if wall else None, dt))

    def ck_acoustic_2nd_half(self, inner, wall=None, dt=0.0):
        self.check(self.lib.sphb200_ck_acoustic_step_2nd_half(inner.h, wall.h if wall else None, dt))

    def total_kinetic_energy(self, body):
        out = C.c_double()
        self.check(self.lib.sphb200_total_kinetic_energy(body.h, C.byref(out)))
        return out.value

    def total_mechanical_energy(self, body, g):
        g3 = (C.c_double * 3)(*list(g) + [0.0] * (3 - len(g)))
        out = C.c_double()
        self.check(self.lib.sphb200_total_mechanical_energy(body.h, g3, C.byref(out)))
        return out.value

    def maximum_speed(self, body):
        out = C.c_double()
        self.check(self.lib.sphb200_maximum_speed(body.h, C.byref(out)))
        return out.value

    def position_bound(self, body, upper=True):
        out = (C.c_double * 3)()
        self.check(self.lib.sphb200_position_bound(body.h, int(upper), out))
        return np.asarray(list(out)[:self.ndim])

    def observe(self, contact, name, width=1):
        out = np.zeros(contact.body.n * width, dtype=np.float64)
        self.check(self.lib.sphb200_observe(contact.h, name.encode(), width, out.ctypes.data_as(C.POINTER(C.c_double))))
        return out.reshape(contact.body.n, width) if width > 1 else out

    def correction_matrix(self, inner):
        self.check(self.lib.sphb200_correction_matrix(inner.h))

    def solid_1st_half(self, inner, decomposed, dt):
        self.check(self.lib.sphb200_solid_1st_half(inner.h, int(decomposed), dt))

    def solid_2nd_half(self, inner, dt):
        self.check(self.lib.sphb200_solid_2nd_half(inner.h, dt))

    def solid_constrain(self, body, original_ids):
        ids = np.ascontiguousarray(original_ids, dtype=np.int64)
        self.check(self.lib.sphb200_solid_constrain(body.h, ids.ctypes.data_as(C.POINTER(C.c_int64)), ids.size))

    def solid_dt(self, body, cfl=0.6):
        out = C.c_double()
        self.check(self.lib.sphb200_solid_dt(body.h, cfl, C.byref(out)))
        return out.value


class Body:
    def __init__(self, system, name, n, dp, h_ratio):
        self.sys, self.lib, self.name = system, system.lib, name
        self.n, self.dp, self.h_ratio = int(n), float(dp), float(h_ratio)
        h = C.c_void_p()
        system.check(self.lib.sphb200_body_create(system.h, name.encode(), self.n, dp, h_ratio, C.byref(h)))
        self.h = h
        self.widths = {"Position": system.ndim, "VolumetricMeasure": 1}

    def set_cell_spacing(self, spacing):
        """widen the cells of this body's cell list (sphb200_body_set_cell_spacing); neighbour sets do not depend on it"""
        self.sys.check(self.lib.sphb200_body_set_cell_spacing(self.h, float(spacing)))

    def set_fluid(self, rho0, c0, mu=0.0):
        self.sys.check(self.lib.sphb200_body_set_fluid(self.h, rho0, c0, mu))

    def set_elastic(self, rho0, youngs, poisson):
        self.sys.check(self.lib.sphb200_body_set_elastic(self.h, rho0, youngs, poisson))

    def set_periodic(self, axis, lo, hi):
        self.sys.check(self.lib.sphb200_body_set_periodic(self.h, axis, lo, hi))

    def info(self):
        out = (C.c_double * 12)()
        self.sys.check(self.lib.sphb200_body_info(self.h, out))
        o = list(out)
        d = self.sys.ndim
        return {"h": o[0], "rc": o[1], "W0": o[2], "sigma0": o[3], "mesh_lo": o[4:4 + d], "spacing": o[7],
                "cells": [int(c) for c in o[8:8 + d]], "n": int(o[11])}

    def register(self, name, width=1, evolving=True):
        self.sys.check(self.lib.sphb200_var_register(self.h, name.encode(), width, int(evolving), None))
        self.widths[name] = width

    def set(self, name, values, evolving=True):
        """register (if new) + upload; values shaped (n,) or (n, width) in the CURRENT particle order"""
        a = np.ascontiguousarray(values, dtype=self.sys.dtype)
        width = 1 if a.ndim == 1 else a.shape[1]
        assert a.shape[0] == self.n, (name, a.shape, self.n)
        self.register(name, width, evolving)
        if self.n:
            self.sys.check(self.lib.sphb200_var_upload(self.h, name.encode(), a.ctypes.data_as(C.c_void_p)))

    def stage(self, name, host_ptr):
        """asynchronous upload of a (pinned) host buffer into the staging mirror; commit_staged() swaps it in"""
        self.sys.check(self.lib.sphb200_var_stage(self.h, name.encode(), C.c_void_p(host_ptr)))

    def commit_staged(self):
        self.sys.check(self.lib.sphb200_body_commit_staged(self.h))

    def fetch_async(self, name, host_ptr):
        """asynchronous device -> host read (output / restart path); SPHSystem.fetch_wait() completes it"""
        self.sys.check(self.lib.sphb200_var_fetch_async(self.h, name.encode(), C.c_void_p(host_ptr)))

    def set_ids(self, original_id):
        ids = np.ascontiguousarray(original_id, dtype=np.int64)
        self.sys.check(self.lib.sphb200_ids_from_host(self.h, ids.ctypes.data_as(C.POINTER(C.c_int64))))

    def width(self, name):
        w = C.c_int()
        self.sys.check(self.lib.sphb200_var_width(self.h, name.encode(), C.byref(w)))
        return w.value

    def get(self, name, width=None):
        width = self.width(name)  # the library is authoritative (dynamics register their own outputs)
        out = np.zeros((self.n, width) if width > 1 else (self.n,), dtype=self.sys.dtype)
        if self.n:
            self.sys.check(self.lib.sphb200_var_download(self.h, name.encode(), out.ctypes.data_as(C.c_void_p)))
        return out

    def device_ptr(self, name):
        p = C.c_void_p()
        self.sys.check(self.lib.sphb200_var_device_ptr(self.h, name.encode(), C.byref(p)))
        return p.value

    def ids(self):
        o = np.zeros(self.n, dtype=np.int64)
        s = np.zeros(self.n, dtype=np.int64)
        self.sys.check(self.lib.sphb200_ids_to_host(self.h, o.ctypes.data_as(C.POINTER(C.c_int64)),
                                                     s.ctypes.data_as(C.POINTER(C.c_int64))))
        return o, s

    def update_cell_list(self):
        self.sys.check(self.lib.sphb200_cell_list_update(self.h))

    def sort(self):
        """ParticleSorting::exec - the cell list update already keeps the body cell-sorted"""
        self.update_cell_list()

    def cell_keys(self):
        lin = np.zeros(self.n, dtype=np.int64)
        mor = np.zeros(self.n, dtype=np.int64)
        self.sys.check(self.lib.sphb200_cell_keys(self.h, lin.ctypes.data_as(C.POINTER(C.c_int64)),
                                                   mor.ctypes.data_as(C.POINTER(C.c_int64))))
        return lin, mor

    def cell_offsets(self):
        nc = C.c_int64()
        self.sys.check(self.lib.sphb200_cell_offsets(self.h, None, C.byref(nc)))
        out = np.zeros(nc.value + 1, dtype=np.int64)
        self.sys.check(self.lib.sphb200_cell_offsets(self.h, out.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(nc)))
        return out

    def periodic_bounding(self, axis):
        self.sys.check(self.lib.sphb200_periodic_bounding(self.h, axis))

    def periodic_cell_list(self, axis):
        """PeriodicConditionUsingCellLinkedList::update_cell_linked_list_: the ghost images are
        generated on the fly by the neighbour search, nothing to insert."""

    def resize(self, n_new):
        self.sys.check(self.lib.sphb200_body_resize(self.h, int(n_new)))
        self.n = int(n_new)

    def set_owned_pages(self, lo, hi):
        self.sys.check(self.lib.sphb200_body_set_owned_pages(self.h, int(lo), int(hi)))

    def owned_range(self):
        b, e = C.c_int64(), C.c_int64()
        self.sys.check(self.lib.sphb200_body_owned_range(self.h, C.byref(b), C.byref(e)))
        return b.value, e.value

    def slab_select(self, has_left, has_right):
        """(stay, to left, to right) row counts of the re-decomposition (sphb200_slab_select)"""
        out = (C.c_int64 * 3)()
        self.sys.check(self.lib.sphb200_slab_select(self.h, int(has_left), int(has_right), out))
        return int(out[0]), int(out[1]), int(out[2])

    def slab_recut(self, lo, hi):
        self.sys.check(self.lib.sphb200_slab_recut(self.h, int(lo), int(hi)))

    def slab_pack(self, side, names, out_ptr):
        arr = (C.c_char_p * len(names))(*[n.encode() for n in names])
        self.sys.check(self.lib.sphb200_slab_pack(self.h, int(side), arr, len(names), C.c_void_p(out_ptr)))

    def slab_rebuild(self, names, left_ptr, n_left, right_ptr, n_right):
        arr = (C.c_char_p * len(names))(*[n.encode() for n in names])
        self.sys.check(self.lib.sphb200_slab_rebuild(self.h, arr, len(names), C.c_void_p(left_ptr), int(n_left), C.c_void_p(right_ptr), int(n_right)))
        self.n = self.info()["n"]

    def keep_range(self, begin, end):
        self.sys.check(self.lib.sphb200_body_keep_range(self.h, int(begin), int(end)))
        self.n = int(end) - int(begin)

    def page_begin(self, page):
        out = C.c_int64()
        self.sys.check(self.lib.sphb200_body_page_begin(self.h, int(page), C.byref(out)))
        return out.value

    def page_offsets(self):
        """page_begin(p) for every page boundary p = 0..pages (sphb200_body_page_offsets)"""
        n = C.c_int64()
        self.sys.check(self.lib.sphb200_body_page_offsets(self.h, None, C.byref(n)))
        out = np.zeros(n.value + 1, dtype=np.int64)
        self.sys.check(self.lib.sphb200_body_page_offsets(self.h, out.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(n)))
        return out

    def record_ptr(self, which):
        p = C.c_void_p()
        self.sys.check(self.lib.sphb200_body_record_ptr(self.h, int(which), C.byref(p)))
        return p.value

    def ids_device_ptr(self):
        a, b = C.c_void_p(), C.c_void_p()
        self.sys.check(self.lib.sphb200_ids_device_ptr(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value


class Relation:
    def __init__(self, system, body, targets, builder=None, part_ids=None):
        self.sys, self.lib, self.body, self.targets = system, system.lib, body, targets
        h = C.c_void_p()
        ids = np.ascontiguousarray(part_ids if part_ids is not None else [], dtype=np.int64)
        idp = ids.ctypes.data_as(C.POINTER(C.c_int64)) if ids.size else None
        if builder == "self":
            system.check(self.lib.sphb200_relation_self_contact_create(body.h, idp, ids.size, C.byref(h)))
        elif builder == "surface":
            arr = (C.c_void_p * len(targets))(*[t.h for t in targets])
            system.check(self.lib.sphb200_relation_surface_contact_create(body.h, arr, len(targets), idp, ids.size, C.byref(h)))
        elif targets is None:
            system.check(self.lib.sphb200_relation_inner_create(body.h, C.byref(h)))
        else:
            arr = (C.c_void_p * len(targets))(*[t.h for t in targets])
            system.check(self.lib.sphb200_relation_contact_create(body.h, arr, len(targets), C.byref(h)))
        self.h = h

    def update(self):
        """SPHRelation::updateConfiguration"""
        self.sys.check(self.lib.sphb200_relation_update(self.h))

    def neighbor_counts(self, k=0):
        """counts[n] only (no index list): cheap at full size"""
        total = C.c_int64()
        counts = np.zeros(self.body.n, dtype=np.int32)
        self.sys.check(self.lib.sphb200_relation_neighbors(self.h, k, counts.ctypes.data_as(C.POINTER(C.c_int32)), None, C.byref(total)))
        return counts, int(total.value)

    def stats(self, k=0):
        out = (C.c_double * 8)()
        self.sys.check(self.lib.sphb200_relation_stats(self.h, k, out))
        o = list(out)
        return {"pairs": int(o[0]), "masked_search": bool(o[1]), "index_list": bool(o[2]), "union_entries": int(o[3]),
                "union_groups": int(o[4]), "group_size": int(o[5]), "n": int(o[6])}

    def neighbors(self, k=0):
        """(counts[n], flat j in list order) of target k - parity/debug"""
        n = self.body.n
        total = C.c_int64()
        self.sys.check(self.lib.sphb200_relation_neighbors(self.h, k, None, None, C.byref(total)))
        counts = np.zeros(n, dtype=np.int32)
        j = np.zeros(total.value, dtype=np.int64)
        self.sys.check(self.lib.sphb200_relation_neighbors(
            self.h, k, counts.ctypes.data_as(C.POINTER(C.c_int32)), j.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(total)))
        return counts, j
