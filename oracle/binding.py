"""ORACLE - TEST INFRASTRUCTURE ONLY.  ctypes binding of oracle/liboracle.so (the CPU restatement of the
reference's classic path) exposing the same Python surface as sphinxsys-learning_b200/host.py, so one
case script can drive either.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product path never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        build()
    lib = C.CDLL(LIB_PATH)
    vp, i, l, d, cp = C.c_void_p, C.c_int, C.c_long, C.c_double, C.c_char_p
    dp, lp, ip = C.POINTER(C.c_double), C.POINTER(C.c_long), C.POINTER(C.c_int)
    sig = {
        "orc_create": (vp, [i, i, dp, dp]), "orc_destroy": (None, [vp]),
        "orc_body_add": (i, [vp, cp, l, d, d]), "orc_body_material": (None, [vp, i, d, d, d]),
        "orc_body_elastic": (None, [vp, i, d, d, d]), "orc_body_periodic": (None, [vp, i, i, d, d]),
        "orc_var": (vp, [vp, i, cp, i, i]), "orc_var_width": (i, [vp, i, cp]), "orc_body_info": (None, [vp, i, dp]),
        "orc_ids": (None, [vp, i, lp, lp]), "orc_cell_keys": (None, [vp, i, lp, lp]),
        "orc_update_cell_list": (None, [vp, i]), "orc_periodic_bounding": (None, [vp, i, i]),
        "orc_periodic_cell_list": (None, [vp, i, i]), "orc_sort": (None, [vp, i]),
        "orc_relation_inner": (i, [vp, i]), "orc_relation_contact": (i, [vp, i, ip, i]),
        "orc_relation_self_contact": (i, [vp, i, lp, l]), "orc_relation_surface_contact": (i, [vp, i, ip, i, lp, l]),
        "orc_relation_update": (None, [vp, i]), "orc_neighbors": (l, [vp, i, i, ip, lp]),
        "orc_search_depth": (i, [vp, i, i]),
        "orc_density_summation": (None, [vp, i, i, i]),
        "orc_integration_1st_half": (None, [vp, i, i, i, d]), "orc_integration_2nd_half": (None, [vp, i, i, i, d]),
        "orc_acoustic_dt": (d, [vp, i, d]), "orc_advection_dt": (d, [vp, i, d, d, i]),
        "orc_gravity_force": (None, [vp, i, dp]), "orc_viscous_force": (None, [vp, i, i]),
        "orc_transport_velocity_correction": (None, [vp, i, d]),
        "orc_total_kinetic_energy": (d, [vp, i]), "orc_total_mechanical_energy": (d, [vp, i, dp]),
        "orc_ck_density_regularization": (None, [vp, i, i, i]), "orc_ck_linear_correction_matrix": (None, [vp, i, i, d]),
        "orc_ck_free_surface_indication": (None, [vp, i, i]), "orc_ck_advection_step_setup": (None, [vp, i]),
        "orc_ck_update_particle_position": (None, [vp, i]), "orc_ck_acoustic_dt": (d, [vp, i, d]), "orc_ck_advection_dt": (d, [vp, i, d, d]),
        "orc_ck_pressure_force_from_fluid": (None, [vp, i, i]), "orc_ck_viscous_force_from_fluid": (None, [vp, i, i]),
        "orc_ck_acoustic_1st_half": (None, [vp, i, i, d]), "orc_ck_acoustic_2nd_half": (None, [vp, i, i, d]), "orc_maximum_speed": (d, [vp, i]), "orc_position_bound": (None, [vp, i, i, dp]), "orc_observe": (None, [vp, i, cp, i, dp]),
        "orc_correction_matrix": (None, [vp, i]), "orc_solid_1st_half": (None, [vp, i, i, d]),
        "orc_solid_2nd_half": (None, [vp, i, d]), "orc_solid_constrain": (None, [vp, i, lp, l]),
        "orc_solid_dt": (d, [vp, i, d]), "orc_num_threads": (i, []), "orc_set_num_threads": (i, [i]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


class SPHSystem:
    backend = "oracle"

    def __init__(self, ndim, lo, hi, dp_ref, real_bytes=8, expand=True, **_):
        self.lib = load()
        self.ndim, self.real_bytes = int(ndim), int(real_bytes)
        self.dtype = np.float64 if real_bytes == 8 else np.float32
        ex = 4.0 * dp_ref if expand else 0.0
        self.lo = np.asarray(lo, dtype=np.float64) - ex
        self.hi = np.asarray(hi, dtype=np.float64) + ex
        lo3 = (C.c_double * 3)(*list(self.lo) + [0.0] * (3 - ndim))
        hi3 = (C.c_double * 3)(*list(self.hi) + [0.0] * (3 - ndim))
        self.h = self.lib.orc_create(ndim, real_bytes, lo3, hi3)
        self.bodies, self.relations = [], []

    def close(self):
        if self.h:
            self.lib.orc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def synchronize(self):
        pass

    launches = 0

    def add_body(self, name, n, dp, h_ratio=1.3):
        b = Body(self, name, n, dp, h_ratio)
        self.bodies.append(b)
        return b

    def inner(self, body):
        return Relation(self, body, None)

    def contact(self, body, targets):
        return Relation(self, body, list(targets))

    def self_contact(self, body, part_ids=None):
        return Relation(self, body, None, builder="self", part_ids=part_ids)

    def surface_contact(self, body, targets, part_ids=None):
        return Relation(self, body, list(targets), builder="surface", part_ids=part_ids)

    def _g3(self, g):
        return (C.c_double * 3)(*list(g) + [0.0] * (3 - len(g)))

    def density_summation(self, inner, contact=None, free_surface=False):
        self.lib.orc_density_summation(self.h, inner.id, contact.id if contact else -1, int(free_surface))

    def integration_1st_half(self, inner, wall=None, riemann=True, dt=0.0):
        self.lib.orc_integration_1st_half(self.h, inner.id, wall.id if wall else -1, int(riemann), dt)

    def integration_2nd_half(self, inner, wall=None, riemann=True, dt=0.0):
        self.lib.orc_integration_2nd_half(self.h, inner.id, wall.id if wall else -1, int(riemann), dt)

    def acoustic_dt(self, body, cfl=0.6):
        return self.lib.orc_acoustic_dt(self.h, body.id, cfl)

    def advection_dt(self, body, U_ref, cfl=0.25, viscous=False):
        return self.lib.orc_advection_dt(self.h, body.id, U_ref, cfl, int(viscous))

    def gravity_force(self, body, g):
        self.lib.orc_gravity_force(self.h, body.id, self._g3(g))

    def viscous_force(self, inner, wall=None):
        self.lib.orc_viscous_force(self.h, inner.id, wall.id if wall else -1)

    def transport_velocity_correction(self, inner, coefficient=0.2):
        self.lib.orc_transport_velocity_correction(self.h, inner.id, coefficient)

    def ck_density_regularization(self, inner, contact=None, free_surface=True):
        self.lib.orc_ck_density_regularization(self.h, inner.id, contact.id if contact else -1, int(free_surface))

    def ck_linear_correction_matrix(self, inner, contact=None, alpha=0.0):
        self.lib.orc_ck_linear_correction_matrix(self.h, inner.id, contact.id if contact else -1, alpha)

    def ck_advection_step_setup(self, body):
        self.lib.orc_ck_advection_step_setup(self.h, body.id)

    def ck_update_particle_position(self, body):
        self.lib.orc_ck_update_particle_position(self.h, body.id)

    def ck_acoustic_dt(self, body, cfl=0.6):
        return self.lib.orc_ck_acoustic_dt(self.h, body.id, cfl)

    def ck_advection_dt(self, body, U_ref, cfl=0.25):
        return self.lib.orc_ck_advection_dt(self.h, body.id, U_ref, cfl)

    def ck_acoustic_1st_half(self, inner, wall=None, dt=0.0):
        self.lib.orc_ck_acoustic_1st_half(self.h, inner.id, wall.id if wall else -1, dt)

    def ck_acoustic_2nd_half(self, inner, wall=None, dt=0.0):
        self.lib.orc_ck_acoustic_2nd_half(self.h, inner.id, wall.id if wall else -1, dt)

    def ck_pressure_force_from_fluid(self, contact, corrected=True):
        self.lib.orc_ck_pressure_force_from_fluid(self.h, contact.id, int(corrected))

    def ck_viscous_force_from_fluid(self, contact, corrected=True):
        self.lib.orc_ck_viscous_force_from_fluid(self.h, contact.id, int(corrected))

    def ck_free_surface_indication(self, inner, contact=None):
        self.lib.orc_ck_free_surface_indication(self.h, inner.id, contact.id if contact else -1)

    def total_kinetic_energy(self, body):
        return self.lib.orc_total_kinetic_energy(self.h, body.id)

    def total_mechanical_energy(self, body, g):
        return self.lib.orc_total_mechanical_energy(self.h, body.id, self._g3(g))

    def maximum_speed(self, body):
        return self.lib.orc_maximum_speed(self.h, body.id)

    def position_bound(self, body, upper=True):
        out = (C.c_double * 3)()
        self.lib.orc_position_bound(self.h, body.id, int(upper), out)
        return np.asarray(list(out)[:self.ndim])

    def observe(self, contact, name, width=1):
        out = np.zeros(contact.body.n * width, dtype=np.float64)
        self.lib.orc_observe(self.h, contact.id, name.encode(), width, out.ctypes.data_as(C.POINTER(C.c_double)))
        return out.reshape(contact.body.n, width) if width > 1 else out

    def correction_matrix(self, inner):
        self.lib.orc_correction_matrix(self.h, inner.id)

    def solid_1st_half(self, inner, decomposed, dt):
        self.lib.orc_solid_1st_half(self.h, inner.id, int(decomposed), dt)

    def solid_2nd_half(self, inner, dt):
        self.lib.orc_solid_2nd_half(self.h, inner.id, dt)

    def solid_constrain(self, body, original_ids):
        # the oracle's fixBodyPart takes CURRENT indices: map through sorted_id
        _, sorted_id = body.ids()
        ids = np.ascontiguousarray(sorted_id[np.asarray(original_ids, dtype=np.int64)], dtype=np.int64)
        self.lib.orc_solid_constrain(self.h, body.id, ids.ctypes.data_as(C.POINTER(C.c_long)), ids.size)

    def solid_dt(self, body, cfl=0.6):
        return self.lib.orc_solid_dt(self.h, body.id, cfl)


class Body:
    def __init__(self, system, name, n, dp, h_ratio):
        self.sys, self.lib, self.name = system, system.lib, name
        self.n, self.dp, self.h_ratio = int(n), float(dp), float(h_ratio)
        self.id = self.lib.orc_body_add(system.h, name.encode(), self.n, dp, h_ratio)
        self.widths = {"Position": system.ndim, "VolumetricMeasure": 1}

    def set_fluid(self, rho0, c0, mu=0.0):
        self.lib.orc_body_material(self.sys.h, self.id, rho0, c0, mu)

    def set_elastic(self, rho0, youngs, poisson):
        self.lib.orc_body_elastic(self.sys.h, self.id, rho0, youngs, poisson)

    def set_periodic(self, axis, lo, hi):
        self.lib.orc_body_periodic(self.sys.h, self.id, axis, lo, hi)

    def info(self):
        out = (C.c_double * 12)()
        self.lib.orc_body_info(self.sys.h, self.id, out)
        o = list(out)
        d = self.sys.ndim
        return {"h": o[0], "rc": o[1], "W0": o[2], "sigma0": o[3], "mesh_lo": o[4:4 + d], "spacing": o[7],
                "cells": [int(c) for c in o[8:8 + d]], "n": self.n}

    def _view(self, name, width, evolving=False):
        p = self.lib.orc_var(self.sys.h, self.id, name.encode(), width, int(evolving))
        ctype = C.c_double if self.sys.real_bytes == 8 else C.c_float
        arr = np.ctypeslib.as_array(C.cast(p, C.POINTER(ctype)), shape=(self.n * width,))
        return arr.reshape(self.n, width) if width > 1 else arr

    def register(self, name, width=1, evolving=True):
        self._view(name, width, evolving)
        self.widths[name] = width

    def set(self, name, values, evolving=True):
        a = np.ascontiguousarray(values, dtype=self.sys.dtype)
        width = 1 if a.ndim == 1 else a.shape[1]
        self.widths[name] = width
        if self.n:
            self._view(name, width, evolving)[...] = a
        else:
            self._view(name, width, evolving)

    def get(self, name, width=None):
        width = self.lib.orc_var_width(self.sys.h, self.id, name.encode()) or width or self.widths.get(name, 1)
        return self._view(name, width).copy() if self.n else np.zeros((0, width) if width > 1 else (0,), self.sys.dtype)

    def ids(self):
        o = np.zeros(self.n, dtype=np.int64)
        s = np.zeros(self.n, dtype=np.int64)
        self.lib.orc_ids(self.sys.h, self.id, o.ctypes.data_as(C.POINTER(C.c_long)), s.ctypes.data_as(C.POINTER(C.c_long)))
        return o, s

    def update_cell_list(self):
        self.lib.orc_update_cell_list(self.sys.h, self.id)

    def sort(self):
        self.lib.orc_sort(self.sys.h, self.id)

    def cell_keys(self):
        lin = np.zeros(self.n, dtype=np.int64)
        mor = np.zeros(self.n, dtype=np.int64)
        self.lib.orc_cell_keys(self.sys.h, self.id, lin.ctypes.data_as(C.POINTER(C.c_long)), mor.ctypes.data_as(C.POINTER(C.c_long)))
        return lin, mor

    def periodic_bounding(self, axis):
        self.lib.orc_periodic_bounding(self.sys.h, self.id, axis)

    def periodic_cell_list(self, axis):
        self.lib.orc_periodic_cell_list(self.sys.h, self.id, axis)


class Relation:
    def __init__(self, system, body, targets, builder=None, part_ids=None):
        self.sys, self.lib, self.body, self.targets = system, system.lib, body, targets
        ids = np.ascontiguousarray(part_ids if part_ids is not None else [], dtype=np.int64)
        idp = ids.ctypes.data_as(C.POINTER(C.c_long)) if ids.size else None
        if builder == "self":
            self.id = self.lib.orc_relation_self_contact(system.h, body.id, idp, ids.size)
        elif builder == "surface":
            arr = (C.c_int * len(targets))(*[t.id for t in targets])
            self.id = self.lib.orc_relation_surface_contact(system.h, body.id, arr, len(targets), idp, ids.size)
        elif targets is None:
            self.id = self.lib.orc_relation_inner(system.h, body.id)
        else:
            arr = (C.c_int * len(targets))(*[t.id for t in targets])
            self.id = self.lib.orc_relation_contact(system.h, body.id, arr, len(targets))

    def update(self):
        self.lib.orc_relation_update(self.sys.h, self.id)

    def neighbors(self, k=0):
        n = self.body.n
        counts = np.zeros(n, dtype=np.int32)
        total = self.lib.orc_neighbors(self.sys.h, self.id, k, counts.ctypes.data_as(C.POINTER(C.c_int)), None)
        j = np.zeros(total, dtype=np.int64)
        self.lib.orc_neighbors(self.sys.h, self.id, k, counts.ctypes.data_as(C.POINTER(C.c_int)), j.ctypes.data_as(C.POINTER(C.c_long)))
        return counts, j
