"""ctypes binding of the C ABI declared in include/sphb200.h.

The library is the product: if libsphb200.so is missing or no CUDA device is present this module
raises - there is no CPU fallback (the CPU oracle lives under oracle/ and is test infrastructure only).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsphb200.so")

OK, ERR_INVALID, ERR_CUDA, ERR_NOT_FOUND, ERR_NO_DEVICE = range(5)


class SphB200Error(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"sphb200 status {status}: {message}")
        self.status = status


_lib = None

_vp, _i, _i64, _d, _cp = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_char_p
_dp = C.POINTER(C.c_double)
_i64p = C.POINTER(C.c_int64)
_vpp = C.POINTER(C.c_void_p)

# name -> (restype, argtypes); every symbol include/sphb200.h declares
SIGNATURES = {
    "sphb200_ctx_create": (_i, [_i, _i, _i, _dp, _dp, _vpp]),
    "sphb200_ctx_destroy": (None, [_vp]),
    "sphb200_ctx_synchronize": (_i, [_vp]),
    "sphb200_last_error": (_cp, [_vp]),
    "sphb200_ctx_stream": (_vp, [_vp]),
    "sphb200_ctx_launch_count": (_i64, [_vp]),
    "sphb200_ctx_set_pair_kernel": (_i, [_vp, _i]),
    "sphb200_body_create": (_i, [_vp, _cp, _i64, _d, _d, _vpp]),
    "sphb200_body_set_cell_spacing": (_i, [_vp, _d]),
    "sphb200_body_set_fluid": (_i, [_vp, _d, _d, _d]),
    "sphb200_body_set_elastic": (_i, [_vp, _d, _d, _d]),
    "sphb200_body_set_periodic": (_i, [_vp, _i, _d, _d]),
    "sphb200_body_info": (_i, [_vp, _dp]),
    "sphb200_var_register": (_i, [_vp, _cp, _i, _i, _vp]),
    "sphb200_var_to_device": (_i, [_vp, _cp]),
    "sphb200_var_to_host": (_i, [_vp, _cp]),
    "sphb200_var_upload": (_i, [_vp, _cp, _vp]),
    "sphb200_var_download": (_i, [_vp, _cp, _vp]),
    "sphb200_var_stage": (_i, [_vp, _cp, _vp]),
    "sphb200_body_commit_staged": (_i, [_vp]),
    "sphb200_ctx_stage_wait": (_i, [_vp]),
    "sphb200_var_fetch_async": (_i, [_vp, _cp, _vp]),
    "sphb200_ctx_fetch_wait": (_i, [_vp]),
    "sphb200_ids_from_host": (_i, [_vp, _i64p]),
    "sphb200_var_device_ptr": (_i, [_vp, _cp, _vpp]),
    "sphb200_var_width": (_i, [_vp, _cp, C.POINTER(C.c_int)]),
    "sphb200_ids_to_host": (_i, [_vp, _i64p, _i64p]),
    "sphb200_cell_list_update": (_i, [_vp]),
    "sphb200_cell_keys": (_i, [_vp, _i64p, _i64p]),
    "sphb200_cell_offsets": (_i, [_vp, _i64p, _i64p]),
    "sphb200_periodic_bounding": (_i, [_vp, _i]),
    "sphb200_relation_inner_create": (_i, [_vp, _vpp]),
    "sphb200_relation_contact_create": (_i, [_vp, _vpp, _i, _vpp]),
    "sphb200_relation_self_contact_create": (_i, [_vp, _i64p, _i64, _vpp]),
    "sphb200_relation_surface_contact_create": (_i, [_vp, _vpp, _i, _i64p, _i64, _vpp]),
    "sphb200_relation_update": (_i, [_vp]),
    "sphb200_relation_neighbors": (_i, [_vp, _i, C.POINTER(C.c_int32), _i64p, _i64p]),
    "sphb200_relation_stats": (_i, [_vp, _i, _dp]),
    "sphb200_density_summation": (_i, [_vp, _vp, _i]),
    "sphb200_integration_1st_half": (_i, [_vp, _vp, _i, _d]),
    "sphb200_integration_2nd_half": (_i, [_vp, _vp, _i, _d]),
    "sphb200_acoustic_dt": (_i, [_vp, _d, _dp]),
    "sphb200_advection_dt": (_i, [_vp, _d, _d, _i, _dp]),
    "sphb200_acoustic_relaxation": (_i, [_vp, _vp, _i, _i, _d, _d, _i, _d, _i, _i, _dp]),
    "sphb200_relaxation_state": (_i, [_vp, _dp]),
    "sphb200_relaxation_begin": (_i, [_vp, _d, _d]),
    "sphb200_substep_ctl": (_i, [_vp, _d, _i, _i]),
    "sphb200_reduce_slot_ptr": (_i, [_vp, _i, _vpp]),
    "sphb200_slab_select": (_i, [_vp, _i, _i, _i64p]),
    "sphb200_slab_recut": (_i, [_vp, _i, _i]),
    "sphb200_slab_pack": (_i, [_vp, _i, C.POINTER(C.c_char_p), _i, _vp]),
    "sphb200_slab_rebuild": (_i, [_vp, C.POINTER(C.c_char_p), _i, _vp, _i64, _vp, _i64]),
    "sphb200_rows_gather": (_i, [_vp, _vp, _i, _vp, _i64, _vp]),
    "sphb200_rows_scatter": (_i, [_vp, _vp, _i, _vp, _i64, _vp]),
    "sphb200_peer_alloc": (_i, [_vp, _i64, _vpp]),
    "sphb200_peer_free": (_i, [_vp, _vp]),
    "sphb200_peer_export": (_i, [_vp, _vp, C.c_char_p]),
    "sphb200_peer_open": (_i, [_vp, C.c_char_p, _vpp]),
    "sphb200_peer_close": (_i, [_vp, _vp]),
    "sphb200_rows_push": (_i, [_vp, _vp, _i, _vp, _i64, _vp, _vp, C.c_uint64]),
    "sphb200_rows_wait_scatter": (_i, [_vp, _vp, _i, _vp, _i64, _vp, _vp, C.c_uint64]),
    "sphb200_peer_timeouts": (_i, [_vp, _i64p]),
    "sphb200_gravity_force": (_i, [_vp, _dp]),
    "sphb200_viscous_force": (_i, [_vp, _vp]),
    "sphb200_transport_velocity_correction": (_i, [_vp, _d]),
    "sphb200_total_kinetic_energy": (_i, [_vp, _dp]),
    "sphb200_total_mechanical_energy": (_i, [_vp, _dp, _dp]),
    "sphb200_maximum_speed": (_i, [_vp, _dp]),
    "sphb200_position_bound": (_i, [_vp, _i, _dp]),
    "sphb200_observe": (_i, [_vp, _cp, _i, _dp]),
    "sphb200_ck_density_regularization": (_i, [_vp, _vp, _i]),
    "sphb200_ck_linear_correction_matrix": (_i, [_vp, _vp, _d]),
    "sphb200_ck_free_surface_indication": (_i, [_vp, _vp]),
    "sphb200_ck_compression_summation": (_i, [_vp, _vp]),
    "sphb200_ck_density_regularize": (_i, [_vp, _i]),
    "sphb200_ck_pressure_force_from_fluid": (_i, [_vp, _i]),
    "sphb200_ck_viscous_force_from_fluid": (_i, [_vp, _i]),
    "sphb200_ck_advection_step_setup": (_i, [_vp]),
    "sphb200_ck_update_particle_position": (_i, [_vp]),
    "sphb200_ck_acoustic_dt": (_i, [_vp, _d, _dp]),
    "sphb200_ck_advection_dt": (_i, [_vp, _d, _d, _dp]),
    "sphb200_ck_acoustic_step_1st_half": (_i, [_vp, _vp, _d]),
    "sphb200_ck_acoustic_step_2nd_half": (_i, [_vp, _vp, _d]),
    "sphb200_correction_matrix": (_i, [_vp]),
    "sphb200_solid_1st_half": (_i, [_vp, _i, _d]),
    "sphb200_solid_2nd_half": (_i, [_vp, _d]),
    "sphb200_solid_constrain": (_i, [_vp, _i64p, _i64]),
    "sphb200_solid_dt": (_i, [_vp, _d, _dp]),
    "sphb200_exclusive_scan_u32": (_i, [_vp, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), _i64, C.POINTER(C.c_uint64)]),
    "sphb200_exclusive_scan_u32_static": (_i, [_vp, _vp, _vp, _i64]),
    "sphb200_ctx_set_stream": (_i, [_vp, _vp]),
    "sphb200_body_resize": (_i, [_vp, _i64]),
    "sphb200_body_set_owned_pages": (_i, [_vp, _i, _i]),
    "sphb200_body_owned_range": (_i, [_vp, _i64p, _i64p]),
    "sphb200_ids_device_ptr": (_i, [_vp, _vpp, _vpp]),
    "sphb200_body_keep_range": (_i, [_vp, _i64, _i64]),
    "sphb200_body_page_begin": (_i, [_vp, _i, _i64p]),
    "sphb200_body_page_offsets": (_i, [_vp, _i64p, _i64p]),
    "sphb200_body_record_ptr": (_i, [_vp, _i, _vpp]),
    "sphb200_integration_1st_half_init": (_i, [_vp, _d]),
    "sphb200_integration_1st_half_interact": (_i, [_vp, _vp, _i, _d]),
}


def load():
    """Load libsphb200.so (built by __graft_entry__.build()) and bind every declared symbol."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
