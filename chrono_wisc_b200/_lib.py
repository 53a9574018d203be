"""ctypes binding of libchrono_b200.so (the C ABI in include/chrono_b200.h).

The library is built in-tree by ``chrono_wisc_b200/csrc/Makefile`` (``__graft_entry__.build()``).  There is no
CPU fallback anywhere in this package: if the shared library is missing or no CUDA device is usable, the calls
raise.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libchrono_b200.so")
CSRC = os.path.join(_HERE, "csrc")


class cb200_settings(C.Structure):
    """Mirror of ``cb200_settings`` (include/chrono_b200.h) = collision_settings + solver_settings of the
    reference (src/chrono_multicore/ChSettings.h:39-246)."""

    _fields_ = [
        ("collision_envelope", C.c_double),
        ("broadphase_grid", C.c_int32),
        ("bins_per_axis", C.c_int32 * 3),
        ("bin_size", C.c_double * 3),
        ("grid_density", C.c_double),
        ("narrowphase_algorithm", C.c_int32),
        ("system_type", C.c_int32),
        ("step_size", C.c_double),
        ("gravity", C.c_double * 3),
        ("contact_force_model", C.c_int32),
        ("adhesion_force_model", C.c_int32),
        ("tangential_displ_mode", C.c_int32),
        ("use_material_properties", C.c_int32),
        ("characteristic_vel", C.c_double),
        ("min_slip_vel", C.c_double),
        ("min_roll_vel", C.c_double),
        ("min_spin_vel", C.c_double),
        ("solver_type", C.c_int32),
        ("solver_mode", C.c_int32),
        ("max_iteration_normal", C.c_int32),
        ("max_iteration_sliding", C.c_int32),
        ("max_iteration_spinning", C.c_int32),
        ("cache_step_length", C.c_int32),
        ("alpha", C.c_double),
        ("contact_recovery_speed", C.c_double),
        ("tolerance", C.c_double),
        ("deterministic", C.c_int32),
        ("test_objective", C.c_int32),
        ("tolerance_objective", C.c_double),
        ("reserved", C.c_int32 * 4),
    ]


class cb200_counts(C.Structure):
    _fields_ = [
        ("num_bodies", C.c_uint32),
        ("num_shapes", C.c_uint32),
        ("bins_per_axis", C.c_uint32 * 3),
        ("num_bins", C.c_uint32),
        ("num_bin_aabb_intersections", C.c_uint32),
        ("num_active_bins", C.c_uint32),
        ("num_possible_collisions", C.c_uint32),
        ("num_potential_contacts", C.c_uint32),
        ("num_rigid_contacts", C.c_uint32),
        ("ct_body_count", C.c_uint32),
        ("solver_iterations", C.c_uint32),
        ("solver_slots", C.c_uint32),
        ("reserved", C.c_uint32 * 2),
        ("min_bounding_point", C.c_double * 3),
        ("max_bounding_point", C.c_double * 3),
        ("bin_size", C.c_double * 3),
        ("solver_residual", C.c_double),
        ("solver_objective", C.c_double),
    ]


# every symbol include/chrono_b200.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_SIGNATURES = {
    "cb200_last_error": (C.c_char_p, []),
    "cb200_device_count": (C.c_int, []),
    "cb200_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "cb200_host_alloc": (C.c_int, [C.c_uint64, C.POINTER(_P)]),
    "cb200_host_free": (C.c_int, [_P]),
    "cb200_destroy": (None, [_P]),
    "cb200_default_settings": (None, [C.POINTER(cb200_settings)]),
    "cb200_set_settings": (C.c_int, [_P, C.POINTER(cb200_settings)]),
    "cb200_append_bodies": (C.c_int, [_P, C.c_int, _P, _P, _P, _P, _P, C.POINTER(C.c_int)]),
    "cb200_set_inertia_full": (C.c_int, [_P, C.c_int, C.c_int, _P, _P]),
    "cb200_append_shapes": (C.c_int, [_P, C.c_int, _P, _P, _P, _P, _P, _P, _P]),
    "cb200_append_shape_vertices": (C.c_int, [_P, C.c_int, _P, _P]),
    "cb200_update_shape_vertices": (C.c_int, [_P, C.c_int, C.c_int, _P]),
    "cb200_set_materials_smc": (C.c_int, [_P, C.c_int, _P]),
    "cb200_set_materials_nsc": (C.c_int, [_P, C.c_int, _P]),
    "cb200_upload_state": (C.c_int, [_P, C.c_int, C.c_int, _P, _P, _P]),
    "cb200_upload_applied_forces": (C.c_int, [_P, C.c_int, C.c_int, _P]),
    "cb200_set_body_flags": (C.c_int, [_P, C.c_int, C.c_int, _P, _P]),
    "cb200_download_body_flags": (C.c_int, [_P, C.c_int, C.c_int, _P, _P]),
    "cb200_set_active_aabb": (C.c_int, [_P, C.c_int, _P, _P]),
    "cb200_set_speed_limits": (C.c_int, [_P, C.c_int, C.c_int, _P, _P]),
    "cb200_update_forces": (C.c_int, [_P]),
    "cb200_run_collision": (C.c_int, [_P, C.c_int]),
    "cb200_run_solver_smc": (C.c_int, [_P]),
    "cb200_run_solver_nsc": (C.c_int, [_P]),
    "cb200_integrate": (C.c_int, [_P]),
    "cb200_do_steps": (C.c_int, [_P, C.c_int]),
    "cb200_synchronize": (C.c_int, [_P]),
    "cb200_get_counts": (C.c_int, [_P, C.POINTER(cb200_counts)]),
    "cb200_download_state": (C.c_int, [_P, C.c_int, C.c_int, _P, _P, _P]),
    "cb200_step_host": (C.c_int, [_P, _P, _P, _P]),
    "cb200_download_aabb": (C.c_int, [_P, _P, _P]),
    "cb200_download_bins": (C.c_int, [_P, _P, _P, _P, _P]),
    "cb200_download_pairs": (C.c_int, [_P, _P]),
    "cb200_download_contacts": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _P]),
    "cb200_download_contact_forces": (C.c_int, [_P, _P, _P]),
    "cb200_download_body_contact_forces": (C.c_int, [_P, _P, _P]),
    "cb200_download_hf": (C.c_int, [_P, _P]),
    "cb200_download_gamma": (C.c_int, [_P, _P, C.c_int]),
    "cb200_download_nsc_rhs": (C.c_int, [_P, _P]),
    "cb200_download_nsc_history": (C.c_int, [_P, _P, _P, C.c_int, C.POINTER(C.c_int)]),
    "cb200_download_nsc_jacobian": (C.c_int, [_P, _P, _P]),
    "cb200_nsc_schur_product": (C.c_int, [_P, C.c_int, _P, _P]),
    "cb200_nsc_project": (C.c_int, [_P, C.c_int, _P]),
    "cb200_get_timer": (C.c_int, [_P, C.c_char_p, C.POINTER(C.c_double)]),
    "cb200_kernel_launches": (C.c_int64, [_P]),
    "cb200_set_bilaterals": (C.c_int, [_P, C.c_int, _P, _P, _P, _P, C.c_int, C.c_double]),
    "cb200_download_bilateral_gamma": (C.c_int, [_P, _P, _P, _P]),
    "cb200_minres_noalloc": (C.c_int, [_P, C.c_int, _P, _P, _P, C.c_int, C.c_double, _P]),
    "cb200_minres_noalloc_max_rows": (C.c_int, []),
    "cb200_filter_contacts": (C.c_int, [_P, C.c_int, _P]),
    "cb200_download_contact_materials": (C.c_int, [_P, C.c_int, _P, _P]),
    "cb200_upload_contact_materials": (C.c_int, [_P, C.c_int, _P]),
    "cb200_report_contacts": (C.c_int, [_P, C.c_int, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cb200_selftest_scan": (C.c_int, [_P, C.c_int64, C.c_uint32, C.c_int, _P]),
    "cb200_device_array": (C.c_int, [_P, C.c_char_p, C.POINTER(_P), C.POINTER(C.c_int64)]),
    "cb200_set_stream": (C.c_int, [_P, _P, C.c_int]),
    "cb200_enqueue_steps": (C.c_int, [_P, C.c_int]),
    "cb200_check": (C.c_int, [_P]),
    "cb200_reset_timers": (C.c_int, [_P]),
    "cb200_upload_body_ids": (C.c_int, [_P, C.c_int, C.c_int, _P]),
    "cb200_download_body_ids": (C.c_int, [_P, C.c_int, C.c_int, _P, _P]),
    "cb200_slab_configure": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int,
                                       C.c_int, C.c_int, C.c_int]),
    "cb200_slab_buffer_doubles": (C.c_int64, [_P]),
    "cb200_slab_pack_dev": (C.c_int, [_P, _P, _P]),
    "cb200_slab_unpack_dev": (C.c_int, [_P, _P, _P]),
    "cb200_comm_unique_id": (C.c_int, [C.c_char_p, _P]),
    "cb200_comm_init_nccl": (C.c_int, [_P, C.c_char_p, C.c_int, C.c_int, _P]),
    "cb200_comm_init_local": (C.c_int, [_P, C.c_int]),
    "cb200_comm_exchange": (C.c_int, [_P, C.c_int]),
    "cb200_comm_steps": (C.c_int, [_P, C.c_int, C.c_int]),
}

_lib = None


def build(verbose=False):
    """Compile libchrono_b200.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    cmd = ["make", "-C", CSRC, "-j8"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:])
        print(res.stderr[-4000:])
    if res.returncode != 0:
        raise RuntimeError("building libchrono_b200.so failed")
    return LIB_PATH


def load():
    """Load the shared library and attach the prototypes; raises if it is missing (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with __graft_entry__.build() (nvcc, sm_100a). "
            "There is no CPU fallback for the chrono_wisc_b200 hot path."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the ABI lost a symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class Cb200Error(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        msg = load().cb200_last_error()
        raise Cb200Error(f"libchrono_b200 error {rc}: {msg.decode() if msg else ''}")
