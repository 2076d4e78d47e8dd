"""ctypes mirror of include/viekf_batch.h (the C ABI of the batch engine).

Struct layouts, enum values and index maps follow the reference's public header
(reference include/vi_ekf.h:87-138) through include/viekf_batch.h.
"""
import ctypes as C
import os
import numpy as np

# index maps (reference include/vi_ekf.h:87-111)
xPOS, xVEL, xATT, xB_A, xB_G, xMU, xZ = 0, 3, 6, 10, 13, 16, 17
dxPOS, dxVEL, dxATT, dxB_A, dxB_G, dxMU, dxZ = 0, 3, 6, 9, 12, 15, 16
# measurement_type_t (reference include/vi_ekf.h:113-125)
ACC, ALT, ATT, POS, VEL, QZETA, FEAT, PIXEL_VEL, DEPTH, INV_DEPTH = range(10)
# meas_result_t (reference include/vi_ekf.h:132-138)
MEAS_SUCCESS, MEAS_GATED, MEAS_NAN, MEAS_INVALID, MEAS_NEW_FEATURE = range(5)
MEAS_NONE = 255
EV_PROP, EV_UPD, EV_INIT, EV_CLEAR, EV_KEEP, EV_STAT, EV_KFRESET = range(7)
STAT_WIDTH = 40
NOALLOC_SCRATCH_DOUBLES = 262144  # VIEKF_NOALLOC_SCRATCH_DOUBLES
FLAG_NAN_STATE, FLAG_DEPTH_NAN_FIXED, FLAG_DEPTH_NEG_FIXED, FLAG_DEPTH_BIG_FIXED, FLAG_NAN_GAIN, FLAG_BACKWARD_DT = 1, 2, 4, 8, 16, 32


def max_x(N):
    return 17 + 5 * N


def max_dx(N):
    return 16 + 3 * N


class Config(C.Structure):
    _fields_ = [
        ("x0", C.c_double * 17), ("P0", C.c_double * 16), ("Qx", C.c_double * 16), ("lambda_", C.c_double * 16),
        ("Qu", C.c_double * 6), ("P0_feat", C.c_double * 3), ("Qx_feat", C.c_double * 3), ("lambda_feat", C.c_double * 3),
        ("cam_center", C.c_double * 2), ("focal_len", C.c_double * 2), ("q_b_c", C.c_double * 4), ("p_b_c", C.c_double * 3),
        ("q_b_u", C.c_double * 4), ("min_depth", C.c_double), ("keyframe_overlap_threshold", C.c_double),
        ("use_drag_term", C.c_int32), ("use_partial_update", C.c_int32), ("use_keyframe_reset", C.c_int32),
        ("reserved", C.c_int32),
    ]

    def copy(self):
        c = Config()
        C.memmove(C.byref(c), C.byref(self), C.sizeof(Config))
        return c


def default_config():
    """The reference's params/ekf.yaml (lines 7-12 x0, 15-21 P0/P0_feat, 24-30 Qx, 33-34 Qu, 37-43 lambda,
    68-70 flags, 73-74 min_depth/overlap, 80-86 camera, 91 q_b_u)."""
    c = Config()
    c.x0[:] = [0, 0, -2, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0.1]
    c.P0[:] = [1e-3] * 9 + [2e-1] * 3 + [1e-1] * 3 + [1e-6]
    c.Qx[:] = [0.0] * 16
    c.lambda_[:] = [1.0] * 9 + [0.1] * 6 + [0.01]
    c.Qu[:] = [1.0, 1.0, 1.0, 4e-4, 4e-4, 4e-4]
    c.P0_feat[:] = [0.01, 0.01, 0.3]
    c.Qx_feat[:] = [0.0, 0.0, 0.0]
    c.lambda_feat[:] = [1.0, 1.0, 0.4]
    c.cam_center[:] = [315.83184814453125, 242.1165771484375]
    c.focal_len[:] = [611.1864013671875, 611.5557861328125]
    c.q_b_c[:] = [0.9974335273839892, 0.019768487288642146, -0.03564306227555538, 0.05886541830542371]
    c.p_b_c[:] = [0.17363129, -0.02205945, 0.05490228]
    c.q_b_u[:] = [0.993760669165504, 0.0497294816014604, 0.0997086508721388, 0.00498959122946198]
    c.min_depth = 1.5
    c.keyframe_overlap_threshold = 0.8
    c.use_drag_term = 1
    c.use_partial_update = 1
    c.use_keyframe_reset = 1
    return c


def load_config(path):
    """VIEKF(param_file) / load (reference vi_ekf.cpp:101-155): the configuration of a params/ekf.yaml-style file."""
    L = load_library()
    c = Config()
    rc = L.viekf_config_load_yaml(str(path).encode(), C.byref(c))
    if rc != 0:
        raise RuntimeError(f"viekf_config_load_yaml({path}): {rc}: {L.viekf_last_error().decode()}")
    return c


# numpy mirror of viekf_event (248 bytes)
EVENT_DTYPE = np.dtype([
    ("kind", "<i4"), ("type", "<i4"), ("id", "<i4"), ("active", "<i4"), ("zdim", "<i4"), ("rdim", "<i4"),
    ("n_ids", "<i4"), ("ids_off", "<i4"), ("noise", "<i4"), ("aux", "<i4"),
    ("t", "<f8"), ("z", "<f8", (8,)), ("R", "<f8", (9,)), ("sigma", "<f8", (8,)),
], align=True)
assert EVENT_DTYPE.itemsize == 248


class Tape(C.Structure):
    _fields_ = [("events", C.c_void_p), ("n_events", C.c_int64), ("id_pool", C.c_void_p), ("n_ids", C.c_int64),
                ("values", C.c_void_p), ("values_stride", C.c_int64), ("seed", C.c_uint64), ("imu_bias_range", C.c_double * 6)]


def lib_path():
    """In-tree CUDA engine.  VIEKF_LIB (tuning aid) selects another build of the same sources, e.g. the phase-timing one."""
    return os.environ.get("VIEKF_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libviekf_b200.so")


_lib = None


def load_library():
    """Loads the CUDA engine.  There is no CPU fallback: a missing library is a hard error."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise RuntimeError(f"viekf_b200: CUDA engine {p} is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                           "there is no CPU fallback")
    L = C.CDLL(p)
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    pd = C.POINTER(C.c_double)
    sigs = {
        "viekf_config_default": (None, [C.POINTER(Config)]),
        "viekf_batch_create": (i32, [C.POINTER(Config), i64, i32, i32, C.POINTER(vp)]),
        "viekf_batch_destroy": (None, [vp]),
        "viekf_last_error": (C.c_char_p, []),
        "viekf_batch_set_x0": (i32, [vp, vp, i32]),
        "viekf_batch_perturb_x0": (i32, [vp, C.c_uint64]),
        "viekf_batch_get_state": (i32, [vp, i64, i64, vp]),
        "viekf_batch_get_state_noalloc": (i32, [vp, i64, i64, vp]),
        "viekf_batch_get_covariance": (i32, [vp, i64, i64, vp]),
        "viekf_batch_get_features": (i32, [vp, i64, i64, vp, vp]),
        "viekf_batch_get_flags": (i32, [vp, i64, i64, vp, vp]),
        "viekf_batch_get_time": (i32, [vp, i64, i64, vp]),
        "viekf_batch_get_global_pose": (i32, [vp, i64, i64, vp, vp]),
        "viekf_batch_get_covariance_diagonal": (i32, [vp, i64, i64, vp]),
        "viekf_batch_get_depths": (i32, [vp, i64, i64, vp]),
        "viekf_batch_get_zetas": (i32, [vp, i64, i64, vp]),
        "viekf_batch_get_qzetas": (i32, [vp, i64, i64, vp]),
        "viekf_batch_get_depth": (i32, [vp, i32, i64, i64, vp]),
        "viekf_batch_get_feat": (i32, [vp, i32, i64, i64, vp]),
        "viekf_batch_set_imu_bias": (i32, [vp, vp, vp]),
        "viekf_batch_set_drag_term": (i32, [vp, i32]),
        "viekf_batch_get_drag_term": (i32, [vp]),
        "viekf_config_load_yaml": (i32, [C.c_char_p, C.POINTER(Config)]),
        "viekf_batch_set_instance_offset": (i32, [vp, i64]),
        "viekf_fp64_peak_probe": (i32, [i32, pd]),
        "viekf_batch_propagate": (i32, [vp, vp, i32, dbl]),
        "viekf_batch_update": (i32, [vp, i32, vp, i32, i32, vp, i32, i32, i32, dbl, vp]),
        "viekf_batch_init_feature": (i32, [vp, vp, i32, i32, dbl, vp]),
        "viekf_batch_clear_feature": (i32, [vp, i32]),
        "viekf_batch_keep_only_features": (i32, [vp, vp, i32]),
        "viekf_batch_keyframe_reset": (i32, [vp]),
        "viekf_batch_queue_enable": (i32, [vp, i32, i32]),
        "viekf_batch_queue_propagate": (i32, [vp, vp, dbl, i32]),
        "viekf_batch_add_measurement": (i32, [vp, dbl, i32, vp, i32, vp, i32, i32, i32, dbl, C.POINTER(i32)]),
        "viekf_batch_handle_measurements": (i32, [vp, vp, vp, vp, i32, C.POINTER(i32)]),
        "viekf_batch_run_tape": (i32, [vp, C.POINTER(Tape), vp, i64, i64]),
        "viekf_batch_upload_tape": (i32, [vp, C.POINTER(Tape), C.POINTER(vp)]),
        "viekf_batch_free_tape": (None, [vp, vp]),
        "viekf_batch_run_dev_tape": (i32, [vp, vp, i64, i64, i32]),
        "viekf_batch_sync": (i32, [vp]),
        "viekf_batch_stream": (vp, [vp]),
        "viekf_batch_launch_count": (i64, [vp]),
        "viekf_batch_stats_reset": (i32, [vp, i32]),
        "viekf_batch_stats": (i32, [vp, vp, i32]),
        "viekf_batch_stats_dev_ptr": (vp, [vp]),
        "viekf_batch_reset": (i32, [vp]),
    }
    for name, (res, args) in sigs.items():
        f = getattr(L, name)  # AttributeError if the library does not export what the header declares
        f.restype = res
        f.argtypes = args
    L._declared = list(sigs)
    _lib = L
    return L
