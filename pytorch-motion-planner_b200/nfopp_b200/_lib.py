"""ctypes binding of libnfopp.so (include/nfopp.h).  The library is the product: if it is missing the
package fails loudly - there is no CPU or PyTorch fallback for the hot path."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnfopp.so")
if os.environ.get("NFOPP_LIB"):          # diagnostics: libnfopp_debug.so (device-side bounds asserts) / libnfopp_trace.so (phase trace)
    LIB_PATH = os.path.join(_HERE, os.environ["NFOPP_LIB"])
MAX_LAYERS = 8


class OnfConfig(C.Structure):
    _fields_ = [("mean", C.c_float), ("sigma", C.c_float), ("use_cos", C.c_int32), ("angle", C.c_int32),
                ("hidden", C.c_int32), ("layers", C.c_int32), ("bias", C.c_int32)]


class OnfLayout(C.Structure):
    _fields_ = [("f_pos", C.c_int32), ("n_angle", C.c_int32), ("f_in", C.c_int32), ("in_dim", C.c_int32),
                ("off_enc_w", C.c_int32), ("off_enc_b", C.c_int32), ("off_ang_b", C.c_int32),
                ("off_ang_f", C.c_int32), ("off_w", C.c_int32 * MAX_LAYERS), ("off_b", C.c_int32 * MAX_LAYERS),
                ("off_w_out", C.c_int32), ("off_b_out", C.c_int32), ("n_params", C.c_int32)]


class PlannerConfig(C.Structure):
    _fields_ = [
        ("onf", OnfConfig), ("problems", C.c_int32), ("constrained", C.c_int32), ("trajectory_length", C.c_int32),
        ("onf_lr", C.c_double), ("onf_beta1", C.c_double), ("onf_beta2", C.c_double), ("traj_lr", C.c_double),
        ("traj_beta1", C.c_double), ("traj_beta2", C.c_double), ("adam_eps", C.c_double),
        ("trajectory_random_offset", C.c_double), ("course_random_offset", C.c_double),
        ("collision_weight", C.c_float), ("init_collision_points", C.c_int32),
        ("reparametrize_trajectory_freq", C.c_int32), ("optimize_collision_model_freq", C.c_int32),
        ("random_field_points", C.c_int32), ("collision_point_count", C.c_int32),
        ("angle_offset", C.c_double), ("angle_weight", C.c_float), ("constraint_deltas_weight", C.c_float),
        ("multipliers_lr", C.c_float), ("boundary_weight", C.c_float), ("collision_multipliers_lr", C.c_float),
        ("collision_beta", C.c_float), ("direction_delta_weight", C.c_float),
        ("checker_kind", C.c_int32), ("env_shared", C.c_int32), ("max_obstacles", C.c_int32),
        ("grid_rows", C.c_int32), ("grid_cols", C.c_int32), ("grid_origin_x", C.c_double),
        ("grid_origin_y", C.c_double), ("grid_cell", C.c_double), ("robot_radius", C.c_double),
        ("box", C.c_double * 4), ("has_bounds", C.c_int32), ("bounds", C.c_double * 4), ("border", C.c_double * 4),
        ("rng_mode", C.c_int32), ("seed", C.c_uint64), ("max_choice_uniforms", C.c_int32), ("onf_impl", C.c_int32)]


_P = C.c_void_p
_I = C.c_int
# name -> (restype, argtypes); every symbol include/nfopp.h declares
SIGNATURES = {
    "nfopp_abi_version": (_I, []),
    "nfopp_last_error": (C.c_char_p, []),
    "nfopp_onf_get_layout": (_I, [C.POINTER(OnfConfig), C.POINTER(OnfLayout)]),
    "nfopp_label_boundary": (_I, [_P, _I, _I, _I, _P, _P, _P]),
    "nfopp_label_circle": (_I, [_P, _I, _I, _I, _P, _P, _I, _I, C.c_double, _P, _P, _P]),
    "nfopp_label_rectangle": (_I, [_P, _I, _I, _P, _P, _I, _I, _P, _P, _P, _P]),
    "nfopp_label_grid": (_I, [_P, _I, _I, _I, _P, _I, _I, _I, C.c_double, C.c_double, C.c_double, _P, _P, _P]),
    "nfopp_onf_scratch_bytes": (_I, [C.POINTER(OnfConfig), _I, _I, C.POINTER(C.c_size_t)]),
    "nfopp_onf_scratch_bytes_n": (_I, [C.POINTER(OnfConfig), _I, _I, _I, _I, C.POINTER(C.c_size_t)]),
    "nfopp_onf_forward": (_I, [C.POINTER(OnfConfig), _P, _P, _P, _I, _I, _P, _I, _P, _P]),
    "nfopp_onf_forward_static": (_I, [C.POINTER(OnfConfig), _P, _P, _P, _I, _I, _P, _P]),
    "nfopp_onf_tiles": (_I, [C.POINTER(OnfConfig), _I, _I]),
    "nfopp_onf_backward": (_I, [C.POINTER(OnfConfig), _P, _P, _P, _I, _I, _I, _P, _P, C.c_float, C.c_float, _P, _P,
                                _P, _I, _P, _P]),
    "nfopp_onf_reduce_grad": (_I, [C.POINTER(OnfConfig), _P, _P, _I, _I, _I, _P, _P]),
    "nfopp_onf_adam": (_I, [C.POINTER(OnfConfig), _P, _P, _P, _P, _P, _I, _I, _I, _I, C.c_double, C.c_double,
                            C.c_double, C.c_double, _P]),
    "nfopp_choice": (_I, [_P, _I, _I, _I, _I, _P, _I, _P, _P, _P]),
    "nfopp_engine_workspace_bytes": (_I, [C.POINTER(PlannerConfig), C.POINTER(C.c_size_t)]),
    "nfopp_engine_create": (_I, [C.POINTER(PlannerConfig), _P, C.c_size_t, C.POINTER(_P)]),
    "nfopp_engine_destroy": (_I, [_P]),
    "nfopp_engine_buffer": (_I, [_P, C.c_char_p, C.POINTER(_P), C.POINTER(C.c_size_t), C.POINTER(C.c_char),
                                 C.POINTER(C.c_int32 * 3)]),
    "nfopp_engine_buffer_count": (_I, [_P]),
    "nfopp_engine_buffer_name": (C.c_char_p, [_P, _I]),
    "nfopp_engine_invalidate_weights": (_I, [_P]),
    "nfopp_engine_get_counter": (_I, [_P, _I, C.POINTER(C.c_int64)]),
    "nfopp_engine_set_counter": (_I, [_P, _I, C.c_int64]),
    "nfopp_engine_warmup": (_I, [_P, _I, _P]),
    "nfopp_engine_step": (_I, [_P, _I, _I, _P]),
    "nfopp_engine_profile_step": (_I, [_P, C.POINTER(C.c_float), _I, _P]),
    "nfopp_engine_kernel_name": (C.c_char_p, [_I]),
    "nfopp_engine_get_paths": (_I, [_P, _P, _P]),
    "nfopp_engine_reparametrize": (_I, [_P, _P]),
    "nfopp_engine_evaluate_paths": (_I, [_P, _I, _P, _P, _P, _P]),
    "nfopp_engine_astar_workspace_bytes": (_I, [_P, _I, _I, C.POINTER(C.c_size_t)]),
    "nfopp_engine_init_astar": (_I, [_P, _I, _I, C.c_double, _P, _P, _P, _P, C.c_size_t, _P, _P]),
    "nfopp_engine_path_statistics": (_I, [_P, C.c_double, _I, _P, _P]),
    "nfopp_postprocess_workspace_bytes": (_I, [_I, _I, _P]),
    "nfopp_postprocess_paths": (_I, [_P, _I, _I, C.c_double, C.c_double, _I, _P, C.c_size_t, _P, _P, _P]),
    "nfopp_engine_set_lanes": (_I, [_P, _I]),
    "nfopp_engine_set_bounds": (_I, [_P, _P, _I, _P]),
    "nfopp_engine_launch_count": (_I, [_P, C.POINTER(C.c_int64)]),
}

_lib = None


class NfoppError(RuntimeError):
    pass


def lib():
    """Loads libnfopp.so once.  Raises (never falls back) when the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NfoppError(f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` "
                             "(nvcc, sm_100a).  nfopp_b200 has no CPU fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        if handle.nfopp_abi_version() != 5:
            raise NfoppError("libnfopp.so ABI version mismatch")
        _lib = handle
    return _lib


def check(rc):
    """nfopp_status -> Python exception (ValueError for bad arguments, like the reference's own errors)."""
    if rc == 0:
        return
    msg = lib().nfopp_last_error().decode()
    if rc == -1:
        raise ValueError(msg)
    if rc == -4:
        raise KeyError(msg)
    raise NfoppError(f"libnfopp error {rc}: {msg}")


def onf_scratch(cfg, problems, impl, device, n=0, want_bwd=True):
    """Device scratch an ONF forward/backward call of `impl` on `n` points per problem needs (tensor paths: bf16
    operand copy of the weights; wide path: also its activation / gradient operand tiles)."""
    import torch
    nbytes = C.c_size_t()
    check(lib().nfopp_onf_scratch_bytes_n(C.byref(cfg), problems, n, impl, int(want_bwd), C.byref(nbytes)))
    return torch.empty(max(int(nbytes.value), 16), dtype=torch.uint8, device=device)


def ptr(t):
    """Device (or host) pointer of a torch tensor / None for ctypes."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(device=None):
    import torch
    if not torch.cuda.is_available():
        raise NfoppError("nfopp_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback for the hot path")
    if device is not None and torch.device(device).type != "cuda":
        raise NfoppError(f"nfopp_b200 runs the hot path on CUDA only, got device={device}")
