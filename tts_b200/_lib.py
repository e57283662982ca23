"""ctypes binding of include/tts_b200.h — the only way the Python boundary reaches the kernels.

There is no CPU fallback: if libtts_b200.so is missing or a call fails, a RuntimeError is raised.
"""
import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libtts_b200.so")
ABI_VERSION = 2
MAX_STAGES = 8
F32, BF16 = 0, 1
OUT_F32, OUT_F16 = 0, 1
STATIC_WORKSPACE_BYTES = 32 << 20   # TTSB200_STATIC_WORKSPACE_BYTES
PRECISIONS = {"fp32": F32, "f32": F32, "float32": F32, "bf16": BF16, "bfloat16": BF16}


class GeneratorConfig(ctypes.Structure):
    _fields_ = [
        ("in_channels", ctypes.c_int32),
        ("out_channels", ctypes.c_int32),
        ("proj_kernel", ctypes.c_int32),
        ("base_channels", ctypes.c_int32),
        ("n_stages", ctypes.c_int32),
        ("upsample_factors", ctypes.c_int32 * MAX_STAGES),
        ("res_kernel", ctypes.c_int32),
        ("num_res_blocks", ctypes.c_int32),
        ("pqmf_bands", ctypes.c_int32),
        ("pqmf_taps", ctypes.c_int32),
        ("act_slope", ctypes.c_float),
    ]


class ForwardArgs(ctypes.Structure):
    _fields_ = [
        ("mel", ctypes.c_void_p),
        ("batch", ctypes.c_int32),
        ("frames", ctypes.c_int32),
        ("pad", ctypes.c_int32),
        ("precision", ctypes.c_int32),
        ("out", ctypes.c_void_p),
        ("workspace", ctypes.c_void_p),
        ("workspace_bytes", ctypes.c_size_t),
        ("stream", ctypes.c_void_p),
        ("skip_pqmf", ctypes.c_int32),
        ("n_taps", ctypes.c_int32),
        ("tap_layers", ctypes.POINTER(ctypes.c_int32)),
        ("tap_out", ctypes.POINTER(ctypes.c_void_p)),
        ("profile", ctypes.c_void_p),
        ("out_dtype", ctypes.c_int32),
        ("out_keep_lo", ctypes.c_int64),
        ("out_keep_len", ctypes.c_int64),
        ("out_batch_stride", ctypes.c_int64),
    ]


class MelConfig(ctypes.Structure):
    _fields_ = [
        ("sample_rate", ctypes.c_int32),
        ("n_fft", ctypes.c_int32),
        ("hop_length", ctypes.c_int32),
        ("win_length", ctypes.c_int32),
        ("num_mels", ctypes.c_int32),
        ("preemphasis", ctypes.c_float),
        ("spec_gain", ctypes.c_float),
        ("ref_level_db", ctypes.c_float),
        ("min_level_db", ctypes.c_float),
        ("max_norm", ctypes.c_float),
        ("signal_norm", ctypes.c_int32),
        ("symmetric_norm", ctypes.c_int32),
        ("clip_norm", ctypes.c_int32),
    ]


class LaunchRecord(ctypes.Structure):
    _fields_ = [
        ("name", ctypes.c_char_p),
        ("layer", ctypes.c_int32),
        ("ms", ctypes.c_float),
        ("flops", ctypes.c_double),
        ("bytes", ctypes.c_double),
    ]


# name -> (restype, argtypes); every symbol include/tts_b200.h declares
SYMBOLS = {
    "ttsb200_abi_version": (ctypes.c_int32, []),
    "ttsb200_last_error": (ctypes.c_char_p, []),
    "ttsb200_device_ok": (ctypes.c_int32, [ctypes.c_int32]),
    "ttsb200_expected_tensors": (ctypes.c_int32, [ctypes.POINTER(GeneratorConfig)]),
    "ttsb200_plan_create": (ctypes.c_int32, [ctypes.POINTER(GeneratorConfig), ctypes.POINTER(ctypes.c_void_p),
                                             ctypes.c_int32, ctypes.c_int32, ctypes.POINTER(ctypes.c_void_p)]),
    "ttsb200_plan_destroy": (None, [ctypes.c_void_p]),
    "ttsb200_output_length": (ctypes.c_int64, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32]),
    "ttsb200_workspace_bytes": (ctypes.c_int64, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                                 ctypes.c_int32]),
    "ttsb200_layer_shape": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                             ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int64)]),
    "ttsb200_generator_forward": (ctypes.c_int32, [ctypes.c_void_p, ctypes.POINTER(ForwardArgs)]),
    "ttsb200_generator_forward_static": (ctypes.c_int32, [ctypes.c_void_p, ctypes.POINTER(ForwardArgs)]),
    "ttsb200_generator_forward_host": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32,
                                                        ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                                        ctypes.c_void_p]),
    "ttsb200_pqmf_synthesis": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32,
                                                ctypes.c_int32, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p]),
    "ttsb200_pqmf_factor_4x63": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_void_p]),
    "ttsb200_pqmf_analysis": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32,
                                               ctypes.c_int32, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p]),
    "ttsb200_profile_create": (ctypes.c_int32, [ctypes.POINTER(ctypes.c_void_p)]),
    "ttsb200_profile_destroy": (None, [ctypes.c_void_p]),
    "ttsb200_profile_reset": (None, [ctypes.c_void_p]),
    "ttsb200_profile_read": (ctypes.c_int32, [ctypes.c_void_p, ctypes.POINTER(LaunchRecord), ctypes.c_int32]),
    "ttsb200_launch_count": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32]),
    "ttsb200_mel_plan_create": (ctypes.c_int32, [ctypes.POINTER(MelConfig), ctypes.c_void_p, ctypes.c_void_p,
                                                 ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32,
                                                 ctypes.POINTER(ctypes.c_void_p)]),
    "ttsb200_mel_plan_destroy": (None, [ctypes.c_void_p]),
    "ttsb200_mel_frames": (ctypes.c_int64, [ctypes.c_void_p, ctypes.c_int64]),
    "ttsb200_mel_forward": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64,
                                             ctypes.c_void_p, ctypes.c_void_p]),
    "ttsb200_frame_stats": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_int32,
                                             ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p]),
    "ttsb200_diag_ffma_tflops": (ctypes.c_double, [ctypes.c_int32]),
    "ttsb200_diag_hostlink_gbs": (ctypes.c_double, [ctypes.c_int32, ctypes.c_int32, ctypes.c_int64, ctypes.c_int32]),
}

_lib = None
_lock = threading.Lock()


def load():
    """dlopen the in-tree library and type every entry point. Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"tts_b200: {LIB_PATH} is missing. Build it with `python -m tts_b200.build` "
                "(needs nvcc; sm_100a only). There is no CPU or PyTorch fallback for this path.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the .so is stale
            fn.restype = res
            fn.argtypes = args
        got = lib.ttsb200_abi_version()
        if got != ABI_VERSION:
            raise RuntimeError(f"tts_b200: ABI version mismatch: library {got}, binding {ABI_VERSION}; rebuild")
        _lib = lib
    return _lib


def last_error():
    return (load().ttsb200_last_error() or b"").decode("utf-8", "replace")


def check(rc):
    """Turn a negative status into the RuntimeError the reference's callers expect."""
    if rc is not None and rc < 0:
        raise RuntimeError(last_error() or f"tts_b200 error {rc}")
    return rc
