"""ctypes binding of librnafold_b200.so (the C ABI declared in include/rnafold_b200.h).

The library is the only fold engine of this package.  If it is missing, or no CUDA device can be
initialised, importing callers get an exception -- there is no CPU fallback.
"""
import ctypes
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RNAFOLD_LIB") or os.path.join(_HERE, "librnafold_b200.so")

EXPORTS = [
    "rnafold_init", "rnafold_shutdown", "rnafold_last_error", "rnafold_version",
    "rnafold_read_parameter_file", "rnafold_export_params", "rnafold_parse_params", "rnafold_set_md", "rnafold_get_md",
    "rnafold_workspace_bytes", "rnafold_fold_batch", "rnafold_fold_hamming_batch", "rnafold_fold_hamming_batch_noalloc",
    "rnafold_last_status", "rnafold_last_launch_info", "rnafold_set_timing", "rnafold_last_kernel_ms",
    "rnafold_measure_onchip_peaks", "rnafold_device_info", "rnafold_wave_smem_bytes", "rnafold_current_device",
    "rnafold_env_create", "rnafold_env_destroy", "rnafold_env_config", "rnafold_env_config_reward", "rnafold_env_set_current", "rnafold_env_set_targets", "rnafold_env_reset",
    "rnafold_env_step", "rnafold_env_stats", "rnafold_env_snapshot", "rnafold_parse_target",
    "rnafold_ensemble_batch", "rnafold_eval_structure_batch", "rnafold_permute_batch", "rnafold_permute_draw", "rnafold_permute_set_order", "rnafold_permute_candidate_order", "rnafold_env_config_permute", "rnafold_env_config_delay", "rnafold_env_flush", "rnafold_env_set_stream",
]

_lock = threading.Lock()
_cdll = None
_device = None


class RnaFoldError(RuntimeError):
    pass


def load_library():
    """dlopen the shared library and declare the prototypes (no device needed)."""
    global _cdll
    with _lock:
        if _cdll is not None:
            return _cdll
        if not os.path.exists(LIB_PATH):
            raise RnaFoldError(
                f"{LIB_PATH} is missing: build it with `python -m rnagent_b200.build` "
                "(this package has no CPU fallback)")
        L = ctypes.CDLL(LIB_PATH)
        vp, ci, cd = ctypes.c_void_p, ctypes.c_int, ctypes.c_double
        L.rnafold_init.argtypes = [ci]
        L.rnafold_last_error.restype = ctypes.c_char_p
        L.rnafold_version.restype = ctypes.c_char_p
        L.rnafold_read_parameter_file.argtypes = [ctypes.c_char_p]
        L.rnafold_export_params.argtypes = [vp, ci]
        L.rnafold_parse_params.argtypes = [ctypes.c_char_p, vp, ci]
        L.rnafold_set_md.argtypes = [ctypes.c_char_p, cd]
        L.rnafold_get_md.argtypes = [ctypes.c_char_p, ctypes.POINTER(cd)]
        L.rnafold_workspace_bytes.argtypes = [ci, ci, ctypes.POINTER(ctypes.c_size_t)]
        L.rnafold_fold_batch.argtypes = [vp, vp, ci, vp, vp, vp]
        L.rnafold_fold_hamming_batch.argtypes = [vp, vp, ci, vp, vp, vp, vp, vp, vp, cd, vp]
        L.rnafold_last_launch_info.argtypes = [ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]
        L.rnafold_fold_hamming_batch_noalloc.argtypes = [vp, vp, ci, ci, vp, vp, vp, vp, vp, cd, vp]
        L.rnafold_last_status.argtypes = [ctypes.POINTER(ctypes.c_int32), ci]
        L.rnafold_set_timing.argtypes = [ci]
        L.rnafold_last_kernel_ms.argtypes = [ctypes.POINTER(ctypes.c_float)]
        L.rnafold_measure_onchip_peaks.argtypes = [ctypes.POINTER(cd), ctypes.POINTER(cd)]
        L.rnafold_wave_smem_bytes.argtypes = [ci, ci]
        L.rnafold_device_info.argtypes = [ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]
        i32p, u64 = ctypes.POINTER(ctypes.c_int32), ctypes.c_uint64
        L.rnafold_env_create.argtypes = [ci, ci, i32p]
        L.rnafold_env_destroy.argtypes = [ci]
        L.rnafold_env_config.argtypes = [ci, ci, ci, u64]
        L.rnafold_env_config_reward.argtypes = [ci, ci]
        L.rnafold_env_set_current.argtypes = [ci, ci]
        L.rnafold_env_set_targets.argtypes = [ci, vp, vp, ci]
        L.rnafold_env_reset.argtypes = [ci, vp]
        L.rnafold_env_step.argtypes = [ci, vp, cd, vp, vp, vp]
        L.rnafold_env_stats.argtypes = [ci, vp, vp, ci]
        L.rnafold_env_snapshot.argtypes = [ci, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32p]
        L.rnafold_parse_target.argtypes = [vp, ci, vp, vp]
        L.rnafold_permute_batch.argtypes = [vp, vp, ci, vp, ci, ci, ci, cd, ci, u64, cd, vp, vp, vp, vp, vp, i32p]
        L.rnafold_permute_draw.argtypes = [u64, ci, ci, ci, ci]
        L.rnafold_ensemble_batch.argtypes = [vp, vp, ci, vp, vp, vp, vp, vp, vp, vp, ctypes.c_double, vp]
        L.rnafold_eval_structure_batch.argtypes = [vp, vp, ci, vp, vp, vp]
        L.rnafold_permute_draw.restype = u64
        L.rnafold_permute_set_order.argtypes = [ci]
        L.rnafold_permute_candidate_order.argtypes = [vp, ci, ci, ci, vp]
        L.rnafold_env_config_permute.argtypes = [ci, ci, ci, ci, ci, cd, ci]
        L.rnafold_env_config_delay.argtypes = [ci, ci]
        L.rnafold_env_flush.argtypes = [ci]
        L.rnafold_env_set_stream.argtypes = [ci, vp, ci]
        _cdll = L
        return L


def check(rc):
    if rc < 0:
        msg = load_library().rnafold_last_error().decode(errors="replace")
        if rc == -4:
            raise NotImplementedError(msg)
        if rc == -1:
            raise ValueError(msg)
        if rc == -3:
            raise OSError(msg)
        raise RnaFoldError(f"rnafold error {rc}: {msg}")
    return rc


def lib(device=None):
    """The initialised library bound to one CUDA device (LOCAL_RANK by default)."""
    global _device
    L = load_library()
    if _device is None:
        if device is None:
            device = int(os.environ.get("RNAFOLD_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        check(L.rnafold_init(int(device)))
        _device = int(device)
    elif device is not None and int(device) != _device:
        raise RnaFoldError(f"library already bound to cuda:{_device}")
    return L


def device_index():
    return _device


def _ptr(x):
    """Raw address of a numpy array, a torch tensor (host or CUDA) or None."""
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    raise TypeError(f"unsupported buffer type {type(x)!r}")


def fold_hamming_raw(seq, off, B, target=None, db=None, mfe=None, hd=None, reward=None, mismatch=None,
                     reward_exp=1.0, stream=None):
    """Thin call-through: arguments are buffers (numpy / torch) or None; see rnafold_b200.h."""
    L = lib()
    if target is None:
        return check(L.rnafold_fold_batch(_ptr(seq), _ptr(off), int(B), _ptr(db), _ptr(mfe), stream))
    return check(L.rnafold_fold_hamming_batch(_ptr(seq), _ptr(off), int(B), _ptr(target), _ptr(db), _ptr(mfe),
                                              _ptr(hd), _ptr(reward), _ptr(mismatch), float(reward_exp), stream))


def last_launch_info():
    L = lib()
    n = ctypes.c_int32(0)
    routed = (ctypes.c_int32 * 3)()
    check(L.rnafold_last_launch_info(ctypes.byref(n), routed))
    return n.value, list(routed)


def last_status(clear=True):
    """Wait for the last (asynchronous) fold call; raises RnaFoldError if a kernel reported a failure."""
    v = ctypes.c_int32(0)
    check(lib().rnafold_last_status(ctypes.byref(v), 1 if clear else 0))
    return v.value


def set_timing(enabled):
    check(lib().rnafold_set_timing(1 if enabled else 0))


def last_kernel_ms():
    ms = ctypes.c_float(0)
    check(lib().rnafold_last_kernel_ms(ctypes.byref(ms)))
    return ms.value


def measure_onchip_peaks():
    """(4-byte shared loads per second, min-plus updates per second) measured on the bound device."""
    a, b = ctypes.c_double(0), ctypes.c_double(0)
    check(lib().rnafold_measure_onchip_peaks(ctypes.byref(a), ctypes.byref(b)))
    return a.value, b.value


def device_info():
    a, b = ctypes.c_int32(0), ctypes.c_int32(0)
    check(lib().rnafold_device_info(ctypes.byref(a), ctypes.byref(b)))
    return a.value, b.value


def export_params():
    L = lib()
    out = np.zeros(70000, dtype=np.int32)
    n = check(L.rnafold_export_params(out.ctypes.data, out.size))
    return out[:n].copy()


def parse_params(path=None):
    """Host-only parse of a parameter file (None = built-in Turner 2004) -> flat int32 dump."""
    L = load_library()
    out = np.zeros(70000, dtype=np.int32)
    n = check(L.rnafold_parse_params(os.fsencode(path) if path else None, out.ctypes.data, out.size))
    return out[:n].copy()
