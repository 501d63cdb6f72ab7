"""oracle/ref_lib.py — TEST INFRASTRUCTURE ONLY. ctypes loader for the compiled, unmodified
reference (oracle/_ref/libgblur_ref.so, see oracle/Makefile and oracle/ref_harness.cpp).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this. The library is prebuilt in the dev container and travels to the GPU box;
/root/reference itself is never read at run time."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")


def build(reference_root: str = "/root/reference") -> bool:
    """Compile oracle/_ref from the reference sources where they lie (dev container only)."""
    if not os.path.isdir(reference_root):
        return available()
    subprocess.check_call(["make", "-s", "-C", _HERE, f"REF={reference_root}"])
    return available()


def available(single_thread: bool = False) -> bool:
    return os.path.exists(_path(single_thread))


def _path(single_thread: bool) -> str:
    return os.path.join(_REF_DIR, "libgblur_ref_st.so" if single_thread else "libgblur_ref.so")


_libs = {}


def _lib(single_thread: bool = False):
    if single_thread in _libs:
        return _libs[single_thread]
    lib = C.CDLL(_path(single_thread))
    u8p, f32p, ip = C.POINTER(C.c_uint8), C.POINTER(C.c_float), C.POINTER(C.c_int)
    lib.ref_hw_threads.restype = C.c_int
    lib.ref_gaussianblur.restype = C.c_double
    lib.ref_gaussianblur.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int]
    lib.ref_kernel_dft.restype = C.c_int
    lib.ref_kernel_dft.argtypes = [C.c_int, C.c_int, C.c_float, ip, ip, ip, ip, ip, f32p, f32p]
    lib.ref_blur_plane_f32.restype = C.c_double
    lib.ref_blur_plane_f32.argtypes = [f32p, C.c_int, C.c_int, C.c_float, f32p]
    lib.ref_blur_plane_u8.restype = C.c_double
    lib.ref_blur_plane_u8.argtypes = [u8p, C.c_int, C.c_int, C.c_float]
    lib.ref_rowpass_plane.restype = C.c_int
    lib.ref_rowpass_plane.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, f32p]
    _libs[single_thread] = lib
    return lib


def hw_threads() -> int:
    return int(_lib().ref_hw_threads())


def _u8(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _f32(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def gaussianblur(img: np.ndarray, sigma: float, apply_to_alpha: bool, single_thread: bool = False,
                 return_seconds: bool = False):
    """gaussianblur::gaussianblur on an HWC u8 array (channels may be any count — the reference
    itself leaves non-{3,4} images untouched). Returns a new array (and wall seconds)."""
    assert img.dtype == np.uint8 and img.ndim == 3
    out = np.ascontiguousarray(img).copy()
    rows, cols, ch = out.shape
    secs = _lib(single_thread).ref_gaussianblur(_u8(out), rows, cols, ch, float(sigma), int(bool(apply_to_alpha)))
    return (out, secs) if return_seconds else out


def kernel_dft(rows: int, cols: int, sigma: float):
    """prepare_kernel_DFT -> dict(pad, n_rowpass, n_colpass, tz_rows, tz_cols, kerf_col, kerf_row)."""
    lib = _lib()
    v = [C.c_int() for _ in range(5)]
    lib.ref_kernel_dft(rows, cols, float(sigma), *[C.byref(x) for x in v], None, None)
    pad, n_rp, n_cp, tz_r, tz_c = [x.value for x in v]
    kc = np.empty(n_rp, dtype=np.float32)
    kr = np.empty(n_cp, dtype=np.float32)
    lib.ref_kernel_dft(rows, cols, float(sigma), None, None, None, None, None, _f32(kc), _f32(kr))
    return dict(pad=pad, n_rowpass=n_rp, n_colpass=n_cp, tz_rows=tz_r, tz_cols=tz_c, kerf_col=kc, kerf_row=kr)


def blur_plane_f32(plane: np.ndarray, sigma: float, want_intermediate: bool = False,
                   return_seconds: bool = False):
    """One fp32 plane through the reference's two process_channel_tiles calls."""
    p = np.ascontiguousarray(plane, dtype=np.float32).copy()
    rows, cols = p.shape
    inter = np.empty((cols, rows), dtype=np.float32) if want_intermediate else None
    secs = _lib().ref_blur_plane_f32(_f32(p), rows, cols, float(sigma), _f32(inter) if want_intermediate else None)
    res = (p, inter.T.copy()) if want_intermediate else p
    return (res, secs) if return_seconds else res


def blur_plane_u8(plane: np.ndarray, sigma: float, return_seconds: bool = False):
    p = np.ascontiguousarray(plane, dtype=np.uint8).copy()
    rows, cols = p.shape
    secs = _lib().ref_blur_plane_u8(_u8(p), rows, cols, float(sigma))
    return (p, secs) if return_seconds else p


def band_strip_u8(plane: np.ndarray, sigma: float, c0: int, ncols: int, band_rows: int = 1024,
                  return_seconds: bool = False):
    """Columns [c0, c0 + ncols) of the blurred plane of a single-channel image that may exceed the reference's own
    int rows*cols envelope (BASELINE.json config 5): pass 1 in row bands, pass 2 on the strip's columns, both through
    the reference's process_channel_tiles with the spectra of the full geometry (oracle/ref_harness.cpp)."""
    import ctypes

    p = np.ascontiguousarray(plane, dtype=np.uint8)
    rows, cols = p.shape
    out = np.empty((rows, ncols), dtype=np.uint8)
    fn = _lib().ref_band_strip_u8
    fn.restype = ctypes.c_double
    secs = fn(_u8(p), rows, cols, ctypes.c_float(float(sigma)), int(c0), int(ncols), int(band_rows), _u8(out))
    return (out, secs) if return_seconds else out


def rowpass_plane(img: np.ndarray, ch: int, sigma: float) -> np.ndarray:
    """fp32 row-pass intermediate of channel ch, image orientation (rows x cols)."""
    rows, cols, c = img.shape
    out = np.empty((rows, cols), dtype=np.float32)
    _lib().ref_rowpass_plane(_u8(np.ascontiguousarray(img)), rows, cols, c, ch, float(sigma), _f32(out))
    return out
