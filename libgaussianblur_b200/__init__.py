"""libgaussianblur_b200 — B200-native (sm_100a) implementation of LibGaussianBlur's hot path.

Python host layer over the C ABI in include/gblur_capi.h (ctypes; torch only supplies device
memory and streams). The reference's own Python surface — module ``gaussianblur`` with
``ImgGeom``, ``Image`` and ``gaussianblur(image, sigma, apply_to_alpha)``
(/root/reference/python/bindings/gaussianblur_py.cpp:7-32) — is the pybind11 module built next to
the library and re-exported here as ``libgaussianblur_b200.gaussianblur_module()``.

There is no CPU implementation in this package: every blur call goes to the CUDA library and
raises if it (or a GPU) is missing.
"""
from __future__ import annotations

import ctypes as C
import importlib.util
import os
import sys
import threading

import numpy as np

from . import build as _build

__all__ = ["Plan", "GblurError", "lib", "gaussianblur", "gaussianblur_module", "host_plan_sizes", "host_spectrum",
           "host_line_plan", "host_kperm", "host_taps", "launch_count", "EXPORTED_SYMBOLS", "HaloPlan", "halo_partition",
           "colsum_noalloc", "NOALLOC_MAX_ROW_BYTES", "host_kernel_choice"]

# every symbol include/gblur_capi.h declares
EXPORTED_SYMBOLS = [
    "gblur_plan_create", "gblur_plan_create_batch", "gblur_plan_create_band", "gblur_plan_destroy",
    "gblur_plan_info", "gblur_plan_spectrum", "gblur_plan_radices", "gblur_plan_last_kernels", "gblur_run_host", "gblur_run_batch_host",
    "gblur_run_device", "gblur_run_batch_device", "gblur_run_stage_device", "gblur_rowpass_device", "gblur_rowpass_split_device", "gblur_colpass_device",
    "gblur_plan_panel_geometry", "gblur_band_panel_alloc", "gblur_ipc_export", "gblur_ipc_open", "gblur_ipc_close",
    "gblur_rowpass_peer_device", "gblur_colpass_panel_device", "gblur_panel_push_device",
    "gblur_halo_partition", "gblur_plan_create_halo", "gblur_halo_info", "gblur_halo_colsum_device", "gblur_halo_run_device",
    "gblur_halo_rowpass_device", "gblur_halo_rowpass2_device", "gblur_halo_colpass_device",
    "gblur_mg_create", "gblur_mg_destroy", "gblur_mg_info", "gblur_mg_run_frames", "gblur_mg_run_bands", "gblur_mg_last_error",
    "gblur_colsum_noalloc",
    "gblur_dump_intermediate", "gblur_launch_count", "gblur_last_error", "gblur_device_count",
    "gblur_host_plan_sizes", "gblur_host_taps", "gblur_host_spectrum", "gblur_host_line_plan", "gblur_host_kperm",
    "gblur_host_kernel_choice",
]


class GblurError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"gblur error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    """The loaded C-ABI library. Fails loudly if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.lib_path()
    if not os.path.exists(path):
        raise ImportError(f"{path} is missing: run `python -m libgaussianblur_b200.build` (needs nvcc); "
                          "there is no CPU fallback")
    L = C.CDLL(path)
    ip, fp, u8p, vp, ll = C.POINTER(C.c_int), C.POINTER(C.c_float), C.c_void_p, C.c_void_p, C.c_longlong
    pp = C.POINTER(C.c_void_p)
    sig = {
        "gblur_plan_create": (C.c_int, [pp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]),
        "gblur_plan_create_batch": (C.c_int, [pp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int]),
        "gblur_plan_create_band": (C.c_int, [pp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int]),
        "gblur_plan_destroy": (None, [vp]),
        "gblur_plan_info": (C.c_int, [vp, ip, ip, ip, ip, ip, ip]),
        "gblur_plan_spectrum": (C.c_int, [vp, C.c_int, fp]),
        "gblur_plan_radices": (C.c_int, [vp, C.c_int, ip, ip, ip]),
        "gblur_plan_last_kernels": (C.c_int, [vp, ip, ip]),
        "gblur_run_host": (C.c_int, [vp, u8p, C.c_int]),
        "gblur_run_batch_host": (C.c_int, [vp, u8p, C.c_int, C.c_int]),
        "gblur_run_device": (C.c_int, [vp, u8p, u8p, C.c_int, vp]),
        "gblur_run_batch_device": (C.c_int, [vp, u8p, u8p, C.c_int, C.c_int, vp]),
        "gblur_run_stage_device": (C.c_int, [vp, u8p, u8p, C.c_int, C.c_int, vp, C.c_int]),
        "gblur_rowpass_device": (C.c_int, [vp, u8p, vp, ll, C.c_int, vp]),
        "gblur_rowpass_split_device": (C.c_int, [vp, u8p, vp, C.c_int, ll, C.c_int, vp]),
        "gblur_colpass_device": (C.c_int, [vp, vp, ll, u8p, ll, C.c_int, vp]),
        "gblur_plan_panel_geometry": (C.c_int, [vp, ip, ip]),
        "gblur_band_panel_alloc": (C.c_int, [vp, C.c_int, pp, C.POINTER(C.c_longlong)]),
        "gblur_ipc_export": (C.c_int, [vp, C.c_char_p]),
        "gblur_ipc_open": (C.c_int, [C.c_char_p, pp]),
        "gblur_ipc_close": (C.c_int, [vp]),
        "gblur_rowpass_peer_device": (C.c_int, [vp, u8p, C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, vp]),
        "gblur_colpass_panel_device": (C.c_int, [vp, vp, u8p, ll, C.c_int, vp]),
        "gblur_panel_push_device": (C.c_int, [vp, C.c_int, vp, vp, C.c_int, C.c_int, vp]),
        "gblur_halo_partition": (C.c_int, [C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, ip, ip, ip, ip]),
        "gblur_plan_create_halo": (C.c_int, [pp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int]),
        "gblur_halo_info": (C.c_int, [vp, ip, ip, ip, ip]),
        "gblur_halo_colsum_device": (C.c_int, [vp, u8p, vp, vp]),
        "gblur_colsum_noalloc": (C.c_int, [C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_float, vp, vp]),
        "gblur_halo_run_device": (C.c_int, [vp, u8p, vp, u8p, vp]),
        "gblur_halo_rowpass_device": (C.c_int, [vp, u8p, C.c_int, C.c_int, vp]),
        "gblur_halo_rowpass2_device": (C.c_int, [vp, u8p, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
        "gblur_halo_colpass_device": (C.c_int, [vp, vp, u8p, vp]),
        "gblur_mg_create": (C.c_int, [pp, C.c_int, ip, C.c_int, C.c_int, C.c_int, C.c_float]),
        "gblur_mg_destroy": (None, [vp]),
        "gblur_mg_info": (C.c_int, [vp, ip, ip, ip, ip]),
        "gblur_mg_run_frames": (C.c_int, [vp, u8p, C.c_int, C.c_int]),
        "gblur_mg_run_bands": (C.c_int, [vp, u8p, C.c_int]),
        "gblur_mg_last_error": (C.c_char_p, []),
        "gblur_dump_intermediate": (C.c_int, [vp, fp, C.c_int]),
        "gblur_launch_count": (C.c_ulonglong, []),
        "gblur_last_error": (C.c_char_p, []),
        "gblur_device_count": (C.c_int, []),
        "gblur_host_plan_sizes": (C.c_int, [C.c_int, C.c_int, C.c_float, ip, ip, ip, ip, ip, ip]),
        "gblur_host_taps": (C.c_int, [C.c_float, C.c_int, fp]),
        "gblur_host_spectrum": (C.c_int, [C.c_float, C.c_int, C.c_int, fp]),
        "gblur_host_line_plan": (C.c_int, [C.c_int, ip, ip, ip, ip, C.POINTER(C.c_longlong), ip]),
        "gblur_host_kperm": (C.c_int, [C.c_float, C.c_int, C.c_int, fp]),
        "gblur_host_kernel_choice": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, ip]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _check(rc: int):
    if rc != 0:
        raise GblurError(rc, lib().gblur_last_error().decode())


def launch_count() -> int:
    return int(lib().gblur_launch_count())


# ---------------------------------------------------------------------------------------------
# host-only semantics (no GPU needed)
# ---------------------------------------------------------------------------------------------
def host_plan_sizes(rows: int, cols: int, sigma: float) -> dict:
    v = [C.c_int() for _ in range(6)]
    _check(lib().gblur_host_plan_sizes(rows, cols, sigma, *[C.byref(x) for x in v]))
    k = ["ksize", "pad", "n_rowpass", "n_colpass", "tz_rows", "tz_cols"]
    return dict(zip(k, [x.value for x in v]))


def host_taps(sigma: float, ksize: int) -> np.ndarray:
    out = np.empty(ksize, dtype=np.float32)
    _check(lib().gblur_host_taps(sigma, ksize, out.ctypes.data_as(C.POINTER(C.c_float))))
    return out


def host_spectrum(sigma: float, ksize: int, n: int) -> np.ndarray:
    out = np.empty(n, dtype=np.float32)
    _check(lib().gblur_host_spectrum(sigma, ksize, n, out.ctypes.data_as(C.POINTER(C.c_float))))
    return out


def host_line_plan(n: int) -> dict:
    rad = (C.c_int * 5)()
    ns, nseq, st, nt = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    smem = C.c_longlong()
    _check(lib().gblur_host_line_plan(n, rad, C.byref(ns), C.byref(nseq), C.byref(st), C.byref(smem), C.byref(nt)))
    return dict(radices=[rad[i] for i in range(ns.value)], nseq=nseq.value, static=bool(st.value),
                smem_bytes=smem.value, nthreads=nt.value)


def host_kperm(sigma: float, ksize: int, n: int) -> np.ndarray:
    out = np.empty(n, dtype=np.float32)
    _check(lib().gblur_host_kperm(sigma, ksize, n, out.ctypes.data_as(C.POINTER(C.c_float))))
    return out


def host_kernel_choice(rows: int, cols: int, channels: int, sigma: float, apply_to_alpha: bool = True) -> int:
    """2: both passes run as overlap-save blocks, 1: the row pass only, 0: whole-line transforms (host only)."""
    v = C.c_int()
    _check(lib().gblur_host_kernel_choice(rows, cols, channels, sigma, int(bool(apply_to_alpha)), C.byref(v)))
    return v.value


# ---------------------------------------------------------------------------------------------
# GPU plan
# ---------------------------------------------------------------------------------------------
def _ptr(t):
    """Device pointer of a torch tensor / int / None."""
    if t is None:
        return None
    if isinstance(t, int):
        return C.c_void_p(t)
    return C.c_void_p(t.data_ptr())


def _stream_ptr(stream):
    if stream is None:
        import torch

        return C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if isinstance(stream, int):
        return C.c_void_p(stream)
    return C.c_void_p(stream.cuda_stream)


class Plan:
    """Immutable GPU plan for one geometry + sigma (the GPU counterpart of KernelDFT,
    /root/reference/include/gaussianblur/helpers.hpp:125-132)."""

    def __init__(self, rows: int, cols: int, channels: int, sigma: float, device: int = 0, max_frames: int = 1,
                 band_rows: int | None = None, band_cols: int | None = None):
        self._h = C.c_void_p()
        L = lib()
        if band_rows is not None or band_cols is not None:
            _check(L.gblur_plan_create_band(C.byref(self._h), device, rows, cols, channels, sigma,
                                            band_rows or rows, band_cols or cols))
        else:
            _check(L.gblur_plan_create_batch(C.byref(self._h), device, rows, cols, channels, sigma, max_frames))
        self.rows, self.cols, self.channels, self.sigma = rows, cols, channels, float(sigma)
        self.device, self.max_frames = device, max_frames
        self.band_rows, self.band_cols = band_rows, band_cols
        v = [C.c_int() for _ in range(6)]
        _check(L.gblur_plan_info(self._h, *[C.byref(x) for x in v]))
        self.ksize, self.pad, self.n_rowpass, self.n_colpass, self.tz_rows, self.tz_cols = [x.value for x in v]

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().gblur_plan_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def spectrum(self, axis: int) -> np.ndarray:
        out = np.empty(self.n_rowpass if axis == 0 else self.n_colpass, dtype=np.float32)
        _check(lib().gblur_plan_spectrum(self._h, axis, out.ctypes.data_as(C.POINTER(C.c_float))))
        return out

    def radices(self, axis: int):
        rad = (C.c_int * 5)()
        n, st = C.c_int(), C.c_int()
        _check(lib().gblur_plan_radices(self._h, axis, rad, C.byref(n), C.byref(st)))
        return [rad[i] for i in range(n.value)], bool(st.value)

    def last_kernels(self):
        """(row pass, column pass) kernels of the last run: "whole_line" or "overlap_save_1024"."""
        r, c = C.c_int(), C.c_int()
        _check(lib().gblur_plan_last_kernels(self._h, C.byref(r), C.byref(c)))
        names = ("whole_line", "overlap_save_1024")
        return names[r.value], names[c.value]

    def processed_channels(self, apply_to_alpha: bool) -> int:
        return 3 if (self.channels == 4 and not apply_to_alpha) else self.channels

    # -- host buffers (H2D + blur + D2H inside the call) --
    def run_host(self, img: np.ndarray, apply_to_alpha: bool = False, n_frames: int | None = None) -> np.ndarray:
        """Blur a C-contiguous uint8 numpy array (or pinned torch CPU tensor) in place."""
        if hasattr(img, "data_ptr"):
            ptr, nbytes = C.c_void_p(img.data_ptr()), img.numel()
        else:
            assert img.dtype == np.uint8 and img.flags.c_contiguous and img.flags.writeable
            ptr, nbytes = C.c_void_p(img.ctypes.data), img.size
        frames = n_frames if n_frames is not None else 1
        assert nbytes == frames * self.rows * self.cols * self.channels, "buffer size does not match the plan"
        _check(lib().gblur_run_batch_host(self._h, ptr, frames, int(bool(apply_to_alpha))))
        return img

    # -- device buffers (torch uint8 CUDA tensors), asynchronous on the current stream --
    def run_device(self, d_in, d_out=None, apply_to_alpha: bool = False, n_frames: int = 1, stream=None):
        d_out = d_in if d_out is None else d_out
        _check(lib().gblur_run_batch_device(self._h, _ptr(d_in), _ptr(d_out), n_frames, int(bool(apply_to_alpha)),
                                            _stream_ptr(stream)))
        return d_out

    def run_stage(self, stages: int, d_in, d_out=None, apply_to_alpha: bool = False, n_frames: int = 1, stream=None):
        """Profiling hook: 1 = row pass only, 2 = column pass only, 3 = both (same kernels as run_device), 4 = the column
        kernel alone (an overlap-save column pass reuses the column terms of its previous run on this plan)."""
        d_out = d_in if d_out is None else d_out
        _check(lib().gblur_run_stage_device(self._h, _ptr(d_in), _ptr(d_out), n_frames, int(bool(apply_to_alpha)),
                                            _stream_ptr(stream), stages))

    def rowpass(self, d_in, d_inter, inter_pitch_floats: int, apply_to_alpha: bool = False, stream=None):
        _check(lib().gblur_rowpass_device(self._h, _ptr(d_in), _ptr(d_inter), inter_pitch_floats,
                                          int(bool(apply_to_alpha)), _stream_ptr(stream)))

    def rowpass_split(self, d_in, d_inter, split_cols: int, split_stride_floats: int, apply_to_alpha: bool = False,
                      stream=None):
        """Row pass writing one contiguous [band_rows, split_cols*Cp] block per column band (all-to-all send buffer)."""
        _check(lib().gblur_rowpass_split_device(self._h, _ptr(d_in), _ptr(d_inter), split_cols, split_stride_floats,
                                                int(bool(apply_to_alpha)), _stream_ptr(stream)))

    # -- fused exchange: panel buffers written by the peers' row passes --
    def panel_alloc(self, apply_to_alpha: bool = False):
        """(device pointer, floats) of a zeroed panel buffer for this column-band plan (owned by the plan)."""
        ptr, n = C.c_void_p(), C.c_longlong()
        _check(lib().gblur_band_panel_alloc(self._h, int(bool(apply_to_alpha)), C.byref(ptr), C.byref(n)))
        return ptr.value, n.value

    def rowpass_peer(self, d_in, peer_ptrs, row_offset: int, apply_to_alpha: bool = False, stream=None):
        arr = (C.c_void_p * len(peer_ptrs))(*peer_ptrs)
        _check(lib().gblur_rowpass_peer_device(self._h, _ptr(d_in), arr, len(peer_ptrs), row_offset,
                                               int(bool(apply_to_alpha)), _stream_ptr(stream)))

    def colpass_panel(self, d_panel: int, d_out, out_pitch_bytes: int, apply_to_alpha: bool = False, stream=None):
        _check(lib().gblur_colpass_panel_device(self._h, C.c_void_p(d_panel), _ptr(d_out), out_pitch_bytes,
                                                int(bool(apply_to_alpha)), _stream_ptr(stream)))

    def colpass(self, d_inter, inter_pitch_floats: int, d_out, out_pitch_bytes: int, apply_to_alpha: bool = False,
                stream=None):
        _check(lib().gblur_colpass_device(self._h, _ptr(d_inter), inter_pitch_floats, _ptr(d_out), out_pitch_bytes,
                                          int(bool(apply_to_alpha)), _stream_ptr(stream)))

    def panel_push(self, src_panel: int, dst_panel: int, slot_begin: int, slot_end: int, apply_to_alpha: bool = False,
                   stream=None):
        """Copy slots [slot_begin, slot_end) of every panel from one panel buffer to another (column-band plan);
        dst may be the IPC mapping of a peer's buffer. Asynchronous on `stream`."""
        _check(lib().gblur_panel_push_device(self._h, int(bool(apply_to_alpha)), C.c_void_p(src_panel),
                                             C.c_void_p(dst_panel), slot_begin, slot_end, _stream_ptr(stream)))

    def dump_intermediate(self, apply_to_alpha: bool = False) -> np.ndarray:
        cp = self.processed_channels(apply_to_alpha)
        out = np.empty((self.rows, self.cols * cp), dtype=np.float32)
        _check(lib().gblur_dump_intermediate(self._h, out.ctypes.data_as(C.POINTER(C.c_float)),
                                             int(bool(apply_to_alpha))))
        return out


def halo_partition(rows: int, cols: int, sigma: float, n_ranks: int, rank: int) -> dict:
    """Rows a rank writes / reads under halo sharding (gblur_halo_partition; host only)."""
    v = [C.c_int() for _ in range(4)]
    _check(lib().gblur_halo_partition(rows, cols, sigma, n_ranks, rank, *[C.byref(x) for x in v]))
    return dict(zip(("out_row0", "out_row1", "in_row0", "in_row1"), (x.value for x in v)))


NOALLOC_MAX_ROW_BYTES = 262144  # include/gblur_capi.h: GBLUR_NOALLOC_MAX_ROW_BYTES


def colsum_noalloc(d_in, rows: int, cols: int, channels: int, sigma: float, d_colsum, device: int = 0, stream=None):
    """gblur_colsum_noalloc: the weighted column sums of a device-resident u8 image into d_colsum (cols * channels int32),
    accumulated in the library's static device scratch — no allocation on the call path. Calls that overlap on the GPU
    race (they share the scratch). No counterpart in the reference."""
    _check(lib().gblur_colsum_noalloc(device, _ptr(d_in), rows, cols, channels, sigma, _ptr(d_colsum), _stream_ptr(stream)))


class HaloPlan:
    """One rank's share of a halo-sharded blur (include/gblur_capi.h: row bands for both passes, overlap-save column
    blocks; the ranks exchange `pad` input rows and one all-reduce of cols * channels integers)."""

    def __init__(self, rows: int, cols: int, channels: int, sigma: float, n_ranks: int, rank: int, device: int = 0):
        self._h = C.c_void_p()
        _check(lib().gblur_plan_create_halo(C.byref(self._h), device, rows, cols, channels, sigma, n_ranks, rank))
        self.rows, self.cols, self.channels, self.sigma, self.device = rows, cols, channels, float(sigma), device
        v = [C.c_int() for _ in range(4)]
        _check(lib().gblur_halo_info(self._h, *[C.byref(x) for x in v]))
        self.out_row0, self.out_row1, self.in_row0, self.in_row1 = [x.value for x in v]
        w = [C.c_int() for _ in range(6)]
        _check(lib().gblur_plan_info(self._h, *[C.byref(x) for x in w]))
        self.pad = w[1].value

    def colsum(self, d_in_rows, d_u, stream=None):
        _check(lib().gblur_halo_colsum_device(self._h, _ptr(d_in_rows), _ptr(d_u), _stream_ptr(stream)))

    def run(self, d_in_rows, d_u_total, d_out, stream=None):
        _check(lib().gblur_halo_run_device(self._h, _ptr(d_in_rows), _ptr(d_u_total), _ptr(d_out), _stream_ptr(stream)))

    def rowpass(self, d_in_rows, row_begin: int, row_end: int, stream=None):
        _check(lib().gblur_halo_rowpass_device(self._h, _ptr(d_in_rows), row_begin, row_end, _stream_ptr(stream)))

    def rowpass2(self, d_in_rows, begin_a: int, end_a: int, begin_b: int, end_b: int, stream=None):
        """row pass of two row ranges in one launch (the halo rows above and below the band)"""
        _check(lib().gblur_halo_rowpass2_device(self._h, _ptr(d_in_rows), begin_a, end_a, begin_b, end_b, _stream_ptr(stream)))

    def colpass(self, d_u_total, d_out, stream=None):
        _check(lib().gblur_halo_colpass_device(self._h, _ptr(d_u_total), _ptr(d_out), _stream_ptr(stream)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().gblur_plan_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close


def ipc_export(d_ptr: int) -> bytes:
    buf = C.create_string_buffer(64)
    _check(lib().gblur_ipc_export(C.c_void_p(d_ptr), buf))
    return buf.raw


def ipc_open(handle: bytes) -> int:
    out = C.c_void_p()
    _check(lib().gblur_ipc_open(C.create_string_buffer(handle, 64), C.byref(out)))
    return out.value


def ipc_close(d_ptr: int) -> None:
    _check(lib().gblur_ipc_close(C.c_void_p(d_ptr)))


def _device_view(obj):
    """(device pointer, shape, device index or None, stream handle or None) when `obj` lives in GPU memory:
    a torch CUDA tensor or anything exporting __cuda_array_interface__ (CuPy, Numba, ...). None for host data."""
    if hasattr(obj, "is_cuda") and hasattr(obj, "data_ptr"):
        if not obj.is_cuda:
            return None
        if str(obj.dtype) != "torch.uint8" or not obj.is_contiguous():
            raise TypeError("device tensor must be contiguous uint8")
        return obj.data_ptr(), tuple(obj.shape), obj.device.index, None
    cai = getattr(obj, "__cuda_array_interface__", None)
    if cai is None:
        return None
    shape = tuple(cai["shape"])
    if cai["typestr"] not in ("|u1", "<u1", "u1"):
        raise TypeError("device array must be uint8")
    strides = cai.get("strides")
    if strides is not None:
        want, acc = [], 1
        for n in reversed(shape):
            want.append(acc)
            acc *= n
        if tuple(strides) != tuple(reversed(want)):
            raise TypeError("device array must be C-contiguous")
    if cai["data"][1]:
        raise TypeError("device array is read-only")
    return int(cai["data"][0]), shape, None, cai.get("stream")


def gaussianblur(img, sigma: float, apply_to_alpha: bool = False, device: int = 0):
    """Blur an HWC (channels 3/4) or HW uint8 image in place on the GPU and return it.

    `img` is a numpy array or a pinned/pageable torch CPU tensor (copied to the GPU and back, pipelined),
    or — zero copy — a torch CUDA tensor / any object with __cuda_array_interface__: the kernels then
    read and write the caller's device memory directly, stream-ordered on the caller's current stream.
    The reference's Python surface moves pixels through list[int] <-> std::vector<uint8_t>
    (/root/reference/python/bindings/gaussianblur_py.cpp:20); this entry point never touches them on the host.
    Error behaviour follows gaussianblur::gaussianblur (/root/reference/src/gaussianblur.cpp:307-317):
    sigma <= 0 or an unsupported channel count leaves the data untouched."""
    if sigma <= 0:
        print("Invalid smoothing factor")
        return img
    dev = _device_view(img)
    shape = dev[1] if dev else tuple(img.shape)
    if len(shape) not in (2, 3):
        raise ValueError("expected an HW or HWC image")
    ch = 1 if len(shape) == 2 else shape[2]
    if ch not in (1, 3, 4):
        print("Unsupported number of channels", file=sys.stderr)
        return img
    if dev:
        ptr, _, index, stream = dev
        dv = device if index is None else index
        with _cached_plan(shape[0], shape[1], ch, sigma, dv) as plan:
            if isinstance(stream, int) and stream > 2:
                st = C.c_void_p(int(stream))
            elif index is None:
                st = C.c_void_p(0)
            else:  # the current stream of the TENSOR's device, not of whatever device happens to be current
                import torch

                st = C.c_void_p(torch.cuda.current_stream(dv).cuda_stream)
            _check(lib().gblur_run_batch_device(plan._h, C.c_void_p(ptr), C.c_void_p(ptr), 1,
                                                int(bool(apply_to_alpha)), st))
        return img
    with _cached_plan(shape[0], shape[1], ch, sigma, device) as plan:
        plan.run_host(img, apply_to_alpha)
    return img


# A few plans kept between gaussianblur() calls (the C++ layer keeps its own, csrc/gaussianblur_api.cpp): building one
# costs table generation, a workspace allocation of the size of the fp32 intermediate and its zero fill — far more
# than a blur. A plan serves one call at a time (its lock is held for the call); work queued on a caller's stream
# stays valid because cached plans are not destroyed behind it — eviction synchronises the device first.
_plan_cache: "dict" = {}
_plan_cache_lock = threading.Lock()
_PLAN_CACHE_CAP = 4


class _PlanLease:
    def __init__(self, entry):
        self.entry = entry

    def __enter__(self):
        self.entry[1].acquire()
        return self.entry[0]

    def __exit__(self, *exc):
        self.entry[1].release()
        return False


def _cached_plan(rows, cols, ch, sigma, device):
    key = (int(device), int(rows), int(cols), int(ch), float(sigma))
    with _plan_cache_lock:
        entry = _plan_cache.pop(key, None)
        if entry is None:
            entry = (Plan(rows, cols, ch, sigma, device=device), threading.Lock())
        _plan_cache[key] = entry  # most recently used last
        while len(_plan_cache) > _PLAN_CACHE_CAP:
            old_key = next(iter(_plan_cache))
            old = _plan_cache[old_key]
            if not old[1].acquire(blocking=False):
                break  # in use: keep it for now
            del _plan_cache[old_key]
            try:
                import torch

                torch.cuda.synchronize(old_key[0])  # nothing queued on a caller's stream may still use its workspace
            except Exception:
                pass
            old[0].close()
            old[1].release()
    return _PlanLease(entry)


def gaussianblur_module():
    """The pybind11 module `gaussianblur` (reference-compatible Python surface)."""
    name = "gaussianblur"
    if name in sys.modules and getattr(sys.modules[name], "__file__", "") == _build.pymod_path():
        return sys.modules[name]
    path = _build.pymod_path()
    if not os.path.exists(path):
        raise ImportError(f"{path} is missing: run `python -m libgaussianblur_b200.build`")
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sys.modules[name] = mod
    return mod
