"""oracle/gblur_oracle.py — TEST INFRASTRUCTURE ONLY. Never imported by the product path.

Float64 numpy restatement of LibGaussianBlur's one hot path (reflect_101 pad -> row real-FFT
convolution -> column real-FFT convolution -> u8), INCLUDING the reference's Nyquist-bin
behaviour. Parity status: PINNED — tests/test_oracle.py checks this restatement against
(a) the compiled, unmodified reference (oracle/_ref/libgblur_ref.so, built by oracle/Makefile
from /root/reference) whenever that library is present and (b) the golden vectors in
tests/golden/ that were minted from that same compiled reference by
tests/golden/make_golden.py.  The reference itself ships no numeric golden vectors for this
path, so "the reference run here" is the pin.

All citations are file:line under /root/reference/.
"""
from __future__ import annotations

import math

import numpy as np

N_MIN = 32


def gaussian_window(sigma: float, max_width: int = 0) -> int:
    """src/gaussianblur.cpp:11-26. radius is computed in double and narrowed to float;
    width = int(radius*2 + 0.5f) in float arithmetic; clamp to max_width; force odd."""
    sigma32 = np.float32(sigma)
    radius = np.float32(max(float(sigma32) * math.sqrt(2.0 * math.log(255.0)) - 1.0, 0.0))
    width = int(np.float32(radius * np.float32(2.0)) + np.float32(0.5))
    if max_width:
        width = min(width, max_width)
    if width % 2 == 0:
        width += 1
    return width


def is_valid_size(n: int) -> bool:
    """src/gaussianblur.cpp:82-89."""
    r = n
    while r >= 5 * N_MIN and r % 5 == 0:
        r //= 5
    while r >= 3 * N_MIN and r % 3 == 0:
        r //= 3
    while r >= 2 * N_MIN and r % 2 == 0:
        r //= 2
    return r == N_MIN


def nearest_transform_size(n: int) -> int:
    """src/gaussianblur.cpp:91-98."""
    n = max(n, N_MIN)
    n = N_MIN * ((n + N_MIN - 1) // N_MIN)
    while not is_valid_size(n):
        n += N_MIN
    return n


def plan_sizes(rows: int, cols: int, sigma: float):
    """src/gaussianblur.cpp:104-123 -> (ksize, pad, n_rowpass, n_colpass, tz_rows, tz_cols).
    n_rowpass is the transform length along the column index (of one row) — the reference
    calls it sizes[1] / kerf_1D_col; n_colpass is sizes[0] / kerf_1D_row."""
    ksize = gaussian_window(sigma, max(rows, cols))
    pad = (ksize - 1) // 2
    sizes = [rows + 2 * pad, cols + 2 * pad]
    tz = [0, 0]
    for i in range(2):
        if not is_valid_size(sizes[i]):
            new = nearest_transform_size(sizes[i])
            tz[i] = new - sizes[i]
            sizes[i] = new
    return ksize, pad, sizes[1], sizes[0], tz[0], tz[1]


def gaussian_taps(sigma: float, width: int) -> np.ndarray:
    """src/gaussianblur.cpp:28-48. The loop variable y, y*y and s are float; -(y*y)/s is a float
    passed to ::exp(double); the double quotient exp/(pi_f*s) is narrowed to float on store; the
    sum is a sequential float accumulate; each tap is multiplied by float(1/sum)."""
    sigma32 = np.float32(sigma)
    mid = np.float32((width - 1) / 2.0)
    s = np.float32(np.float32(2.0) * sigma32 * sigma32)
    pis = np.float32(np.float32(math.pi) * s)
    taps = np.empty(width, dtype=np.float32)
    y = np.float32(-mid)
    i = 0
    while y <= mid and i < width:
        arg = np.float32(-(np.float32(y * y)) / s)
        taps[i] = np.float32(math.exp(float(arg)) / float(pis))
        y = np.float32(y + np.float32(1.0))
        i += 1
    acc = np.float32(0.0)
    for t in taps:
        acc = np.float32(acc + t)
    inv = np.float32(np.float32(1.0) / acc)
    return (taps * inv).astype(np.float32)


def kernel_spectrum(sigma: float, ksize: int, n: int) -> np.ndarray:
    """src/gaussianblur.cpp:52-59,145-146: taps zero-padded to n, rotated so the centre tap sits
    at index 0, forward real FFT. Returns the n//2+1 REAL gains in float64 (the kernel is even,
    so its spectrum is real; the reference's fp32 pffft result differs by <= 1.7e-7 abs)."""
    taps = gaussian_taps(sigma, ksize).astype(np.float64)
    k = np.zeros(n, dtype=np.float64)
    half = ksize // 2
    for i in range(ksize):
        k[(i - half) % n] = taps[i]
    return np.fft.rfft(k).real


def kerf_ordered(sigma: float, ksize: int, n: int) -> np.ndarray:
    """The spectrum in pffft's 'ordered' real layout [K0, K_{N/2}, Re1, Im1, ...]
    (include/pffft_pommier/pffft.h:143-146) as fp32, imaginary parts exactly 0."""
    kr = kernel_spectrum(sigma, ksize, n)
    out = np.zeros(n, dtype=np.float32)
    out[0] = kr[0]
    out[1] = kr[n // 2]
    out[2::2] = kr[1 : n // 2]
    return out


def effective_gains(sigma: float, ksize: int, n: int) -> np.ndarray:
    """Per-bin multiplier actually applied by pffft_sorted_optimized_convolution
    (src/gaussianblur.cpp:62-79): bins 0..n/2-1 get K[k]/n; slot 1 of the ordered layout is the
    Nyquist bin, and the loop multiplies it by kernel[0] = K[0] (the quirk)."""
    g = kernel_spectrum(sigma, ksize, n).copy()
    g[n // 2] = g[0]
    return g


def pad_reflect101(x: np.ndarray, pad: int, n: int) -> np.ndarray:
    """src/gaussianblur.cpp:217-228 along the last axis: left = x[pad..1], middle = x,
    right = x[T-2 .. T-1-pad], trailing zeros up to n."""
    t = x.shape[-1]
    if pad > t - 1:
        raise ValueError("precondition pad <= tile-1 violated (reference reads out of bounds)")
    out = np.zeros(x.shape[:-1] + (n,), dtype=np.float64)
    out[..., pad : pad + t] = x
    if pad:
        out[..., :pad] = x[..., pad:0:-1]
        out[..., pad + t : pad + t + pad] = x[..., t - 2 : t - 2 - pad : -1] if t - 2 - pad >= 0 else x[..., t - 2 :: -1]
    return out


def conv_lines(x: np.ndarray, sigma: float, ksize: int, pad: int, n: int, chunk: int = 256) -> np.ndarray:
    """One pass of process_channel_tiles (src/gaussianblur.cpp:202-244) along the last axis of a
    2-D float array, all lines independent: pad -> rfft -> x gains (with quirk) -> irfft -> crop."""
    g = effective_gains(sigma, ksize, n)
    t = x.shape[-1]
    out = np.empty_like(x, dtype=np.float64)
    for a in range(0, x.shape[0], chunk):
        b = pad_reflect101(x[a : a + chunk].astype(np.float64), pad, n)
        y = np.fft.irfft(np.fft.rfft(b, axis=-1) * g, n=n, axis=-1)
        out[a : a + chunk] = y[..., pad : pad + t]
    return out


def blur_plane(plane: np.ndarray, sigma: float, return_intermediate: bool = False):
    """Rows then columns on one float plane (src/gaussianblur.cpp:261-280). The reference keeps
    the pass-1 result in fp32; rounding it here keeps the restatement closest to it."""
    rows, cols = plane.shape
    ksize, pad, n_rp, n_cp, _, _ = plan_sizes(rows, cols, sigma)
    inter = conv_lines(plane, sigma, ksize, pad, n_rp).astype(np.float32)
    res = conv_lines(np.ascontiguousarray(inter.T), sigma, ksize, pad, n_cp).T
    if return_intermediate:
        return res, inter
    return res


def to_u8(v: np.ndarray) -> np.ndarray:
    """helpers.hpp:251-252,270: uint8_t(float + 0.5f) — fp32 add, truncation. Values are clamped
    to [0,255] here (the reference's behaviour outside that range is undefined)."""
    f = v.astype(np.float32) + np.float32(0.5)
    return np.clip(np.trunc(f), 0, 255).astype(np.uint8)


def gaussianblur(img: np.ndarray, sigma: float, apply_to_alpha: bool) -> np.ndarray:
    """gaussianblur::gaussianblur (src/gaussianblur.cpp:307-325) on an HWC u8 array; returns a
    new array. sigma<=0 or channels not in {3,4}: image returned unchanged (:310-317,194-197).
    A 2-D array is treated as the additive 1-channel case."""
    if sigma <= 0:
        return img.copy()
    if img.ndim == 2:
        return to_u8(blur_plane(img.astype(np.float64), sigma))
    rows, cols, ch = img.shape
    if ch not in (3, 4):
        return img.copy()
    out = img.copy()
    nproc = 4 if (ch == 4 and apply_to_alpha) else 3
    for c in range(nproc):
        out[..., c] = to_u8(blur_plane(img[..., c].astype(np.float64), sigma))
    return out


def synth_u8(shape, seed: int) -> np.ndarray:
    """Deterministic synthetic input: px[i] = fmix32(lo(i) ^ hi(i)*0x9E3779B9 ^ seed) >> 24."""
    n = int(np.prod(shape))
    i = np.arange(n, dtype=np.uint64)
    h = (i & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    hi = (i >> np.uint64(32)).astype(np.uint32)
    h = h ^ (hi * np.uint32(0x9E3779B9)) ^ np.uint32(seed & 0xFFFFFFFF)
    h ^= h >> np.uint32(16)
    h *= np.uint32(0x85EBCA6B)
    h ^= h >> np.uint32(13)
    h *= np.uint32(0xC2B2AE35)
    h ^= h >> np.uint32(16)
    return (h >> np.uint32(24)).astype(np.uint8).reshape(shape)
