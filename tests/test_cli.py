"""Desktop command line (examples/desktop/main.cpp): same usage, messages and <name>_pffft.<ext> output naming
as the reference's example (/root/reference/examples/desktop/main.cpp:52-99). Argument handling is checked
without a GPU; the blur itself (-m gpu) is compared with the library call on the same pixels."""
import os
import struct
import subprocess

import numpy as np
import pytest

from libgaussianblur_b200 import build as _build
from oracle import gblur_oracle as go


def _cli(*args, cwd=None):
    return subprocess.run([_build.cli_path(), *args], capture_output=True, text=True, cwd=cwd, timeout=300)


def _write_bmp(path, img):
    h, w, ch = img.shape
    stride = (w * ch + 3) & ~3
    rows = np.zeros((h, stride), dtype=np.uint8)
    bgr = img[..., [2, 1, 0] + ([3] if ch == 4 else [])]
    rows[:, : w * ch] = bgr.reshape(h, w * ch)[::-1]
    hdr = struct.pack("<2sIHHI", b"BM", 54 + stride * h, 0, 0, 54)
    hdr += struct.pack("<IiiHHIIiiII", 40, w, h, 1, ch * 8, 0, stride * h, 2835, 2835, 0, 0)
    open(path, "wb").write(hdr + rows.tobytes())


def _read_bmp(path):
    f = open(path, "rb").read()
    off, = struct.unpack_from("<I", f, 10)
    w, h = struct.unpack_from("<ii", f, 18)
    bpp, = struct.unpack_from("<H", f, 28)
    ch = bpp // 8
    stride = (w * ch + 3) & ~3
    rows = np.frombuffer(f, dtype=np.uint8, count=stride * abs(h), offset=off).reshape(abs(h), stride)
    px = rows[:, : w * ch].reshape(abs(h), w, ch)
    if h > 0:
        px = px[::-1]
    return np.ascontiguousarray(px[..., [2, 1, 0] + ([3] if ch == 4 else [])])


def test_cli_usage_and_argument_errors(tmp_path):
    _build.build()
    r = _cli("--help")
    assert r.returncode == 0 and r.stdout.startswith("Usage: gaussianblur <smoothing_factor> <input_file>")
    r = _cli()
    assert r.returncode == 1 and "Invalid number of arguments." in r.stderr
    r = _cli("0", "x.bmp")
    assert r.returncode == 1 and "Invalid smoothing factor" in r.stdout
    r = _cli("2.0", str(tmp_path / "missing.bmp"))
    assert r.returncode == 1 and "Failed to load image" in r.stderr
    (tmp_path / "a.png").write_bytes(b"\x89PNG")
    r = _cli("2.0", str(tmp_path / "a.png"))
    assert r.returncode == 1 and "not supported" in r.stderr
    gray = tmp_path / "g.pam"
    gray.write_bytes(b"P7\nWIDTH 4\nHEIGHT 4\nDEPTH 1\nMAXVAL 255\nTUPLTYPE GRAYSCALE\nENDHDR\n" + bytes(16))
    r = _cli("1.0", str(gray))
    assert r.returncode == 1 and "only RGB and RGBA images are supported" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("fmt,ch,alpha", [("bmp", 3, 0), ("bmp", 4, 1), ("ppm", 3, 0), ("pam", 4, 0), ("pam", 4, 1)])
def test_cli_blur_equals_library(gb, tmp_path, fmt, ch, alpha):
    rows, cols, sigma = 97, 131, 3.0  # 131*3 bytes per BMP row: padded to a multiple of four
    img = go.synth_u8((rows, cols, ch), 7 + ch)
    src = tmp_path / f"in.{fmt}"
    if fmt == "bmp":
        _write_bmp(src, img)
    elif fmt == "ppm":
        src.write_bytes(b"P6\n# made by the test\n%d %d\n255\n" % (cols, rows) + img.tobytes())
    else:
        src.write_bytes(b"P7\nWIDTH %d\nHEIGHT %d\nDEPTH %d\nMAXVAL 255\nTUPLTYPE RGB_ALPHA\nENDHDR\n" % (cols, rows, ch)
                        + img.tobytes())
    r = _cli(str(sigma), str(src), str(alpha))
    assert r.returncode == 0, r.stderr
    assert f"Source image: {src} {cols}x{rows} ({ch})" in r.stdout
    out_path = tmp_path / f"in_pffft.{fmt}"
    assert out_path.exists()
    if fmt == "bmp":
        got = _read_bmp(out_path)
    else:
        raw = out_path.read_bytes()
        got = np.frombuffer(raw[-rows * cols * ch:], dtype=np.uint8).reshape(rows, cols, ch)
    want = gb.gaussianblur(img.copy(), sigma, bool(alpha))
    assert np.array_equal(got, want)
    if ch == 4 and not alpha:
        assert np.array_equal(got[..., 3], img[..., 3])
