"""Mint golden vectors from the compiled, UNMODIFIED reference (oracle/_ref/libgblur_ref.so, built
by oracle/Makefile from /root/reference). Run in the dev container only:

    python tests/golden/make_golden.py

The reference's own tests hold no numeric golden vectors for the blur path, so
these are the pin: inputs are regenerated from seeds (oracle.gblur_oracle.synth_u8 or the named
adversarial patterns below), outputs are stored in golden_small.npz (full arrays, small cases) and
golden_digests.json (sha256 of larger outputs + kernel-plan facts + spectra samples).
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import gblur_oracle as go  # noqa: E402
from oracle import ref_lib  # noqa: E402

# (name, rows, cols, channels, sigma, apply_to_alpha, pattern)
SMALL_CASES = [
    ("rgb_3x3_s3", 3, 3, 3, 3.0, False, "synth"),
    ("rgb_4x4_s2", 4, 4, 3, 2.0, False, "synth"),
    ("rgb_17x33_s1", 17, 33, 3, 1.0, False, "synth"),
    ("rgba_17x33_s1_alpha", 17, 33, 4, 1.0, True, "synth"),
    ("rgba_17x33_s1_noalpha", 17, 33, 4, 1.0, False, "synth"),
    ("rgb_40x56_s0p3", 40, 56, 3, 0.3, False, "synth"),
    ("rgba_64x48_s3_altcols", 64, 48, 4, 3.0, True, "altcols"),
    ("rgb_64x48_s3_altrows", 64, 48, 3, 3.0, False, "altrows"),
    ("rgba_64x48_s3_checker", 64, 48, 4, 3.0, True, "checker"),
    ("rgb_64x48_s5_const255", 64, 48, 3, 5.0, False, "const255"),
    ("rgb_64x48_s5_impulse", 64, 48, 3, 5.0, False, "impulse"),
    ("rgba_96x80_s5_hramp", 96, 80, 4, 5.0, True, "hramp"),
    ("rgb_96x80_s5_vramp", 96, 80, 3, 5.0, False, "vramp"),
    ("rgb_100x75_s8", 100, 75, 3, 8.0, False, "synth"),
]
DIGEST_CASES = [
    ("rgba_270x480_s5_alpha", 270, 480, 4, 5.0, True, "synth"),
    ("rgb_270x480_s5", 270, 480, 3, 5.0, False, "synth"),
    ("rgba_270x480_s20_altcols", 270, 480, 4, 20.0, True, "altcols"),
    ("rgb_1080x1920_s5", 1080, 1920, 3, 5.0, False, "synth"),
    ("rgb_1080x1920_s10", 1080, 1920, 3, 10.0, False, "synth"),
    ("rgba_540x960_s50_alpha", 540, 960, 4, 50.0, True, "synth"),
]
PLAN_CASES = [(1080, 1920, 5.0), (4320, 7680, 20.0), (1080, 1920, 10.0), (32768, 32768, 50.0),
              (32768, 65536, 150.0), (10240, 10240, 5.0), (3, 3, 3.0), (4, 4, 2.0), (17, 33, 1.0), (270, 480, 5.0)]


def make_input(rows, cols, ch, pattern, seed=0x1234):
    if pattern == "synth":
        return go.synth_u8((rows, cols, ch), seed)
    img = np.zeros((rows, cols, ch), dtype=np.uint8)
    r = np.arange(rows)[:, None, None]
    c = np.arange(cols)[None, :, None]
    if pattern == "altcols":
        img[:] = np.where(c % 2 == 0, 0, 255)
    elif pattern == "altrows":
        img[:] = np.where(r % 2 == 0, 0, 255)
    elif pattern == "checker":
        img[:] = np.where((r + c) % 2 == 0, 0, 255)
    elif pattern == "const255":
        img[:] = 255
    elif pattern == "impulse":
        img[rows // 2, cols // 2, :] = 255
    elif pattern == "hramp":
        img[:] = (c * 255 // max(cols - 1, 1)).astype(np.uint8)
    elif pattern == "vramp":
        img[:] = (r * 255 // max(rows - 1, 1)).astype(np.uint8)
    else:
        raise ValueError(pattern)
    return img


def main():
    assert ref_lib.build(), "reference oracle could not be built"
    small = {}
    for name, rows, cols, ch, sigma, alpha, pat in SMALL_CASES:
        img = make_input(rows, cols, ch, pat)
        small[name] = ref_lib.gaussianblur(img, sigma, alpha)
        small[name + "__rowpass_ch0"] = ref_lib.rowpass_plane(img, 0, sigma)
    # 1-channel route (process_channel_tiles x2), SURVEY.md 8c
    plane = go.synth_u8((48, 64), 0x1234 + 3)
    small["gray_48x64_s4"] = ref_lib.blur_plane_u8(plane, 4.0)
    np.savez_compressed(os.path.join(HERE, "golden_small.npz"), **small)

    digests = {"outputs": {}, "plans": {}, "spectra": {}}
    for name, rows, cols, ch, sigma, alpha, pat in DIGEST_CASES:
        out = ref_lib.gaussianblur(make_input(rows, cols, ch, pat), sigma, alpha)
        digests["outputs"][name] = {
            "sha256": hashlib.sha256(out.tobytes()).hexdigest(),
            "sum": int(out.astype(np.uint64).sum()),
            "sample": out.reshape(-1)[:: max(1, out.size // 64)][:64].tolist(),
        }
    for rows, cols, sigma in PLAN_CASES:
        if rows * cols <= 4320 * 7680:
            k = ref_lib.kernel_dft(rows, cols, sigma)
            key = f"{rows}x{cols}_s{sigma}"
            digests["plans"][key] = {x: int(k[x]) for x in ("pad", "n_rowpass", "n_colpass", "tz_rows", "tz_cols")}
            kc = k["kerf_col"]
            idx = sorted(set([0, 1, 2, 3, 4, 5, 10, 11, 100, 101, len(kc) // 2, len(kc) - 2, len(kc) - 1]) & set(range(len(kc))))
            digests["spectra"][key] = {"idx": idx, "kerf_col": [float(kc[i]) for i in idx]}
        else:  # only the cheap facts for the gigapixel geometries (host arithmetic, no image)
            key = f"{rows}x{cols}_s{sigma}"
            k = ref_lib.kernel_dft(rows, cols, sigma)
            digests["plans"][key] = {x: int(k[x]) for x in ("pad", "n_rowpass", "n_colpass", "tz_rows", "tz_cols")}
    with open(os.path.join(HERE, "golden_digests.json"), "w") as fh:
        json.dump(digests, fh, indent=1, sort_keys=True)
    print("wrote", len(small), "small arrays and", len(digests["outputs"]), "digests")


if __name__ == "__main__":
    main()
