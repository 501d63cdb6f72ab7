"""In-tree build of the sm_100a library (no JIT cache, no install step).

    python -m libgaussianblur_b200.build [--force] [--verbose]

Outputs (git-ignored, but they travel to the GPU box with the snapshot):
    libgaussianblur_b200/lib/libgblur_b200.so          C ABI (include/gblur_capi.h) + C++ API
    libgaussianblur_b200/lib/gaussianblur.<abi>.so     pybind11 module mirroring the reference's
    libgaussianblur_b200/lib/gaussianblur              desktop command line (examples/desktop/main.cpp)
Every CUDA translation unit is compiled with
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo
static_impl.cu is compiled once per compile-time FFT shape (static_shapes.h) so the shapes
build in parallel.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import os
import re
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_build")
LIB = os.path.join(HERE, "lib")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC", "-Xptxas", "-v",
          "-I", os.path.join(ROOT, "include"), "-I", CSRC]


def _static_ids():
    txt = open(os.path.join(CSRC, "static_shapes.h")).read()
    return [int(m) for m in re.findall(r"^\s*X\((\d+),", txt, flags=re.M)]


def _sources_digest(extra: str = "") -> str:
    h = hashlib.sha256(extra.encode())
    for d in (CSRC, os.path.join(ROOT, "include"), os.path.join(ROOT, "include", "gaussianblur"),
              os.path.join(ROOT, "examples", "desktop")):
        if not os.path.isdir(d):
            continue
        for f in sorted(os.listdir(d)):
            p = os.path.join(d, f)
            if os.path.isfile(p):
                h.update(f.encode())
                h.update(open(p, "rb").read())
    return h.hexdigest()


def _deps_digest(src: str, cmd) -> str:
    """sha256 of the command line, the source and every quoted/bracketed header it reaches inside
    csrc/ and include/ (so a change to one kernel file rebuilds only the objects that see it)."""
    roots = [CSRC, os.path.join(ROOT, "include")]
    seen, todo = set(), [src]
    h = hashlib.sha256(" ".join(cmd).encode())
    while todo:
        f = todo.pop()
        if f in seen or not os.path.isfile(f):
            continue
        seen.add(f)
        txt = open(f, "rb").read()
        h.update(os.path.basename(f).encode())
        h.update(txt)
        for inc in re.findall(rb'^\s*#\s*include\s*[<"]([^>"]+)[>"]', txt, flags=re.M):
            name = inc.decode()
            for d in [os.path.dirname(f)] + roots:
                todo.append(os.path.normpath(os.path.join(d, name)))
    return h.hexdigest()


def _run_cached(cmd, obj, src):
    """compile `src` to `obj` unless an object built from identical inputs is already there"""
    tag = obj + ".dep"
    digest = _deps_digest(src, cmd)
    if os.path.exists(obj) and os.path.exists(tag) and open(tag).read().strip() == digest:
        return None
    r = _run(cmd, obj + ".log")
    with open(tag, "w") as fh:
        fh.write(digest)
    return r


def _run(cmd, log):
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(log, "w") as fh:
        fh.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"build step failed ({log}):\n{r.stderr[-4000:]}")
    return r


def lib_path() -> str:
    return os.path.join(LIB, "libgblur_b200.so")


def pymod_path() -> str:
    return os.path.join(LIB, "gaussianblur" + sysconfig.get_config_var("EXT_SUFFIX"))


def cli_path() -> str:
    return os.path.join(LIB, "gaussianblur")


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIB, exist_ok=True)
    stamp = os.path.join(LIB, "build.stamp")
    digest = _sources_digest()
    if force:
        for f in os.listdir(OBJ):
            if f.endswith(".dep"):
                os.remove(os.path.join(OBJ, f))
    if (not force and os.path.exists(lib_path()) and os.path.exists(pymod_path()) and os.path.exists(cli_path())
            and os.path.exists(stamp)
            and open(stamp).read().strip() == digest):
        return lib_path()
    if not os.path.exists(NVCC):
        raise RuntimeError(f"nvcc not found at {NVCC}; cannot build the sm_100a library")

    jobs = []  # (object, cmd, source)
    for name in ("kernels.cu", "generic_row.cu", "generic_col.cu", "cluster_kernels.cu", "ols_kernels.cu", "noalloc.cu"):
        o = os.path.join(OBJ, name.replace(".cu", ".o"))
        jobs.append((o, [NVCC, *ARCH, *COMMON, "-c", os.path.join(CSRC, name), "-o", o], os.path.join(CSRC, name)))
    for sid in _static_ids():
        o = os.path.join(OBJ, f"static_{sid}.o")
        jobs.append((o, [NVCC, *ARCH, *COMMON, f"-DGBLUR_TU_SHAPE={sid}", "-c",
                         os.path.join(CSRC, "static_impl.cu"), "-o", o], os.path.join(CSRC, "static_impl.cu")))
    for name in ("capi.cpp", "gaussianblur_api.cpp", "multi_gpu.cpp"):
        o = os.path.join(OBJ, name.replace(".cpp", ".o"))
        jobs.append((o, [NVCC, *COMMON, "-c", os.path.join(CSRC, name), "-o", o],
                     os.path.join(CSRC, name)))  # host-only TU

    # development aid: GBLUR_BUILD_ONLY=ols_kernels,kernels recompiles just those objects and links the others as
    # they are (stale but ABI-compatible); the stamp is not written, so the next plain build checks everything again
    only = [t for t in os.environ.get("GBLUR_BUILD_ONLY", "").split(",") if t]
    if only:
        missing = [o for o, _, _ in jobs if not os.path.exists(o) and os.path.basename(o)[:-2] not in only]
        if missing:
            raise RuntimeError("GBLUR_BUILD_ONLY needs a complete earlier build; missing " + ", ".join(missing))
        all_jobs = jobs
        jobs = [j for j in jobs if os.path.basename(j[0])[:-2] in only]
    workers = max(1, min(len(jobs), os.cpu_count() or 1))
    with cf.ThreadPoolExecutor(workers) as ex:
        futs = {ex.submit(_run_cached, cmd, o, src): o for o, cmd, src in jobs}
        for f in cf.as_completed(futs):
            f.result()
            if verbose:
                print("built", os.path.basename(futs[f]), flush=True)

    if only:
        jobs = all_jobs
    objs = [o for o, _, _ in jobs]
    _run([NVCC, *ARCH, "-shared", "-Xcompiler", "-fPIC", "-cudart", "static", *objs, "-o", lib_path()],
         os.path.join(OBJ, "link.log"))

    # pybind11 module with the reference's Python surface (python/bindings/gaussianblur_py.cpp)
    import pybind11

    py_inc = sysconfig.get_paths()["include"]
    cxx = os.environ.get("CXX", "g++")
    _run([cxx, "-std=c++17", "-O2", "-fPIC", "-shared", "-fvisibility=hidden",
          "-I", os.path.join(ROOT, "include"), "-I", pybind11.get_include(), "-I", py_inc,
          os.path.join(CSRC, "pybind_module.cpp"), "-L", LIB, "-lgblur_b200", "-Wl,-rpath,$ORIGIN",
          "-o", pymod_path()], os.path.join(OBJ, "pybind.log"))
    # desktop command line with the reference's usage (examples/desktop/main.cpp)
    _run([cxx, "-std=c++17", "-O2", "-I", os.path.join(ROOT, "include"),
          os.path.join(ROOT, "examples", "desktop", "main.cpp"), "-L", LIB, "-lgblur_b200", "-Wl,-rpath,$ORIGIN",
          "-o", cli_path()], os.path.join(OBJ, "cli.log"))
    with open(stamp, "w") as fh:
        fh.write("partial" if only else digest)
    return lib_path()


def ptxas_summary():
    """(kernel, registers, spill bytes) from the last build's logs."""
    rows = []
    for f in sorted(os.listdir(OBJ)):
        if not f.endswith(".o.log"):
            continue
        txt = open(os.path.join(OBJ, f)).read()
        for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'.*?(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads.*?Used (\d+) registers", txt, flags=re.S):
            rows.append((f, m.group(1), int(m.group(5)), int(m.group(3)), int(m.group(4))))
    return rows


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
    if a.verbose:
        for r in ptxas_summary():
            print(r)
