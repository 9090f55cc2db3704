"""In-tree build of libotedama_b200.so (nvcc, sm_100a only).

The library is the product: there is no CPU fallback, and using ``otedama_b200``
on a box where the .so is missing raises instead of degrading.  Every scan-kernel
variant is its own translation unit (csrc/scan_variant.cu with -DOTD_VIDX=i), so
the objects compile in parallel; objects are cached under csrc/_obj by content hash.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import re
import subprocess
import sys

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
CSRC = os.path.join(_PKG, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB_PATH = os.path.join(_PKG, "libotedama_b200.so")
HEADERS = [os.path.join(CSRC, f) for f in ("sha256_core.cuh", "scan_core.cuh", "scan_kernel.cuh", "job_kernels.cuh",
                                           "intpeak.cuh", "scan_variants.h", "scan_vs_core.cuh", "scan_vs_kernel.cuh")] + \
    [os.path.join(_ROOT, "include", "otedama_b200.h")]
MAIN_SRC = os.path.join(CSRC, "otedama_b200.cu")
VARIANT_SRC = os.path.join(CSRC, "scan_variant.cu")
VS_VARIANT_SRC = os.path.join(CSRC, "scan_vs_variant.cu")

# Performance studies only: extra nvcc flags for the version-shared variant objects (e.g. "-Xptxas -regUsageLevel=4")
# and an alternative output path, so that differently compiled libraries can sit side by side (OTD_LIB selects one).
VS_EXTRA_FLAGS = os.environ.get("OTD_VS_EXTRA_FLAGS", "").split()
LIB_OUT = os.environ.get("OTD_LIB_OUT", "")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def variant_indices() -> list:
    txt = open(os.path.join(CSRC, "scan_variants.h")).read()
    return [int(m) for m in re.findall(r"^\s*X\((\d+),", txt, flags=re.M)]


def vs_variant_indices() -> list:
    txt = open(os.path.join(CSRC, "scan_variants.h")).read()
    return [int(m) for m in re.findall(r"^\s*Y\((\d+),", txt, flags=re.M)]


def _digest(paths, extra="") -> str:
    h = hashlib.sha256(extra.encode())
    for p in paths:
        h.update(open(p, "rb").read())
    return h.hexdigest()[:16]


def _compile(src, out, defines, log):
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + defines + ["-c", "-o", out, src]
    if src == VS_VARIANT_SRC:
        cmd[1:1] = VS_EXTRA_FLAGS
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log.append(" ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + proc.stdout + proc.stderr)
    return out


STAMP_PATH = LIB_PATH + ".srchash"


def source_hash() -> str:
    return _digest(HEADERS + [MAIN_SRC, VARIANT_SRC, VS_VARIANT_SRC], " ".join(NVCC_FLAGS))


def needs_build() -> bool:
    """Content based (not mtime based): snapshots and checkouts shuffle mtimes."""
    if not os.path.exists(LIB_PATH) or not os.path.exists(STAMP_PATH):
        return True
    return open(STAMP_PATH).read().strip() != source_hash()


def build(force: bool = False, verbose: bool = False, jobs: int = 0) -> str:
    if LIB_OUT:
        force = True      # a side-by-side experimental library: always (re)link, never touch the stamp
    if not force and not needs_build():
        return LIB_PATH
    os.makedirs(OBJ, exist_ok=True)
    log: list = []
    tasks = []
    key_main = _digest(HEADERS + [MAIN_SRC], " ".join(NVCC_FLAGS))
    tasks.append((MAIN_SRC, os.path.join(OBJ, f"main_{key_main}.o"), []))
    key_var = _digest(HEADERS[:3] + [HEADERS[5], VARIANT_SRC], " ".join(NVCC_FLAGS))
    for i in variant_indices():
        tasks.append((VARIANT_SRC, os.path.join(OBJ, f"variant{i}_{key_var}.o"), [f"-DOTD_VIDX={i}"]))
    key_vs = _digest(HEADERS[:3] + HEADERS[5:8] + [VS_VARIANT_SRC], " ".join(NVCC_FLAGS + VS_EXTRA_FLAGS))
    for i in vs_variant_indices():
        tasks.append((VS_VARIANT_SRC, os.path.join(OBJ, f"vsvariant{i}_{key_vs}.o"), [f"-DOTD_VSIDX={i}"]))
    todo = [t for t in tasks if (force and not LIB_OUT) or not os.path.exists(t[1])]
    jobs = jobs or min(len(todo) or 1, os.cpu_count() or 1)
    with cf.ThreadPoolExecutor(max_workers=jobs) as ex:
        futs = [ex.submit(_compile, src, out, defs, log) for src, out, defs in todo]
        for f in futs:
            f.result()
    keep = {t[1] for t in tasks}
    for f in os.listdir(OBJ):  # drop stale objects
        p = os.path.join(OBJ, f)
        if p not in keep and not LIB_OUT and not VS_EXTRA_FLAGS:
            os.remove(p)
    nvcc = os.environ.get("NVCC", "nvcc")
    out_path = LIB_OUT or LIB_PATH
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", out_path] + [t[1] for t in tasks] + ["-ldl"]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log.append(" ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    with open(os.path.join(_PKG, "build.log"), "w") as fh:
        fh.write("\n".join(log))
    if proc.returncode != 0:
        raise RuntimeError("link failed:\n" + proc.stdout + proc.stderr)
    if not LIB_OUT:
        with open(STAMP_PATH, "w") as fh:
            fh.write(source_hash() + "\n")
    if verbose:
        print("\n".join(log))
    return out_path


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
