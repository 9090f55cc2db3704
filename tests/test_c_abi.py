"""include/otedama_b200.h consumed from plain C (what cgo sees): the header must
compile as C99 with -Wall -Werror -pedantic, and a C program linked against the
library must get the same answers as the Python harness.  Without a GPU the
compute entry points must refuse; with one (gpu-marked test) they must work."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "_cabi", "abi_smoke.c")
EXE = os.path.join(ROOT, "tests", "_cabi", "abi_smoke")


@pytest.fixture(scope="module")
def exe(product):
    libdir = os.path.join(ROOT, "otedama_b200")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                           "-o", EXE, SRC, "-L", libdir, "-lotedama_b200", f"-Wl,-rpath,{libdir}"])
    return EXE


def test_header_is_valid_c99():
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-fsyntax-only", "-x", "c",
                           os.path.join(ROOT, "include", "otedama_b200.h")])


def test_c_program_host_entry_points(exe):
    from otedama_b200 import capi
    if capi.lib().otd_device_count() > 0:
        pytest.skip("a CUDA device is present (covered by the gpu-marked test)")
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.startswith("OK (no device"), out.stdout + out.stderr


@pytest.mark.gpu
def test_c_program_scan_and_worker_on_the_gpu(exe):
    out = subprocess.run([exe, "gpu"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and out.stdout.startswith("OK (device present"), out.stdout + out.stderr
