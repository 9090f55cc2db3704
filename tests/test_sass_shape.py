"""Guards on the compiled default scan kernel (sm_100a SASS, read with cuobjdump — no GPU needed).

The kernel's speed depends on code-generation choices that a small source change can flip silently
(DESIGN.md §3.2): the adds must be IMADs whose multiplier 1 sits in a UNIFORM register
(`IMAD R, R, UR, R`, two vector-register reads; with the 1 in a vector register the same kernel is 14 % slower),
no spills, 64 registers at most (4 blocks of 256 threads per SM), and the unrolled body has to stay below the
instruction-cache cliff."""
import glob
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "otedama_b200", "csrc", "_obj")


@pytest.fixture(scope="module")
def sass(product):
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    objs = glob.glob(os.path.join(OBJ, "variant0_*.o"))
    assert len(objs) == 1, objs
    txt = subprocess.run(["cuobjdump", "-sass", "-res-usage", objs[0]], capture_output=True, text=True, check=True).stdout
    body = txt[txt.index("Function : _ZN3otd11scan_kernel"):]
    return txt, body


def test_default_kernel_resources(sass):
    txt, _ = sass
    m = re.search(r"Function _ZN3otd11scan_kernel\S*:\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", txt)
    assert m, txt[:400]
    reg, stack, shared, local = map(int, m.groups())
    assert reg <= 64 and shared == 0 and local == 0
    assert stack <= 128          # only the rare path's digest array (confirm_candidate takes a pointer)


def test_default_kernel_instruction_shape(sass):
    _, body = sass
    ops = re.findall(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)\s*([^;]*);", body, flags=re.M)
    n = len(ops)
    assert 3000 < n < 3450, n       # whole function incl. the rare path; the hot loop (2851 of them, 46 KB) must
                                    # stay under the ~52 KB instruction-cache cliff
    imad = [args for op, args in ops if op == "IMAD"]
    uniform_one = [a for a in imad if re.match(r"R\d+, R\d+(\.reuse)?, UR\d+, R\d+", a)]
    vector_mul = [a for a in imad if re.match(r"R\d+, R\d+(\.reuse)?, R\d+(\.reuse)?, R\d+", a)]
    assert len(uniform_one) > 800, (len(uniform_one), len(vector_mul))
    assert len(vector_mul) < 10, len(vector_mul)
    count = lambda name: sum(1 for op, _ in ops if op.split(".")[0] == name)   # noqa: E731
    assert count("SHF") >= 1176 and count("LOP3") >= 628      # nothing of the algorithm went missing
    assert count("SHF") + count("LOP3") < 1804 * 1.08         # ... and nothing much was added to the ALU pipe
    assert count("IADD3") + count("VIADD") + count("IADD") < 160   # the adds live on the FMA pipe
    # no tensor core, TMA, shared memory or barrier; the only global load is the stop word, once per block
    assert not any(op.startswith(("HMMA", "UTCHMMA", "UTMALDG", "LDS", "STS", "BAR")) for op, _ in ops)
    assert sum(1 for op, _ in ops if op.startswith("LDG")) == 1


# ---- the version-shared scan kernel (default variant) ----------------------------------------------------------------
@pytest.fixture(scope="module")
def vs_sass(product):
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    objs = glob.glob(os.path.join(OBJ, "vsvariant0_*.o"))
    assert len(objs) == 1, objs
    txt = subprocess.run(["cuobjdump", "-sass", "-res-usage", objs[0]], capture_output=True, text=True, check=True).stdout
    body = txt[txt.index("Function : _ZN3otd14scan_vs_kernel"):]
    return txt, body


def test_version_shared_kernel_resources(vs_sass):
    """No spills; registers at most 64, so that four 256-thread blocks stay resident per SM although the shipped
    launch bounds (256, 3) would allow 85; the 46-word schedule column per thread is the only shared memory
    (4 x 47,104 B fit the SM's 228 KB)."""
    txt, _ = vs_sass
    m = re.search(r"Function _ZN3otd14scan_vs_kernel\S*:\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", txt)
    assert m, txt[:400]
    reg, stack, shared, local = map(int, m.groups())
    assert reg <= 64 and local == 0 and stack <= 128
    assert shared - 46 * 256 * 4 in (0, 1024)             # cuobjdump adds the 1 KB the driver reserves per block


def test_version_shared_kernel_instruction_shape(vs_sass):
    """The schedule is written once (46 STS) and read once per version instance (46 LDS) — not hoisted into
    registers across the version loop, which would spill —, the per-version body carries no schedule arithmetic
    (SHF + LOP3 within 3 % of 1500 + 304 for the per-nonce part), and the adds sit on the FMA pipe as IMADs with a
    uniform-register 1."""
    _, body = vs_sass
    ops = re.findall(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)\s*([^;]*);", body, flags=re.M)
    count = lambda name: sum(1 for op, _ in ops if op.split(".")[0] == name)   # noqa: E731
    assert count("STS") == 46 and count("LDS") == 46
    assert count("LDL") + count("STL") < 64                   # only the rare path's digest array
    assert 1176 <= count("SHF") <= 1176 + 80                  # 948 per version + 228 per nonce (+ the looped rare path)
    assert 628 <= count("LOP3") <= 628 * 1.10
    imad = [args for op, args in ops if op == "IMAD"]
    uniform_one = [a for a in imad if re.match(r"R\d+, R\d+(\.reuse)?, UR\d+, R\d+", a)]
    vector_mul = [a for a in imad if re.match(r"R\d+, R\d+(\.reuse)?, R\d+(\.reuse)?, R\d+", a)]
    # (the ~100 vector-register multipliers are the once-per-nonce schedule expansion, 1/32 of the dynamic stream)
    assert len(uniform_one) > 700 and len(vector_mul) < 120, (len(uniform_one), len(vector_mul))
    assert not any(op.startswith(("HMMA", "UTCHMMA", "UTMALDG", "BAR")) for op, _ in ops)
