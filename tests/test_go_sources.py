"""The Go files under go/ cannot be compiled here (no Go toolchain in the image), so the mistakes a compiler would
catch at the cgo boundary are checked textually instead: every C symbol, constant, type and struct field the Go
code names must exist in include/otedama_b200.h, braces must balance, and the cgo preamble must include what the
code calls (C.memcpy needs <string.h> — a defect of the round-1 shim)."""
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = open(os.path.join(ROOT, "include", "otedama_b200.h")).read()
GO_FILES = sorted(glob.glob(os.path.join(ROOT, "go", "**", "*.go"), recursive=True))


def _strip_comments(src):
    src = re.sub(r"/\*.*?\*/", lambda m: m.group(0) if "#include" in m.group(0) else "", src, flags=re.S)
    return "\n".join(l.split("//")[0] if not l.lstrip().startswith("//go:build") else l for l in src.splitlines())


def _struct_fields(name):
    m = re.search(r"typedef struct(?: \w+)? \{(.*?)\} " + name + ";", HEADER, flags=re.S)
    assert m, name
    body = re.sub(r"/\*.*?\*/", "", m.group(1), flags=re.S)
    fields = set()
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):
            fm = re.search(r"(\w+)\s*(\[[^\]]*\])?\s*$", part.strip())
            if fm:
                fields.add(fm.group(1))
    return fields


def test_go_files_exist_and_balance():
    names = {os.path.relpath(p, os.path.join(ROOT, "go")) for p in GO_FILES}
    for want in ("internal/miner/gpu_worker.go", "internal/hal/cuda_driver.go", "internal/config/gpu_config.go",
                 "internal/engine/gpu_setup.go", "internal/engine/gpu_v1.go", "internal/engine/gpu_v2.go",
                 "internal/poolproto/stratumv1/notify_ext.go"):
        assert want in names, want
    for p in GO_FILES:
        code = _strip_comments(open(p).read())
        code = re.sub(r'"(?:\\.|[^"\\])*"|`[^`]*`', '""', code)          # string literals out of the way
        for a, b in ("{}", "()", "[]"):
            assert code.count(a) == code.count(b), (p, a, code.count(a), code.count(b))
        assert re.search(r"^package \w+", code, flags=re.M), p


def test_every_c_name_used_from_go_exists_in_the_header():
    funcs = set(re.findall(r"\b(otd_[a-z0-9_]+)\s*\(", re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)))
    types = set(re.findall(r"\}\s*(otd_\w+_t)\s*;", HEADER)) | set(re.findall(r"typedef struct \w+ (otd_\w+_t);", HEADER))
    consts = set(re.findall(r"#define (OTD_\w+)", HEADER))
    seen = 0
    for p in GO_FILES:
        src = open(p).read()
        if 'import "C"' not in src:
            continue
        code = _strip_comments(src)
        for name in re.findall(r"\bC\.(otd_\w+)", code):
            assert name in funcs or name in types, (p, name)
            seen += 1
        for name in re.findall(r"\bC\.(OTD_\w+)", code):
            assert name in consts, (p, name)
        if "C.memcpy" in code:
            assert "#include <string.h>" in src, p
        if "C.free" in code or "C.malloc" in code:
            assert "#include <stdlib.h>" in src, p
    assert seen > 15


def test_struct_fields_named_in_go_exist_in_the_header():
    worker = _strip_comments(open(os.path.join(ROOT, "go", "internal", "miner", "gpu_worker.go")).read())
    work = _struct_fields("otd_work_t") | {f"header.{f}" for f in _struct_fields("otd_header_t")}
    for f in re.findall(r"\bcw\.((?:header\.)?\w+)", worker):
        assert f in work or f == "header", f
    share = _struct_fields("otd_share_t")
    for f in re.findall(r"\bbuf\[i\]\.(\w+)", worker):
        assert f in share, f
    cfg = _struct_fields("otd_worker_config_t")
    for f in re.findall(r"\bc\.(ring_capacity|queue_capacity|launch_log2|kernel_variant|\w+_capacity)\b", worker):
        assert f in cfg, f
    stats = _struct_fields("otd_stats_t")
    m = re.search(r"func \(w \*GPUWorker\) Stats\(\) Stats \{(.*?)\n\}", worker, flags=re.S)
    for f in re.findall(r"\bs\.(\w+)", m.group(1)):
        assert f in stats, f
    driver = _strip_comments(open(os.path.join(ROOT, "go", "internal", "hal", "cuda_driver.go")).read())
    info = _struct_fields("otd_device_info_t")
    for f in re.findall(r"\binfo\.(\w+)", driver):
        assert f in info, f
