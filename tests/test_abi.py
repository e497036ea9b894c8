"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every symbol the
headers declare; the restated boundary structs keep the reference layout.  No compute call is made here."""
import ctypes as C
import os
import re
import subprocess

import pytest

from lighthouse2_b200 import core

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    if not os.path.exists(core.LIB_PATH):
        core.build_library()
    return core.LIB_PATH


def test_library_exports_every_declared_symbol(lib_path):
    hdr = open(os.path.join(ROOT, "include", "rendercore_b200.h")).read()
    declared = set(re.findall(r"B200_API\s+[\w\s\*:]+?\b(b200_\w+|CreateCore|DestroyCore)\s*\(", hdr))
    assert declared, "header parse failed"
    assert declared == set(core.C_ABI_SYMBOLS), declared ^ set(core.C_ABI_SYMBOLS)
    lib = C.CDLL(lib_path)
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in include/rendercore_b200.h but not exported"


def test_no_oracle_or_cpu_fallback_in_product():
    """The product must not reference the oracle (it is test infrastructure only)."""
    pkg = os.path.join(ROOT, "lighthouse2_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".h", ".cuh")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.lower() or f == "__init__.py", f"{f} mentions the oracle"
                assert "liboracle" not in text


def test_create_fails_loudly_without_gpu(lib_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(core.B200Error):
        core.RenderCoreB200(0)


def test_stats_struct_matches_header(lib_path):
    hdr = open(os.path.join(ROOT, "include", "rendercore_b200.h")).read()
    body = hdr[hdr.index("typedef struct b200_stats {"):hdr.index("} b200_stats;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.replace("typedef struct b200_stats {", "").strip()
        if not decl:
            continue
        names += [n.strip() for n in re.sub(r"^(uint32_t|int32_t|float)\s+", "", decl).split(",")]
    assert names == [n for n, _ in core.Stats._fields_]
    assert C.sizeof(core.Stats) == 4 * len(names)
    assert C.sizeof(core.ViewPyramid) == 68


def test_abi_layout_compiles(tmp_path):
    """include/lh2_abi.h pins sizeof/offsetof of every boundary type with static_asserts; when the
    reference headers are present the same numbers are checked against them too."""
    src = tmp_path / "abi.cpp"
    src.write_text('#include "lh2_abi.h"\nint main(){ return sizeof(lh2abi::CoreTri) == 176 ? 0 : 1; }\n')
    exe = tmp_path / "abi"
    subprocess.run(["g++", "-std=c++17", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    assert subprocess.run([str(exe)]).returncode == 0
    ref = "/root/reference/lib"
    if os.path.isdir(ref):
        chk = os.path.join(ROOT, "tests", "cpp", "abi_vs_reference.cpp")
        inc = [f"-I{ROOT}/oracle/shim", f"-I{ref}/RenderCore_Bart", f"-I{ref}/platform", f"-I{ref}/RenderSystem", f"-I{ref}/GLFW/include",
               f"-I{ref}/glad/include", f"-I{ref}/half2.1.0", f"-I{ref}/zlib", f"-I{ROOT}/include"]
        exe2 = tmp_path / "abi_ref"
        subprocess.run(["g++", "-std=c++17", "-fpermissive", "-w", "-DGLFW_INCLUDE_NONE", "-include", "cfloat", "-include", "cstring",
                        *inc, chk, "-o", str(exe2)], check=True)
        assert subprocess.run([str(exe2)]).returncode == 0
