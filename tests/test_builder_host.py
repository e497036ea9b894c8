"""Host BVH builder (lighthouse2_b200/csrc/bvh_builder.cpp) without a GPU: tests/cpp/builder_check.cpp is compiled against the
builder's source and verifies, for several kinds of input, what the traversal kernels rely on (every primitive once, the node layout of
wide_bvh.h, child planes that contain the geometry below them) and that the tree is the same for any number of build threads.

The builder replaces BVH2::Rebuild (lib/RenderCore_Bart/bvh.cpp:57-256); its tree need not be the reference's (the reference's
traversal visits every box a ray overlaps, bvh.cpp:281-298, so any conservative hierarchy gives the same closest hit) -- which is why
the checks here are structural, and the hits are compared on the GPU (tests/test_parity_gpu.py)."""

import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "lighthouse2_b200", "csrc")


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("builder") / "builder_check")
    # the host flags of lighthouse2_b200/csrc/Makefile
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-pthread", "-I", CSRC, os.path.join(ROOT, "tests", "cpp", "builder_check.cpp"),
                    os.path.join(CSRC, "bvh_builder.cpp"), "-o", exe], check=True)
    return exe


def run(exe, case, n, threads, max_leaf=3):
    out = subprocess.run([exe, case, str(n), str(threads), str(max_leaf)], capture_output=True, text=True, timeout=300)
    fields = dict(kv.split("=", 1) for kv in out.stdout.split())
    assert out.returncode == 0 and fields["ok"] == "1", "%s %d x%d: %s %s" % (case, n, threads, out.stdout, out.stderr)
    return fields


@pytest.mark.parametrize("case,n", [("soup", 150_000), ("grid", 160_000), ("clusters", 120_000), ("same", 5_000)])
def test_tree_is_valid_and_independent_of_the_thread_count(checker, case, n):
    one = run(checker, case, n, 1)
    for threads in (2, 5, 8):
        many = run(checker, case, n, threads)
        assert many["hash"] == one["hash"] and many["nodes"] == one["nodes"], (case, threads)
    assert int(one["leaves"]) >= n / 3  # leaves hold at most three triangles


@pytest.mark.parametrize("n", [0, 1, 2, 3, 4, 7, 8, 9, 25, 1000])
def test_small_inputs(checker, n):
    for case in ("soup", "same"):
        f = run(checker, case, n, 4)
        if n <= 3:
            assert f["nodes"] == "1"


def test_single_primitive_leaves_for_the_top_level(checker):
    # the top level over instance boxes is built with one primitive per leaf (rendercore.cpp)
    f = run(checker, "clusters", 2000, 1, max_leaf=1)
    assert f["leaves"] == "2000"
