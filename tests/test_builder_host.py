"""Host BVH builder (lighthouse2_b200/csrc/bvh_builder.cpp) without a GPU: tests/cpp/builder_check.cpp is compiled against the
builder's source and verifies, for several kinds of input, what the traversal kernels rely on (every primitive once, the node layout of
wide_bvh.h, child planes that contain the geometry below them) and that the tree is the same for any number of build threads.

The builder replaces BVH2::Rebuild (lib/RenderCore_Bart/bvh.cpp:57-256); its tree need not be the reference's (the reference's
traversal visits every box a ray overlaps, bvh.cpp:281-298, so any conservative hierarchy gives the same closest hit) -- which is why
the checks here are structural, and the hits are compared on the GPU (tests/test_parity_gpu.py)."""

import os
import subprocess

import numpy as np
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


def test_radix_subtree_index_rule():
    """The bookkeeping behind device builder 3 (build.cu, "SAH inside the radix tree"): in Karras' radix tree over n sorted keys the
    inner node i has the sorted range [first, last] with i one of its ends, its children are the split position g and g + 1, and the
    inner nodes of its subtree are exactly the positions of its range without the OTHER end.  That is what lets a rebuilt subtree
    live in the index places of the one it replaces.  Restated here in plain Python (radixTreeKernel's algorithm) and checked on
    random keys with duplicates."""
    rng = np.random.default_rng(17)
    for n in (2, 3, 7, 64, 257, 600):
        keys = np.sort(rng.integers(0, 1 << 12, n)).astype(np.uint64)  # few distinct values: many ties, broken by the index as in prefixLen

        def delta(i, j):
            if j < 0 or j >= n:
                return -1
            a, b = int(keys[i]), int(keys[j])
            if a != b:
                return 64 - (a ^ b).bit_length()
            return 64 + 32 - (i ^ j).bit_length()
        left, right, first, last = {}, {}, {}, {}
        for i in range(n - 1):
            d = 1 if delta(i, i + 1) - delta(i, i - 1) >= 0 else -1
            dmin = delta(i, i - d)
            lmax = 2
            while delta(i, i + lmax * d) > dmin:
                lmax *= 2
            length, t = 0, lmax // 2
            while t >= 1:
                if delta(i, i + (length + t) * d) > dmin:
                    length += t
                t //= 2
            j = i + length * d
            dnode = delta(i, j)
            s, div = 0, 2
            while True:
                t = (length + div - 1) // div
                if delta(i, i + (s + t) * d) > dnode:
                    s += t
                if t <= 1:
                    break
                div *= 2
            g = i + s * d + min(d, 0)
            first[i], last[i] = min(i, j), max(i, j)
            left[i] = ("leaf", g) if first[i] == g else ("node", g)
            right[i] = ("leaf", g + 1) if last[i] == g + 1 else ("node", g + 1)

        def inner_ids(i):
            out = {i}
            for kind, c in (left[i], right[i]):
                if kind == "node":
                    out |= inner_ids(c)
            return out
        assert first[0] == 0 and last[0] == n - 1
        for i in range(n - 1):
            assert i in (first[i], last[i])
            other = last[i] if i == first[i] else first[i]
            assert inner_ids(i) == set(range(first[i], last[i] + 1)) - {other}, (n, i)
