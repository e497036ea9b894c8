"""Generates the golden hit buffers under tests/golden/ from the REFERENCE ITSELF.

Run in the build container (needs /root/reference to compile oracle/_ref):
    python tests/golden/make_golden.py
Every golden t/prim comes from oracle/_ref/liboracle_ref.so = the unmodified lib/RenderCore_Bart/bvh.cpp +
common.h; the instance ids (which the reference API cannot return) come from the plain-C port after it has
been checked bit-identical on (t, prim) for the same rays.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cases  # noqa: E402
import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    oracle.build()
    assert oracle.have_ref(), "oracle/_ref/liboracle_ref.so missing: /root/reference is required to make goldens"
    todo = [cases.case_c1(), cases.case_soup(), cases.case_instanced(), cases.case_instanced(rotate=True), cases.case_axis_aligned(),
            cases.case_single_triangle()]
    for case in todo:
        O, D = case["rays"]
        ref = cases.load_oracle(case, "ref")
        port = cases.load_oracle(case, "port")
        t, prim, _ = ref.trace(O, D, nthreads=8)
        tp, primp, instp = port.trace(O, D, nthreads=8)
        assert np.array_equal(t.view(np.uint32), tp.view(np.uint32)) and np.array_equal(prim, primp), case["name"]
        sha = hashlib.sha256(t.tobytes() + prim.tobytes() + instp.tobytes()).hexdigest()
        path = os.path.join(HERE, case["name"] + ".npz")
        np.savez_compressed(path, t=t, prim=prim, inst=instp, sha256=np.frombuffer(bytes.fromhex(sha), np.uint8))
        print(f"{case['name']:>10}: {len(t)} rays, {int((prim >= 0).sum())} hits, sha256 {sha[:16]}  -> {os.path.relpath(path, ROOT)} "
              f"({os.path.getsize(path) / 1024:.0f} KiB)")


if __name__ == "__main__":
    main()
