"""Generates tests/golden/nested_reference.json.  Run in the build container (needs /root/reference):

    python tests/golden/make_nested_golden.py

Digests of highOrderFlx(1:nVertLevels, :) computed by THE REFERENCE'S OWN code for the nested_loops mini-app -- its C++ forms
(nested_yakl_c.cpp:184-306) compiled by g++, oracle/ref_yakl_py.py -- on seeded problems built as `program nested` builds
them (codesign-kernels_b200/nested.py::NestedProblem), at the namelist sizes (nested.nml:1-7) and at the small shapes of
tests/test_zz_nested.py.  The translated Fortran forms (oracle/ref_nested_f90_py.py) must give the same digests when the
script runs; tests/test_zz_nested.py requires the hand-written oracle to hash to them -- a check that runs anywhere.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

CASES = {  # name: (nEdges, nCells, nVertLevels, nAdv, packn, seed)
    "namelist": (25600, 2800, 100, 10, 8, 1),  # nested.nml:1-7 with F90_PACK_SIZE = 8 (Makefile:110)
    "namelist_shape_small": (640, 70, 100, 10, 1, 3),
    "odd_levels": (257, 33, 37, 7, 1, 3),
    "single_edge_single_cell": (1, 1, 5, 3, 1, 3),
}


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    from codesign_kernels_b200 import nested
    import ref_nested_f90_py as f90
    import ref_yakl_py as yref

    assert yref.available() and f90.available(), "needs /root/reference (or built oracle/_ref libraries)"
    out = {"source": "nested_loops/nested_yakl_c.cpp:184-306 compiled by g++ (oracle/_ref/libnested_yakl_ref.so); the translated Fortran forms "
                     "nested.F90:495-564, 593-630 give the same digests", "cases": {}}
    for name, (ne, nc, nl, na, pk, seed) in CASES.items():
        prob = nested.NestedProblem(nEdges=ne, nCells=nc, nVertLevels=nl, nAdv=na, nIters=1, packn=pk, seed=seed)
        r = yref.RefProblem(prob)
        o, g = r.original_loop()[:nl], r.gpu_optimized()[:nl]
        assert np.array_equal(f90.run(prob, "original")[:nl], o) and np.array_equal(f90.run(prob, "gpu_optimized")[:nl], g)
        out["cases"][name] = {"nEdges": ne, "nCells": nc, "nVertLevels": nl, "nAdv": na, "packn": pk, "seed": seed,
                              "tracer_sha256": digest(prob.tracerCur), "original_sha256": digest(o), "gpu_optimized_sha256": digest(g)}
    with open(os.path.join(ROOT, "tests", "golden", "nested_reference.json"), "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out["cases"]["namelist"]))


if __name__ == "__main__":
    main()
