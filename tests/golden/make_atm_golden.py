"""Generates tests/golden/atm_translated.json.  Run in the build container (needs /root/reference):

    python tests/golden/make_atm_golden.py

Golden digests of the hot path computed by THE REFERENCE'S OWN three modules (atmosphere/biharmonic_wk_kernel.F90: CPU
F90:100-202, GPU1 F90:209-362, GPU2 F90:369-536), translated mechanically from the Fortran to C by oracle/f90_to_c.py and
compiled here (oracle/ref_atm_py.py; this image has no Fortran compiler).  Inputs: the reference's own LCG (F90:48-58, 77-91)
at its default size (F90:11-17) and at two ragged sizes, plus the two relative L2 norms the reference publishes
(atmosphere/results/atm.*), recomputed from the translated modules.  tests/test_zzy_atm_ref.py requires the hand-written oracle
to hash to the same digests -- a check that runs anywhere, with or without oracle/_ref.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

CASES = [(16, 72, 40), (7, 33, 1), (3, 16, 2)]  # (nelemd, nlev, qsize); nlev = 33 is no multiple of kchunk = 8


def digest(a):
    return hashlib.sha256(np.asfortranarray(a).tobytes(order="F")).hexdigest()


def main():
    import oracle_py as orc
    import ref_atm_py as ref

    orc.build()
    assert ref.available(), "needs /root/reference (or a built oracle/_ref/libatm_f90ref.so)"
    out = {"source": "atmosphere/biharmonic_wk_kernel.F90 translated by oracle/f90_to_c.py (unit mode), gcc -O2 -ffp-contract=off",
           "inputs": "reference LCG, F90:48-58 + F90:77-91 (old = 11, 1301*old+97 mod 131072)", "cases": []}
    for nelemd, nlev, qsize in CASES:
        dvv, elem, q = orc.initialize_data(nelemd, nlev, qsize)
        cpu = ref.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
        g1 = ref.biharmonic_wk_scalar_gpu1(elem, q.copy(order="F"), dvv)
        g2 = ref.biharmonic_wk_scalar_gpu2(elem, q.copy(order="F"), dvv)
        out["cases"].append({"nelemd": nelemd, "nlev": nlev, "qsize": qsize, "input_sha256": digest(q),
                             "cpu_sha256": digest(cpu), "gpu1_sha256": digest(g1), "gpu2_sha256": digest(g2),
                             "gpu1_vs_cpu_rel_l2": "%.16E" % orc.compute_l2norm(g1, cpu), "gpu2_vs_cpu_rel_l2": "%.16E" % orc.compute_l2norm(g2, cpu)})
    # a nets:nete sub-range through the CPU signature (F90:186-191)
    dvv, elem, q = orc.initialize_data(9, 12, 5)
    sub = ref.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, 3, 7)
    out["nets_nete_case"] = {"nelemd": 9, "nlev": 12, "qsize": 5, "nets": 3, "nete": 7, "cpu_sha256": digest(sub)}
    path = os.path.join(ROOT, "tests", "golden", "atm_translated.json")
    with open(path, "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out["cases"][0]))


if __name__ == "__main__":
    main()
