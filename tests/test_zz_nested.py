"""nested_loops mini-app: the MPAS-Ocean high-order flux loop behind the reference's
own C interface (nested_loops/cke.hpp:7-23).

PARITY PINNED ON THE REFERENCE COMPILED HERE: oracle/_ref/libnested_yakl_ref.so is the reference's
own C++ (nested_loops/nested_yakl_c.cpp:142-306 on the vendored YAKL headers, host back end) built by
`make -C oracle ref` from the sources where they lie.  oracle/cke_oracle.c is checked against it bit
for bit (and with the reference's own criterion, errTol = 1e-10, nested.F90:267-279), and the -m gpu
tests compare the CUDA path with it directly.  The reference holds no stored outputs (inputs come from
the compiler's random_number), so "the reference run here on the same inputs" is the pin.

This file sorts last on purpose: the -m gpu tests here are the newest in the suite.
"""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


@pytest.fixture(scope="module")
def nested(bwk):
    from codesign_kernels_b200 import nested as m

    return m


@pytest.fixture(scope="module")
def cko():
    import cke_oracle_py

    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, stdout=subprocess.DEVNULL)
    return cke_oracle_py


@pytest.fixture(scope="module")
def yref():
    """The reference's own C++ (oracle/_ref).  Built here from /root/reference; on the GPU box the prebuilt .so travels."""
    import ref_yakl_py

    if not ref_yakl_py.available():
        pytest.skip("oracle/_ref/libnested_yakl_ref.so absent and /root/reference not present to build it")
    return ref_yakl_py


def numpy_reference(prob):
    """Independent restatement of nested.F90:119-157 (vectorised over edges and levels, loop over
    the contributing cells in order) -- same operation order per point, so bitwise comparable."""
    nv, ne = prob.nvldim, prob.nEdges
    ntf = prob.normalThicknessFlux
    wgt = ntf * prob.advMaskHighOrder
    sgn = np.copysign(1.0, ntf)
    out = np.zeros((nv, ne), order="F")
    lev = np.arange(1, nv + 1)[:, None]
    for i in range(prob.nAdv):
        use = i < prob.nAdvCellsForEdge
        c = np.where(use, prob.advCellsForEdge[i], 1) - 1
        coef1 = prob.advCoefs[i][None, :]
        coef3 = (prob.advCoefs3rd[i] * prob.coef3rdOrder)[None, :]
        inside = (lev >= prob.minLevelCell[c][None, :]) & (lev <= prob.maxLevelCell[c][None, :]) & use[None, :]
        inside &= lev <= prob.nVertLevels
        term = (prob.tracerCur[:, c] * wgt) * (coef1 + coef3 * sgn)
        out = np.where(inside, out + term, out)
    return out


CASES = {
    # name: (nEdges, nCells, nVertLevels, nAdv, packn)
    "namelist_shape_small": (640, 70, 100, 10, 1),
    "gpu_target_pack8": (300, 40, 100, 10, 8),  # Makefile:110 F90_PACK_SIZE=8 -> nvldim 104
    "odd_levels_scalar_path": (257, 33, 37, 7, 1),  # nvldim odd -> the one-level-per-thread kernels
    "single_edge_single_cell": (1, 1, 5, 3, 1),
    "no_contributing_cells": (19, 4, 12, 0, 1),
}


def make(nested, name, ragged=False, seed=3):
    ne, nc, nl, na, pk = CASES[name]
    prob = nested.NestedProblem(nEdges=ne, nCells=nc, nVertLevels=nl, nAdv=na, nIters=1, packn=pk, seed=seed)
    if ragged and na:
        rng = np.random.default_rng(seed + 100)
        prob.nAdvCellsForEdge = rng.integers(0, na + 1, size=ne).astype(np.int32)
        prob.minLevelCell = rng.integers(1, 4, size=nc).astype(np.int32)  # ice-shelf case, nested.F90:62
        lev = np.arange(1, prob.nvldim + 1)[:, None]
        active = (lev >= prob.minLevelCell[None, :]) & (lev <= prob.maxLevelCell[None, :])
        prob.tracerCur = np.asfortranarray(np.where(active, prob.tracerCur, 0.0))
        prob.cellMask = np.asfortranarray(active.astype(np.float64))
        prob.advMaskHighOrder[: prob.nVertLevels] = rng.integers(0, 2, size=(prob.nVertLevels, ne)).astype(np.float64)
        prob.normalThicknessFlux[3 % prob.nVertLevels, :: 5] = 0.0  # sign(1, 0) = +1
    return prob


# ------------------------------------------------------------------------------------------ CPU


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("ragged", [False, True])
def test_oracle_forms_agree_like_the_reference_check(nested, cko, name, ragged):
    prob = make(nested, name, ragged)
    ref = cko.original_cpu(prob)
    assert np.array_equal(ref, numpy_reference(prob)), "C restatement differs from the numpy restatement"
    for label, flx in (
        ("gpu_optimized", cko.gpu_optimized(prob)),
        ("impl1", cko.impl1(prob)[0]),
        ("impl2", cko.impl2(prob)),
    ):
        bad, worst = cko.check(flx, ref, prob, nested.ERR_TOL)
        assert bad == 0, (label, bad, worst)
        # with 0/1 masks and tracers that vanish outside the level range the forms are equal bit for bit
        assert np.array_equal(flx[: prob.nVertLevels], ref[: prob.nVertLevels]), label


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("ragged", [False, True])
def test_oracle_is_pinned_on_the_reference_compiled_here(nested, cko, yref, name, ragged):
    """oracle/cke_oracle.c against the reference's own yakl_original_loop / yakl_gpu_optimized
    (nested_yakl_c.cpp:184-243, 260-306) on the same inputs: bit for bit on the nVertLevels rows the
    reference writes (it leaves the F90_PACK_SIZE padding rows of highOrderFlx unset), and zero
    failures of the reference's own errTol test."""
    prob = make(nested, name, ragged)
    nv = prob.nVertLevels
    r = yref.RefProblem(prob)
    ref_orig, ref_opt = r.original_loop(), r.gpu_optimized()
    assert np.array_equal(cko.original_cpu(prob)[:nv], ref_orig[:nv])
    assert np.array_equal(cko.gpu_optimized(prob)[:nv], ref_opt[:nv])
    assert np.array_equal(cko.impl2(prob)[:nv], ref_orig[:nv])  # cke_impl2 = the level-range form
    assert np.array_equal(cko.impl1(prob)[0][:nv], ref_opt[:nv])  # cke_impl1 = the cellMask form (0/1 masks)
    bad, worst = cko.check(ref_opt, ref_orig, prob, nested.ERR_TOL)
    assert bad == 0, worst


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("ragged", [False, True])
def test_translated_fortran_forms_equal_the_compiled_cpp_reference(nested, cko, yref, name, ragged):
    """The Fortran loop forms of `program nested` (run_original_cpu_directive nested.F90:495-564, run_gpu_optimized_directive
    nested.F90:593-630), translated mechanically by oracle/f90_to_c.py, against the reference's own C++ forms of the same loops
    compiled by g++ (nested_yakl_c.cpp:184-306): bit for bit.  Both sides are reference code; what this checks is the
    TRANSLATOR -- the tool that also produces oracle/_ref for the two Fortran-only paths (atmosphere, MPDATA) -- against
    compiled reference code.  It also reproduces, on the translation, that coef3rdOrder = 2.14 is a default-real literal."""
    import ref_nested_f90_py as f90

    if not f90.available():
        pytest.skip("oracle/_ref/libnested_f90ref.so is absent and /root/reference is not here to build it")
    prob = make(nested, name, ragged)
    nv = prob.nVertLevels
    r = yref.RefProblem(prob)
    t_orig, t_opt = f90.run(prob, "original"), f90.run(prob, "gpu_optimized")
    assert np.array_equal(t_orig[:nv].view(np.uint64), r.original_loop()[:nv].view(np.uint64))
    assert np.array_equal(t_opt[:nv].view(np.uint64), r.gpu_optimized()[:nv].view(np.uint64))
    assert np.array_equal(t_orig[:nv], cko.original_cpu(prob)[:nv]) and np.array_equal(t_opt[:nv], cko.gpu_optimized(prob)[:nv])
    # the reference's own acceptance test between its two Fortran forms (errTol, nested.F90:267-279)
    bad, worst = cko.check(t_opt, t_orig, prob, nested.ERR_TOL)
    assert bad == 0, worst
    # the translation takes coef3rdOrder from the module's own `parameter :: coef3rdOrder = 2.14` (nested_vars.F90:32), the C++
    # side gets the value the problem carries: equal bits above mean the problem's float32(2.14) is what the Fortran uses
    assert prob.coef3rdOrder == float(np.float32(2.14)) != 2.14


def test_oracle_hashes_to_the_committed_digests_of_the_reference(nested, cko):
    """tests/golden/nested_reference.json (made by tests/golden/make_nested_golden.py from the reference's C++ compiled here; the
    translated Fortran forms gave the same bytes) against the hand-written oracle: runs anywhere, with or without oracle/_ref."""
    import hashlib
    import json

    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "nested_reference.json")))
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()  # noqa: E731
    for name, c in gold["cases"].items():
        prob = nested.NestedProblem(nEdges=c["nEdges"], nCells=c["nCells"], nVertLevels=c["nVertLevels"], nAdv=c["nAdv"], nIters=1,
                                    packn=c["packn"], seed=c["seed"])
        assert sha(prob.tracerCur) == c["tracer_sha256"], name  # the seeded problem is the one the digests were made on
        nv = prob.nVertLevels
        assert sha(cko.original_cpu(prob)[:nv]) == c["original_sha256"], name
        assert sha(cko.gpu_optimized(prob)[:nv]) == c["gpu_optimized_sha256"], name


def test_oracle_vs_reference_with_general_masks(nested, cko, yref):
    """Non-0/1 cellMask and tracers that do not vanish outside the level range: the two reference forms differ from
    each other, and each oracle form still follows its own reference form bit for bit."""
    prob = make(nested, "gpu_target_pack8", ragged=True)
    rng = np.random.default_rng(5)
    prob.cellMask = np.asfortranarray(rng.random(prob.cellMask.shape))
    prob.tracerCur = np.asfortranarray(15.0 * rng.random(prob.tracerCur.shape))
    nv = prob.nVertLevels
    r = yref.RefProblem(prob)
    ref_orig, ref_opt = r.original_loop(), r.gpu_optimized()
    assert not np.array_equal(ref_orig[:nv], ref_opt[:nv])
    assert np.array_equal(cko.original_cpu(prob)[:nv], ref_orig[:nv])
    assert np.array_equal(cko.gpu_optimized(prob)[:nv], ref_opt[:nv])
    # (cke_impl2 multiplies by cellMask inside the level range, cke_impl2.cpp:58-59, so with a fractional mask it is
    # neither of the two YAKL forms; it is pinned on them through the 0/1-mask cases above.)


def test_oracle_vs_reference_at_the_namelist_sizes(nested, cko, yref):
    """nested.nml:1-7 as shipped, both paddings (F90_PACK_SIZE 1 and 8)."""
    for packn in (1, 8):
        prob = nested.NestedProblem(packn=packn, seed=11)
        nv = prob.nVertLevels
        r = yref.RefProblem(prob)
        ref_orig, ref_opt = r.original_loop(), r.gpu_optimized()
        assert np.array_equal(cko.original_cpu(prob)[:nv], ref_orig[:nv])
        assert np.array_equal(cko.gpu_optimized(prob)[:nv], ref_opt[:nv])
        bad, worst = cko.check(ref_opt, ref_orig, prob, nested.ERR_TOL)
        assert bad == 0 and worst == 0.0


def test_oracle_known_answer_by_hand(nested, cko):
    """Two cells, one edge, three levels, worked by hand from nested.F90:130-155."""
    prob = nested.NestedProblem(nEdges=1, nCells=2, nVertLevels=3, nAdv=2, nIters=1, seed=0)
    prob.minLevelCell[:] = [1, 2]
    prob.maxLevelCell[:] = [3, 2]
    prob.tracerCur = np.asfortranarray(np.array([[1.0, 0.0], [2.0, 5.0], [4.0, 0.0]]))
    prob.cellMask = np.asfortranarray((prob.tracerCur != 0).astype(np.float64))
    prob.advCellsForEdge = np.asfortranarray(np.array([[1], [2]], dtype=np.int32))
    prob.advCoefs = np.asfortranarray(np.array([[0.5], [0.25]]))
    prob.advCoefs3rd = np.asfortranarray(np.array([[1.0], [2.0]]))
    prob.coef3rdOrder = 2.0
    prob.normalThicknessFlux = np.asfortranarray(np.array([[3.0], [-2.0], [1.0]]))
    prob.advMaskHighOrder = np.asfortranarray(np.array([[1.0], [1.0], [0.0]]))
    # level 1: 1*3*(0.5+2)            = 7.5
    # level 2: 2*(-2)*(0.5-2) + 5*(-2)*(0.25-4) = 6 + 37.5 = 43.5
    # level 3: mask 0                 = 0
    want = np.array([[7.5], [43.5], [0.0]])
    assert np.array_equal(cko.original_cpu(prob), want)
    assert np.array_equal(cko.impl1(prob)[0], want)
    assert np.array_equal(cko.impl2(prob), want)


def test_coef3rdorder_is_the_float32_literal(nested):
    # nested_vars.F90:32 assigns the default-real literal 2.14 to a double
    p = nested.NestedProblem(nEdges=1, nCells=1, nVertLevels=3, nAdv=1)
    assert p.coef3rdOrder == 2.1400001049041748 and p.coef3rdOrder != 2.14


def test_generator_follows_the_reference_recipe(nested):
    p = nested.NestedProblem(nEdges=500, nCells=400, nVertLevels=100, nAdv=10, packn=8, seed=1)
    assert p.nvldim == 104 and p.nvlpk == 13  # nested_vars.F90:107-108
    assert p.minLevelCell.min() == 1 and p.maxLevelCell.min() >= 3 and p.maxLevelCell.max() == 100
    assert 0.35 < np.mean(p.maxLevelCell == 100) < 0.65  # "a good fraction (half?) at the max depth"
    assert p.advCellsForEdge.min() >= 1 and p.advCellsForEdge.max() <= 400
    assert np.all((p.tracerCur != 0) <= (p.cellMask == 1)) and p.tracerCur.max() < 15.0
    assert np.all(p.tracerCur[100:] == 0) and np.all(p.normalThicknessFlux[100:] == 0)
    assert np.abs(p.normalThicknessFlux).max() <= 7.5 and np.all(p.advMaskHighOrder[:100] == 1)


def test_library_exports_every_declared_symbol(nested):
    L = nested.lib()
    names = nested.declared_symbols()
    assert {"kokkos_init", "kokkos_finalize", "cke_init", "cke_get_results", "cke_cleanup", "cke_impl1_run",
            "cke_impl2_run"} <= set(names)
    for n in names:
        assert hasattr(L, n), n


def test_header_repeats_the_reference_interface(nested):
    """Argument lists of include/cke_b200.h == nested_loops/cke.hpp:7-23 (only where the reference
    tree is present: this container, never the GPU box)."""
    ref = "/root/reference/nested_loops/cke.hpp"
    if not os.path.exists(ref):
        pytest.skip("reference tree not present")

    def protos(text):
        text = re.sub(r"/\*.*?\*/|//[^\n]*", "", text, flags=re.S)
        out = {}
        for m in re.finditer(r"void\s+((?:cke|kokkos)_\w+)\s*\(([^)]*)\)\s*;", text):
            args = [re.sub(r"\s+", " ", a).strip() for a in m.group(2).split(",")]
            args = [re.sub(r"\s*\*\s*", "* ", a).strip() for a in args if a not in ("", "void")]
            out[m.group(1)] = args
        return out

    mine = protos(open(os.path.join(ROOT, "include", "cke_b200.h")).read())
    theirs = protos(open(ref).read())
    assert len(theirs) == 7
    for name, args in theirs.items():
        assert mine.get(name) == args, (name, mine.get(name), args)


def test_no_cpu_fallback_without_a_device(nested):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    prob = make(nested, "single_edge_single_cell")
    nl = nested.NestedLoops()
    with pytest.raises(nested.CkeError, match="no CUDA device|no CPU path|sm_"):
        nl.cke_init(prob)
    with pytest.raises(nested.CkeError, match="before a successful cke_init"):
        nl.cke_impl1_run()


@pytest.fixture(scope="module")
def emul():
    drv = os.path.join(ROOT, "tests", "driver")
    subprocess.run(["make", "-C", drv, "libcke_emul.so"], check=True, stdout=subprocess.DEVNULL)
    L = C.CDLL(os.path.join(drv, "libcke_emul.so"))
    i32, p = C.c_int, C.c_void_p
    L.cke_emul_iteration.argtypes = [i32] * 10 + [p] * 10 + [C.c_double, p]
    L.cke_emul_iteration.restype = i32

    def run(prob, impl, grid, block, edges_per_cta, tracer=None, unroll=5):
        f = lambda x, dt: np.asfortranarray(x, dtype=dt)  # noqa: E731
        tr = np.array(prob.tracerCur if tracer is None else tracer, dtype=np.float64, order="F", copy=True)
        out = np.full((prob.nvldim, prob.nEdges), np.nan, order="F")
        keep = [f(prob.nAdvCellsForEdge, np.int32), f(prob.minLevelCell, np.int32), f(prob.maxLevelCell, np.int32),
                f(prob.advCellsForEdge, np.int32), tr, f(prob.normalThicknessFlux, np.float64),
                f(prob.advMaskHighOrder, np.float64), f(prob.cellMask, np.float64), f(prob.advCoefs, np.float64),
                f(prob.advCoefs3rd, np.float64)]
        rc = L.cke_emul_iteration(unroll, impl, grid, block, edges_per_cta, prob.nEdges, prob.nCells, prob.nVertLevels,
                                  prob.nvldim, prob.nAdv, *[k.ctypes.data for k in keep], prob.coef3rdOrder,
                                  out.ctypes.data)
        assert rc == 0
        return out, tr

    return run


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("ragged", [False, True])
def test_emulated_original_loop_form_against_the_reference_compiled_here(nested, yref, emul, name, ragged):
    """The device source's level-range form WITHOUT the cellMask factor (what libcke_b200's yakl_original_loop launches)
    against the reference's own yakl_original_loop, with general masks and tracers that do not vanish outside the ranges."""
    prob = make(nested, name, ragged)
    rng = np.random.default_rng(9)
    prob.cellMask = np.asfortranarray(rng.random(prob.cellMask.shape))
    prob.tracerCur = np.asfortranarray(15.0 * rng.random(prob.tracerCur.shape))
    nv = prob.nVertLevels
    got, tr = emul(prob, 3, 3, 64, 4)
    assert np.array_equal(tr, prob.tracerCur)
    assert np.array_equal(got[:nv], yref.RefProblem(prob).original_loop()[:nv])


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("ragged", [False, True])
@pytest.mark.parametrize("geom", [(3, 32, 8, 5), (5, 64, 4, 1), (1, 16, 1, 10), (2, 256, 14, 5)])
def test_device_source_emulated_on_host_is_bit_identical_to_the_oracle(nested, cko, emul, name, ragged, geom):
    """The kernel source of libcke_b200.so (cke_kernels.cuh), compiled for the host with threads and
    a barrier standing in for a CTA, against the oracle: grid-stride over edge blocks, staging,
    the two-levels-per-thread and scalar paths, ragged nAdvCellsForEdge, level ranges."""
    grid, block, epc, unroll = geom
    prob = make(nested, name, ragged)
    if prob.nEdges > 300:
        prob.nEdges = 300  # keep the thread emulation quick: the first 300 edges of the same arrays
        for a in ("nAdvCellsForEdge",):
            setattr(prob, a, getattr(prob, a)[:300].copy())
        for a in ("advCellsForEdge", "advCoefs", "advCoefs3rd", "normalThicknessFlux", "advMaskHighOrder", "highOrderFlx"):
            setattr(prob, a, np.asfortranarray(getattr(prob, a)[:, :300]))
    want1, want_tr = cko.impl1(prob)
    got1, got_tr = emul(prob, 1, grid, block, epc, unroll=unroll)
    assert np.array_equal(got1, want1)
    assert np.array_equal(got_tr, want_tr)
    got2, _ = emul(prob, 2, grid, block, epc, unroll=unroll)
    assert np.array_equal(got2, cko.impl2(prob))
    bad, worst = cko.check(got1, cko.original_cpu(prob), prob, nested.ERR_TOL)
    assert bad == 0, worst


def test_emulated_device_source_with_general_masks(nested, cko, emul):
    """Masks that are not 0/1 and tracers that do not vanish outside the level range: the forms
    then differ from each other, and the device source must still follow ITS form bit for bit
    (impl1 masks in place every iteration, impl2 honours the level range)."""
    prob = make(nested, "odd_levels_scalar_path", ragged=True)
    rng = np.random.default_rng(9)
    prob.cellMask = np.asfortranarray(rng.random(prob.cellMask.shape))
    prob.tracerCur = np.asfortranarray(15.0 * rng.random(prob.tracerCur.shape))
    w1, tr1 = cko.impl1(prob)
    g1, gtr1 = emul(prob, 1, 4, 32, 8)
    assert np.array_equal(g1, w1) and np.array_equal(gtr1, tr1)
    w1b, tr2 = cko.impl1(prob, tracer=tr1)  # second iteration: mask applied twice
    g1b, gtr2 = emul(prob, 1, 4, 32, 8, tracer=gtr1)
    assert np.array_equal(g1b, w1b) and np.array_equal(gtr2, tr2) and not np.array_equal(w1b, w1)
    g2, _ = emul(prob, 2, 4, 32, 8)
    assert np.array_equal(g2, cko.impl2(prob))


# ------------------------------------------------------------------------------------------ GPU


def _run_gpu(nested, prob, impl, iters=1):
    nl = nested.NestedLoops()
    nl.kokkos_init()
    nl.cke_init(prob, nIters=iters)
    try:
        (nl.cke_impl1_run if impl == 1 else nl.cke_impl2_run)()
        launches = nl.launch_count()
        out = nl.cke_get_results(np.full((prob.nvldim, prob.nEdges), np.nan, order="F"))
        tracer = nl.get_tracer()
        again = nl.cke_get_results(np.full((prob.nvldim, prob.nEdges), np.nan, order="F"))
    finally:
        nl.cke_cleanup()
        nl.kokkos_finalize()
    return out, tracer, again, launches


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("ragged", [False, True])
def test_gpu_bit_identical_to_the_oracle(nested, cko, name, ragged):
    prob = make(nested, name, ragged)
    ref = cko.original_cpu(prob)
    want1, want_tr = cko.impl1(prob)
    out, tracer, again, launches = _run_gpu(nested, prob, 1)
    assert launches == (2 if prob.nEdges else 0)
    assert np.array_equal(out, want1)
    assert np.array_equal(tracer, want_tr)
    assert np.all(again == 0.0)  # cke_get_results zeroes the device copy, cke.cpp:159
    bad, worst = cko.check(out, ref, prob, nested.ERR_TOL)  # the reference's own test, nested.F90:417-431
    assert bad == 0, worst
    out2, _, _, launches2 = _run_gpu(nested, prob, 2)
    assert launches2 == 1
    assert np.array_equal(out2, cko.impl2(prob))
    bad, worst = cko.check(out2, ref, prob, nested.ERR_TOL)
    assert bad == 0, worst


@pytest.mark.gpu
def test_gpu_namelist_problem_many_iterations(nested, cko):
    """nested.nml:2-6 as shipped (25,600 edges, 2,800 cells, 100 levels, nAdv 10), at the GPU
    target's padding (F90_PACK_SIZE=8, Makefile:110), nIters iterations inside one run call."""
    prob = nested.NestedProblem(packn=8, seed=11)
    ref = cko.original_cpu(prob)
    want1, want_tr = cko.impl1(prob)
    out, tracer, _, launches = _run_gpu(nested, prob, 1, iters=7)
    assert launches == 14
    assert np.array_equal(out, want1) and np.array_equal(tracer, want_tr)  # 0/1 masks: iterations are idempotent
    bad, worst = cko.check(out, ref, prob, nested.ERR_TOL)
    assert bad == 0, worst
    out2, _, _, _ = _run_gpu(nested, prob, 2, iters=3)
    assert np.array_equal(out2, cko.impl2(prob))


@pytest.mark.gpu
def test_gpu_general_masks_follow_each_form(nested, cko):
    prob = make(nested, "gpu_target_pack8", ragged=True)
    rng = np.random.default_rng(5)
    prob.cellMask = np.asfortranarray(rng.random(prob.cellMask.shape))
    prob.tracerCur = np.asfortranarray(15.0 * rng.random(prob.tracerCur.shape))
    w, tr = cko.impl1(prob)
    w, tr = cko.impl1(prob, tracer=tr)  # two iterations: the mask is applied twice
    out, tracer, _, _ = _run_gpu(nested, prob, 1, iters=2)
    assert np.array_equal(out, w) and np.array_equal(tracer, tr)
    out2, _, _, _ = _run_gpu(nested, prob, 2)
    assert np.array_equal(out2, cko.impl2(prob))


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("ragged", [False, True])
def test_gpu_bit_identical_to_the_reference_compiled_here(nested, yref, name, ragged):
    """The CUDA path against oracle/_ref directly (no restatement in the loop): cke_impl1 == the reference's
    yakl_gpu_optimized and cke_impl2 == its yakl_original_loop, bit for bit on the rows the reference writes."""
    prob = make(nested, name, ragged)
    nv = prob.nVertLevels
    r = yref.RefProblem(prob)
    out1, _, _, _ = _run_gpu(nested, prob, 1)
    assert np.array_equal(out1[:nv], r.gpu_optimized()[:nv])
    out2, _, _, _ = _run_gpu(nested, prob, 2)
    assert np.array_equal(out2[:nv], r.original_loop()[:nv])


@pytest.mark.gpu
def test_gpu_namelist_problem_against_the_reference_compiled_here(nested, yref):
    """nested.nml:1-7 as shipped at the GPU target's padding; errTol test of nested.F90:267-279 against the
    reference's own original loop, and in fact bit-identical."""
    prob = nested.NestedProblem(packn=8, seed=11)
    nv = prob.nVertLevels
    r = yref.RefProblem(prob)
    ref_orig = r.original_loop()
    out1, _, _, _ = _run_gpu(nested, prob, 1, iters=3)
    out2, _, _, _ = _run_gpu(nested, prob, 2, iters=3)
    for out in (out1, out2):
        rel = np.abs(out[:nv] - ref_orig[:nv]) / np.maximum(np.abs(ref_orig[:nv]), 1e-300)
        assert (rel <= nested.ERR_TOL).all()
        assert np.array_equal(out[:nv], ref_orig[:nv])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["namelist_shape_small", "gpu_target_pack8", "odd_levels_scalar_path", "single_edge_single_cell"])
def test_gpu_yakl_entry_points_against_the_reference_compiled_here(nested, yref, name):
    """libcke_b200's yakl_init / yakl_init_arrays / yakl_original_loop / yakl_gpu_optimized (the reference's YAKL-side
    interface, nested_yakl_c.cpp:130-306, arrays published through the c_* globals) against the SAME entry points of the
    reference compiled here, on general masks: bit for bit on the rows the reference writes; the host tracer is input only;
    two calls in a row give the same result (the device tracer is re-uploaded by every call)."""
    prob = make(nested, name, ragged=True)
    rng = np.random.default_rng(7)
    prob.cellMask = np.asfortranarray(rng.random(prob.cellMask.shape))
    prob.tracerCur = np.asfortranarray(15.0 * rng.random(prob.tracerCur.shape))
    nv = prob.nVertLevels
    ref = yref.RefProblem(prob)
    yk = nested.YaklLoops(prob)
    try:
        for _ in range(2):
            assert np.array_equal(yk.yakl_gpu_optimized()[:nv], ref.gpu_optimized()[:nv])
            assert np.array_equal(yk.yakl_original_loop()[:nv], ref.original_loop()[:nv])
        tr2 = np.asfortranarray(3.0 * rng.random(prob.tracerCur.shape))
        assert np.array_equal(yk.yakl_original_loop(tr2)[:nv], ref.original_loop(tr2)[:nv])
    finally:
        yk.cleanup()


@pytest.mark.gpu
def test_gpu_rejects_bad_connectivity(nested):
    prob = make(nested, "single_edge_single_cell")
    prob.advCellsForEdge[0, 0] = 7
    nl = nested.NestedLoops()
    with pytest.raises(nested.CkeError, match="outside 1.."):
        nl.cke_init(prob)


# ---- oracle-free properties (hold at any size): the flux is linear in the tracer, and scaling by a power of two is exact


def _linearity_holds(run, prob):
    """run(prob) -> highOrderFlx.  flux(4 * tracer) == 4 * flux(tracer) bit for bit (every product and sum
    scales exactly by a power of two), flux(0) == 0, and reversing the sign of normalThicknessFlux with the
    3rd-order coefficients zeroed reverses the sign of the flux exactly."""
    import copy

    base = run(prob)
    p4 = copy.copy(prob)
    p4.tracerCur = np.asfortranarray(4.0 * prob.tracerCur)
    assert np.array_equal(run(p4), 4.0 * base)
    p0 = copy.copy(prob)
    p0.tracerCur = np.asfortranarray(0.0 * prob.tracerCur)
    assert not np.any(run(p0))
    pm = copy.copy(prob)
    pm.advCoefs3rd = np.asfortranarray(0.0 * prob.advCoefs3rd)
    plus = run(pm)
    pm2 = copy.copy(pm)
    pm2.normalThicknessFlux = np.asfortranarray(-pm.normalThicknessFlux)
    assert np.array_equal(run(pm2), -plus)
    assert np.any(base != 0.0)


def test_emulated_device_source_is_linear_in_the_tracer(nested, emul):
    prob = make(nested, "gpu_target_pack8", ragged=True)
    _linearity_holds(lambda p: emul(p, 1, 4, 64, 14, unroll=5)[0], prob)
    _linearity_holds(lambda p: emul(p, 2, 4, 64, 14, unroll=5)[0], prob)


@pytest.mark.gpu
def test_gpu_flux_is_linear_in_the_tracer_at_the_namelist_sizes(nested):
    """No oracle in the loop: nested.nml sizes, both implementations."""
    prob = nested.NestedProblem(packn=8, seed=21)
    _linearity_holds(lambda p: _run_gpu(nested, p, 1)[0], prob)
    _linearity_holds(lambda p: _run_gpu(nested, p, 2)[0], prob)


# ---- race check of the staged kernel: host threads + barrier under ThreadSanitizer


def _tsan_run(tmp_path, emul_source=None, tag="real"):
    drv = os.path.join(ROOT, "tests", "driver")
    exe = str(tmp_path / f"cke_tsan_{tag}")
    cmd = ["/usr/bin/g++", "-O1", "-g", "-fsanitize=thread", "-ffp-contract=off", "-std=c++20", "-pthread", "-I" + drv, "-o", exe,
           os.path.join(drv, "cke_tsan.cpp")]
    if emul_source:
        cmd.insert(1, f'-DCKE_EMUL_SOURCE="{emul_source}"')
    b = subprocess.run(cmd, capture_output=True, text=True)
    if b.returncode != 0:
        pytest.skip("ThreadSanitizer build not available here: " + b.stderr[-200:])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600, env=dict(os.environ, TSAN_OPTIONS="halt_on_error=0 exitcode=0"))
    return r.returncode, r.stdout + r.stderr


def test_device_source_has_no_shared_memory_race_under_tsan(tmp_path):
    """The kernel as shipped is clean; the same kernel without the barrier between staging the edge records and
    using them is reported -- so the check would see a missing __syncthreads."""
    rc, out = _tsan_run(tmp_path)
    assert rc == 0 and out.count(" rc 0 ") == 4, out[-500:]
    assert "ThreadSanitizer" not in out, out[-800:]
    src = open(os.path.join(ROOT, "codesign-kernels_b200", "csrc", "cke_kernels.cuh")).read()
    at = src.index("// ---- phase B")
    cut = src.rfind("__syncthreads();", 0, at)
    (tmp_path / "cke_kernels_nobarrier.cuh").write_text(src[:cut] + "/* barrier removed */" + src[cut + len("__syncthreads();"):])
    emu = open(os.path.join(ROOT, "tests", "driver", "cke_emul.cpp")).read()
    inc = '#include "../../codesign-kernels_b200/csrc/cke_kernels.cuh"'
    assert inc in emu
    (tmp_path / "cke_emul_nobarrier.cpp").write_text(emu.replace(inc, f'#include "{tmp_path}/cke_kernels_nobarrier.cuh"'))
    rc, out = _tsan_run(tmp_path, str(tmp_path / "cke_emul_nobarrier.cpp"), tag="nobarrier")
    assert "ThreadSanitizer: data race" in out
