"""CPU: the DSS plan (host logic of libbwk_b200) against the oracle's node-wise DSS on the
synthetic cubed sphere.  No GPU, no compute kernels."""
import numpy as np
import pytest


@pytest.mark.parametrize("ne", [1, 2, 3, 4, 8])
def test_cubed_sphere_node_counts(bwk, oracle, ne):
    g = bwk.cubed_sphere_gdof(ne)
    assert np.array_equal(g, oracle.cubed_sphere_gdof(ne))
    n = 6 * ne * ne
    d = oracle.Dss(n, g)
    assert d.unique_nodes() == 54 * ne * ne + 2
    m = d.multiplicity_hist()
    assert m[3] == 8 and m[5:].sum() == 0  # exactly the 8 cube vertices have multiplicity 3
    assert m[1] * 1 + m[2] * 2 + m[3] * 3 + m[4] * 4 == 16 * n
    if ne == 4:
        assert (m[1], m[2], m[4]) == (384, 384, 90)
    assert bwk.Plan(n, g).unique_nodes == 54 * ne * ne + 2


@pytest.mark.parametrize("ne,nranks", [(1, 1), (2, 1), (3, 1), (4, 1), (2, 2), (3, 4), (4, 8), (5, 3), (8, 8)])
def test_plan_gather_equals_nodewise_dss_bit_for_bit(bwk, oracle, ne, nranks):
    n = 6 * ne * ne
    g = bwk.cubed_sphere_gdof(ne)
    rng = np.random.default_rng(ne * 100 + nranks)
    field = rng.standard_normal((4, 4, 1, 1, n))
    field = np.asfortranarray(field)
    ref = oracle.Dss(n, g).apply(field.copy(order="F")).reshape(-1, order="F")
    got = np.empty(16 * n)
    for r in range(nranks):
        pl = bwk.Plan(n, g, r, nranks)
        got[16 * pl.local_first : 16 * (pl.local_first + pl.local_count)] = pl.apply_host(field.reshape(-1, order="F"))
    assert np.array_equal(got, ref)  # same values, same summation order


@pytest.mark.parametrize("ne,nranks", [(2, 2), (3, 4), (4, 8), (6, 5)])
def test_send_and_recv_lists_agree_between_ranks(bwk, ne, nranks):
    n = 6 * ne * ne
    g = bwk.cubed_sphere_gdof(ne)
    plans = [bwk.Plan(n, g, r, nranks) for r in range(nranks)]
    total = 0
    for a in range(nranks):
        assert plans[a].send_list(a)[0].size == 0 and plans[a].recv_list(a)[0].size == 0
        for b in range(nranks):
            se, sr = plans[a].send_list(b)
            re_, rr = plans[b].recv_list(a)
            assert np.array_equal(se, re_) and np.array_equal(sr, rr)
            lo, cnt = plans[a].local_first, plans[a].local_count
            assert np.all((se >= lo) & (se < lo + cnt))
            assert set(np.unique(sr)) <= {0, 3, 4, 5}
            total += se.size
    assert total > 0
    # every element that needs a halo is flagged dependent, every feeder boundary
    assert all(p.dependent_elems > 0 and p.boundary_elems > 0 for p in plans)


def test_single_rank_has_no_halo(bwk):
    g = bwk.cubed_sphere_gdof(3)
    pl = bwk.Plan(54, g)
    assert pl.dependent_elems == 0 and pl.boundary_elems == 0


def test_open_boundary_mesh_is_supported(bwk, oracle):
    # a single face of the cube = planar patch with open edges (multiplicity-1 boundary points)
    ne = 3
    g = bwk.cubed_sphere_gdof(ne)[: 16 * ne * ne]
    n = ne * ne
    field = np.asfortranarray(np.random.default_rng(1).standard_normal((4, 4, 1, 1, n)))
    ref = oracle.Dss(n, g).apply(field.copy(order="F")).reshape(-1, order="F")
    got = bwk.Plan(n, g).apply_host(field.reshape(-1, order="F"))
    assert np.array_equal(got, ref)


def test_unsupported_mesh_is_rejected_loudly(bwk):
    # five elements around one corner node: not a structured quad corner
    g = np.arange(16 * 5, dtype=np.int64) + 1000
    g[[0, 16, 32, 48, 64]] = 7
    with pytest.raises(bwk.BwkError):
        bwk.Plan(5, g)


@pytest.mark.parametrize("ne,nranks", [(1, 1), (3, 1), (4, 1), (8, 1), (30, 1), (6, 4), (8, 8)])
def test_strips_tile_the_elements_and_are_west_east_chains(bwk, ne, nranks):
    """Sweep-2 work units: every local element is in exactly one strip, strips hold <= 4 consecutive
    elements, and consecutive members share the edge (E column of e == W column of e+1, same
    orientation) -- read off the gdof ids, independently of the plan."""
    n = 6 * ne * ne
    g = bwk.cubed_sphere_gdof(ne).reshape(n, 4, 4)  # [e][j][i]
    for r in range(nranks):
        pl = bwk.Plan(n, g.reshape(-1), r, nranks)
        lo, cnt = pl.local_first, pl.local_count
        f, c = pl.strips(0)
        assert np.all((c >= 1) & (c <= 4))
        assert f[0] == 0 and np.array_equal(f[1:], (f + c)[:-1]) and f[-1] + c[-1] == cnt
        for a, k in zip(f, c):
            for e in range(lo + a, lo + a + k - 1):
                assert np.array_equal(g[e][:, 3], g[e + 1][:, 0])
        if nranks == 1 and ne % 4 == 0:
            assert np.all(c == 4)  # face rows are cut into full strips
        if nranks == 1:
            # "regular" (fully asynchronous kernel path): every face-interior strip must be regular, and
            # a regular strip must have, for every member, S and N neighbours that face it with one of
            # their natural j-rows in the same orientation (so that TMA can stage the row) -- read off
            # the ids, independently of the plan.  The strip ends are free (face edges qualify).
            reg = pl.strip_regular
            rows = {}
            for e in range(n):
                rows.setdefault(tuple(g[e][0, :]), []).append(e)
                rows.setdefault(tuple(g[e][3, :]), []).append(e)
            nreg = 0
            for a, k, r in zip(f, c, reg):
                ie = (np.arange(a, a + k) % (ne * ne)) % ne
                je = (np.arange(a, a + k) % (ne * ne)) // ne
                interior = np.all((ie > 0) & (ie < ne - 1) & (je > 0) & (je < ne - 1))
                if interior:
                    assert r, (a, k, r)
                if r:
                    nreg += 1
                    for e in range(a, a + k):
                        assert len([x for x in rows[tuple(g[e][0, :])] if x != e]) == 1  # S neighbour: same ids, same order, as a j-row
                        assert len([x for x in rows[tuple(g[e][3, :])] if x != e]) == 1
            if ne >= 8:
                # rows 1..ne-2 of every face are entirely regular now (face-edge columns included)
                assert nreg >= 6 * (ne - 2) * ((ne + 3) // 4)
        fd, cd = pl.strips(1)
        fi, ci = pl.strips(2)
        # the dependent / independent lists partition the strips of list 0 (strips stay whole: a strip is dependent as soon
        # as one member reads a received row), so every dependent element sits in a dependent strip
        assert cd.sum() >= pl.dependent_elems and cd.sum() + ci.sum() == cnt
        assert sorted(zip(f.tolist(), c.tolist())) == sorted(zip(fd.tolist() + fi.tolist(), cd.tolist() + ci.tolist()))
        covered = np.zeros(cnt, dtype=int)
        for ff, cc in ((fd, cd), (fi, ci)):
            for a, k in zip(ff, cc):
                covered[a : a + k] += 1
        assert np.all(covered == 1)


@pytest.mark.parametrize("ne,nlev,qsize", [(2, 3, 2), (4, 5, 1)])
def test_patch_oracle_equals_full_oracle(oracle, ne, nlev, qsize):
    """tests/patch_oracle.py (used to check BASELINE's 32 GB / 258 GB configurations element by element) gives the bits of
    the whole-sphere oracle, at face-interior elements, face edges and cube corners."""
    from patch_oracle import patch_oracle_full

    n = 6 * ne * ne
    dvv, elem, q = oracle.initialize_data(n, nlev, qsize)
    dss = oracle.Dss(n, oracle.cubed_sphere_gdof(ne))
    ref = dss.biharmonic_full(elem, dss.rspheremp(elem), q, dvv)
    gd = oracle.cubed_sphere_gdof(ne).reshape(n, 16)
    for e in sorted({0, ne - 1, ne * ne - 1, ne * ne, n - 1, n // 2, 2 * ne * ne + ne + 1}):
        got = patch_oracle_full(oracle, ne, nlev, qsize, gd, e)
        assert np.array_equal(got[..., 0], ref[..., e]), e


@pytest.mark.parametrize("ne,nranks", [(4, 1), (8, 1), (30, 1), (16, 2), (12, 5), (120, 1), (120, 8)])
def test_warp_strip_groups_are_whole_face_rows_chained_west_to_east(bwk, ne, nranks):
    """The grouping install_plan feeds the warp-strip sweep (bwk_plan.cpp: group_strips), checked against the
    gdof ids alone: the work units partition the strips of the list; a unit of k chained strips holds 4k consecutive
    elements of ONE row of one face, every consecutive pair linked E column -> W column in the natural orientation, and the
    members' S (and N) neighbours are consecutive elements too, all facing with the same one of their rows.  (The NM = 8
    form once chained strips across the end of a face row: wrong values at ne = 4.)"""
    n = 6 * ne * ne
    g = bwk.cubed_sphere_gdof(ne).reshape(n, 4, 4)  # [e][j][i]
    owner = {}
    for e in range(n):
        for j in (0, 3):
            owner.setdefault(tuple(g[e][j, :]), []).append((e, j))  # natural rows
            owner.setdefault(tuple(g[e][j, ::-1]), []).append((e, 10 + j))  # anything else: reversed rows, columns
            owner.setdefault(tuple(g[e][:, j]), []).append((e, 20 + j))
            owner.setdefault(tuple(g[e][::-1, j]), []).append((e, 30 + j))

    def facing(e, j):  # the other element holding row j of e on one of its edges, and how
        o = [x for x in owner[tuple(g[e][j, :])] if x[0] != e]
        assert len(o) == 1
        return o[0]

    ranks = range(nranks) if ne < 100 else (0, nranks - 1)
    for r in ranks:
        pl = bwk.Plan(n, g.reshape(-1), r, nranks)
        lo, cnt = pl.local_first, pl.local_count
        for which in (0, 1, 2):
            sf, sc = pl.strips(which)
            strips = dict(zip(sf.tolist(), sc.tolist()))
            for members in (4, 8, 16):
                gf, gk = pl.strip_groups(which, members)
                assert np.all(gk <= members // 4)
                seen = []
                for a, k in zip(gf.tolist(), gk.tolist()):
                    if k == 0:  # left to the round-1 strip kernel: one strip of the list, untouched
                        seen.append(a)
                        continue
                    for s in range(k):
                        assert strips[a + 4 * s] == 4
                        seen.append(a + 4 * s)
                    el = np.arange(lo + a, lo + a + 4 * k)
                    assert len(set((el // ne).tolist())) == 1  # one face, one row of elements
                    for e in el[:-1]:
                        assert np.array_equal(g[e][:, 3], g[e + 1][:, 0])
                    for j in (0, 3):
                        nb = [facing(int(e), j) for e in el]
                        ids = np.array([x[0] for x in nb])
                        local = (ids >= lo) & (ids < lo + cnt)
                        assert local.all() or not local.any()  # a unit's rows come from tmp or from the halo, never both
                        if local.all():  # rows read from tmp in place: one natural row of consecutive elements
                            assert len({x[1] for x in nb}) == 1 and nb[0][1] in (0, 3)
                            assert np.array_equal(ids, ids[0] + np.arange(4 * k))
                assert sorted(seen) == sorted(strips)
            if which == 0 and nranks == 1 and ne == 120:
                gf, gk = pl.strip_groups(0, 8)
                assert (gk == 2).sum() == 10680 and (gk == 1).sum() == 0 and (gk == 0).sum() == 240  # headline config
            if which == 0 and ne == 4:
                assert np.all(pl.strip_groups(0, 16)[1] <= 1)  # a face row is one strip: nothing to chain
