// bwk_plan.cpp -- host-side DSS plan builder (no CUDA).
//
// Input: global degree-of-freedom ids per GLL point (HOMME's elem%gdofP idea).  Output: for
// every local element the rows of neighbouring elements its 12 boundary points must add
// (the role of edgeVpack/edgeVunpack putmap/getmap in upstream HOMME edge_mod_base.F90; the
// reference atmosphere/biharmonic_wk_kernel.F90 has no DSS, only the comment F90:518-519),
// plus the per-peer send/recv row lists of the halo exchange (bndry_exchangeV).
#include "bwk_internal.h"

#include <algorithm>
#include <cstring>
#include <set>

namespace bwk {

void partition(int64_t n, int rank, int nranks, int64_t *first, int64_t *count)
{
    // floor(r*n/P) boundaries: contiguous blocks whose sizes differ by at most one
    const int64_t a = (int64_t)(((__int128)n * rank) / nranks);
    const int64_t b = (int64_t)(((__int128)n * (rank + 1)) / nranks);
    *first = a;
    *count = b - a;
}

int owner_of(int64_t n, int nranks, int64_t e)
{
    int r = (int)(((__int128)(e + 1) * nranks - 1) / n); // first guess, then fix up
    if (r < 0) r = 0;
    if (r >= nranks) r = nranks - 1;
    for (;;) {
        int64_t f, c;
        partition(n, r, nranks, &f, &c);
        if (e < f) --r;
        else if (e >= f + c) ++r;
        else return r;
    }
}

// integer-lattice cube, no orientation tables.
void cubed_sphere_gdof(int ne, int64_t *gdof)
{
    const int64_t L = 3 * (int64_t)ne, L1 = L + 1;
    for (int f = 0; f < 6; ++f)
        for (int je = 0; je < ne; ++je)
            for (int ie = 0; ie < ne; ++ie) {
                const int64_t e = ((int64_t)f * ne + je) * ne + ie;
                for (int j = 0; j < 4; ++j)
                    for (int i = 0; i < 4; ++i) {
                        const int64_t A = 3 * ie + i, B = 3 * je + j;
                        int64_t x, y, z;
                        if (f == 0) { x = L; y = A; z = B; }
                        else if (f == 1) { x = L - A; y = L; z = B; }
                        else if (f == 2) { x = 0; y = L - A; z = B; }
                        else if (f == 3) { x = A; y = 0; z = B; }
                        else if (f == 4) { x = A; y = B; z = L; }
                        else { x = A; y = B; z = 0; }
                        gdof[16 * e + i + 4 * j] = (x * L1 + y) * L1 + z;
                    }
            }
}

namespace {

struct Groups {
    // group g holds members member[start[g] .. start[g+1]) (global point ids, ascending)
    std::vector<int64_t> start, member;
    std::vector<int32_t> group_of; // [16*nelem] (int32 is enough up to 2^31 groups)
    int64_t ngroups = 0;
    void build(int64_t npts, const int64_t *gdof)
    {
        std::vector<std::pair<int64_t, int64_t>> kp((size_t)npts);
        for (int64_t p = 0; p < npts; ++p) kp[(size_t)p] = {gdof[p], p};
        std::sort(kp.begin(), kp.end());
        member.resize((size_t)npts);
        group_of.resize((size_t)npts);
        start.clear();
        for (int64_t p = 0; p < npts; ++p) {
            if (p == 0 || kp[(size_t)p].first != kp[(size_t)p - 1].first) start.push_back(p);
            member[(size_t)p] = kp[(size_t)p].second;
            group_of[(size_t)kp[(size_t)p].second] = (int32_t)(start.size() - 1);
        }
        start.push_back(npts);
        ngroups = (int64_t)start.size() - 1;
    }
    const int64_t *begin(int64_t gp) const { return &member[(size_t)start[(size_t)group_of[(size_t)gp]]]; }
    const int64_t *end(int64_t gp) const { return &member[(size_t)start[(size_t)group_of[(size_t)gp] + 1]]; }
};

struct EdgeNbr {
    int64_t elem = -1; // global neighbour element, -1 none
    int edge = 0;      // neighbour's edge facing us
    int rev = 0;
};
struct DiagNbr {
    int64_t elem = -1;
    int pt = 0; // corner point of that element (0,3,12,15)
};
struct ElemTopo {
    EdgeNbr edge[4];
    DiagNbr diag[4];
    int64_t corner_gp[4][4]; // contributors per corner as global point ids: own, S/N nbr, W/E nbr, diag (-1 absent)
};

const int kCornerPt[4] = {0, 3, 12, 15};
// adjacent (S/N edge, t) and (W/E edge, t) of each corner
const int kCornerSN[4] = {EDGE_S, EDGE_S, EDGE_N, EDGE_N};
const int kCornerSNt[4] = {0, 3, 0, 3};
const int kCornerWE[4] = {EDGE_W, EDGE_E, EDGE_W, EDGE_E};
const int kCornerWEt[4] = {0, 0, 3, 3};

// classifies a mid-edge point of an element: which edge it is on and its position along it
bool mid_edge_of(int pt, int *edge, int *t)
{
    const int i = pt % 4, j = pt / 4;
    const bool ib = (i == 0 || i == 3), jb = (j == 0 || j == 3);
    if (ib == jb) return false; // corner or interior
    if (jb) { *edge = (j == 0) ? EDGE_S : EDGE_N; *t = i; }
    else { *edge = (i == 0) ? EDGE_W : EDGE_E; *t = j; }
    return true;
}

bool elem_topology(const Groups &G, const int64_t *gdof, int64_t e, ElemTopo *T, std::string *err)
{
    for (int ed = 0; ed < 4; ++ed) {
        EdgeNbr nb;
        int64_t other[2] = {-1, -1};
        for (int m = 0; m < 2; ++m) {
            const int64_t gp = 16 * e + edge_pt(ed, 1 + m);
            const int64_t *b = G.begin(gp), *en = G.end(gp);
            const int64_t n = en - b;
            if (n > 2) { *err = "mid-edge point shared by more than two elements"; return false; }
            for (const int64_t *q = b; q != en; ++q)
                if (*q != gp) other[m] = *q;
        }
        if ((other[0] < 0) != (other[1] < 0)) { *err = "edge with a half-shared neighbour"; return false; }
        if (other[0] >= 0) {
            if (other[0] / 16 != other[1] / 16) { *err = "edge shared with two different elements"; return false; }
            nb.elem = other[0] / 16;
            if (nb.elem == e) { *err = "element is its own neighbour"; return false; }
            int e1, t1, e2, t2;
            if (!mid_edge_of((int)(other[0] % 16), &e1, &t1) || !mid_edge_of((int)(other[1] % 16), &e2, &t2) || e1 != e2 ||
                t1 + t2 != 3 || (t1 != 1 && t1 != 2)) {
                *err = "neighbour's matching points are not one of its edges";
                return false;
            }
            nb.edge = e1;
            nb.rev = (t1 == 2);
            // the end points must coincide as well
            for (int t = 0; t < 4; t += 3) {
                const int tn = nb.rev ? 3 - t : t;
                if (gdof[16 * e + edge_pt(ed, t)] != gdof[16 * nb.elem + edge_pt(nb.edge, tn)]) {
                    *err = "edge end points do not match the neighbour's";
                    return false;
                }
            }
        }
        T->edge[ed] = nb;
    }
    for (int c = 0; c < 4; ++c) {
        const int64_t gp = 16 * e + kCornerPt[c];
        int64_t covered[2] = {-1, -1};
        const EdgeNbr &a = T->edge[kCornerSN[c]];
        if (a.elem >= 0) covered[0] = 16 * a.elem + edge_pt(a.edge, a.rev ? 3 - kCornerSNt[c] : kCornerSNt[c]);
        const EdgeNbr &b2 = T->edge[kCornerWE[c]];
        if (b2.elem >= 0) covered[1] = 16 * b2.elem + edge_pt(b2.edge, b2.rev ? 3 - kCornerWEt[c] : kCornerWEt[c]);
        if (covered[0] >= 0 && covered[0] == covered[1]) { *err = "corner covered twice by the same point"; return false; }
        DiagNbr d;
        for (const int64_t *q = G.begin(gp); q != G.end(gp); ++q) {
            if (*q == gp || *q == covered[0] || *q == covered[1]) continue;
            if (d.elem >= 0) { *err = "corner node shared by more than four elements (unsupported)"; return false; }
            d.elem = *q / 16;
            d.pt = (int)(*q % 16);
            if (d.pt != 0 && d.pt != 3 && d.pt != 12 && d.pt != 15) { *err = "diagonal contributor is not a corner point"; return false; }
            if (d.elem == e) { *err = "element touches itself at a corner"; return false; }
        }
        T->diag[c] = d;
        T->corner_gp[c][0] = gp;
        T->corner_gp[c][1] = covered[0];
        T->corner_gp[c][2] = covered[1];
        T->corner_gp[c][3] = d.elem >= 0 ? 16 * d.elem + d.pt : -1;
    }
    return true;
}

// rows of OTHER elements that element e's gather reads
void sources_of(const ElemTopo &T, std::vector<RowRef> *out)
{
    for (int ed = 0; ed < 4; ++ed)
        if (T.edge[ed].elem >= 0) out->push_back({T.edge[ed].elem, edge_rowid(T.edge[ed].edge)});
    for (int c = 0; c < 4; ++c)
        if (T.diag[c].elem >= 0) out->push_back({T.diag[c].elem, T.diag[c].pt % 4 == 0 ? ROW_W : ROW_E});
}

} // namespace

Plan build_plan(int64_t nelem_global, const int64_t *gdof, int rank, int nranks)
{
    Plan pl;
    pl.nelem_global = nelem_global;
    pl.rank = rank;
    pl.nranks = nranks;
    partition(nelem_global, rank, nranks, &pl.first, &pl.count);
    pl.send.assign((size_t)nranks, {});
    pl.recv.assign((size_t)nranks, {});

    Groups G;
    G.build(16 * nelem_global, gdof);
    pl.unique_nodes = G.ngroups;

    const int64_t lo = pl.first, hi = pl.first + pl.count;
    std::vector<ElemTopo> topo((size_t)pl.count);
    for (int64_t le = 0; le < pl.count; ++le)
        if (!elem_topology(G, gdof, lo + le, &topo[(size_t)le], &pl.error)) {
            pl.error += " (element " + std::to_string(lo + le) + ")";
            return pl;
        }

    // ---- receive lists: remote rows my elements read, per owner, sorted (= wire order)
    std::vector<RowRef> src;
    for (int64_t le = 0; le < pl.count; ++le) {
        src.clear();
        sources_of(topo[(size_t)le], &src);
        for (const RowRef &r : src)
            if (r.elem < lo || r.elem >= hi) pl.recv[(size_t)owner_of(nelem_global, nranks, r.elem)].push_back(r);
    }
    pl.recv_offset.assign((size_t)nranks + 1, 0);
    for (int p = 0; p < nranks; ++p) {
        auto &v = pl.recv[(size_t)p];
        std::sort(v.begin(), v.end());
        v.erase(std::unique(v.begin(), v.end()), v.end());
        pl.recv_offset[(size_t)p + 1] = pl.recv_offset[(size_t)p] + (int64_t)v.size();
    }
    auto halo_slot = [&](const RowRef &r) -> int32_t {
        const int p = owner_of(nelem_global, nranks, r.elem);
        const auto &v = pl.recv[(size_t)p];
        const auto it = std::lower_bound(v.begin(), v.end(), r);
        return (int32_t)(pl.recv_offset[(size_t)p] + (it - v.begin()));
    };

    // ---- send lists: for every remote element that shares a node with one of mine, the rows
    //      of MY elements its gather reads (same function, so both sides agree by construction)
    if (nranks > 1) {
        std::set<int64_t> consumers;
        for (int64_t le = 0; le < pl.count; ++le)
            for (int ed = 0; ed < 4; ++ed)
                for (int t = 0; t < 4; ++t) {
                    const int64_t gp = 16 * (lo + le) + edge_pt(ed, t);
                    for (const int64_t *q = G.begin(gp); q != G.end(gp); ++q) {
                        const int64_t e2 = *q / 16;
                        if (e2 < lo || e2 >= hi) consumers.insert(e2);
                    }
                }
        ElemTopo T;
        for (int64_t e2 : consumers) {
            std::string err;
            if (!elem_topology(G, gdof, e2, &T, &err)) { pl.error = err + " (remote element " + std::to_string(e2) + ")"; return pl; }
            src.clear();
            sources_of(T, &src);
            const int dest = owner_of(nelem_global, nranks, e2);
            for (const RowRef &r : src)
                if (r.elem >= lo && r.elem < hi) pl.send[(size_t)dest].push_back(r);
        }
        for (int p = 0; p < nranks; ++p) {
            auto &v = pl.send[(size_t)p];
            std::sort(v.begin(), v.end());
            v.erase(std::unique(v.begin(), v.end()), v.end());
        }
    }

    // ---- device descriptors
    pl.dss.resize((size_t)pl.count);
    std::vector<uint8_t> is_boundary((size_t)pl.count, 0), is_dependent((size_t)pl.count, 0);
    for (int64_t le = 0; le < pl.count; ++le) {
        const ElemTopo &T = topo[(size_t)le];
        ElemDss d;
        std::memset(&d, 0, sizeof d);
        for (int ed = 0; ed < 4; ++ed) {
            const EdgeNbr &nb = T.edge[ed];
            d.edge_src[ed] = -1;
            if (nb.elem < 0) continue;
            d.edge_rev[ed] = (uint8_t)nb.rev;
            if (nb.elem >= lo && nb.elem < hi) {
                d.edge_src[ed] = (int32_t)(nb.elem - lo);
                if (nb.edge == EDGE_S || nb.edge == EDGE_N) { d.edge_kind[ed] = SRC_TMP; d.edge_row[ed] = (uint8_t)(nb.edge == EDGE_S ? 0 : 3); }
                else { d.edge_kind[ed] = SRC_TMPCOL; d.edge_row[ed] = (uint8_t)(nb.edge == EDGE_W ? 0 : 1); }
            } else {
                d.edge_kind[ed] = SRC_HALO;
                d.edge_row[ed] = 0;
                d.edge_src[ed] = halo_slot({nb.elem, edge_rowid(nb.edge)});
                is_dependent[(size_t)le] = 1;
            }
        }
        for (int c = 0; c < 4; ++c) {
            const DiagNbr &dg = T.diag[c];
            d.diag_src[c] = -1;
            if (dg.elem >= 0) {
                d.diag_pos[c] = (uint8_t)(dg.pt / 4); // position along the W/E column = j
                const int row = dg.pt % 4 == 0 ? ROW_W : ROW_E;
                if (dg.elem >= lo && dg.elem < hi) { d.diag_kind[c] = SRC_TMPCOL; d.diag_row[c] = (uint8_t)(row == ROW_W ? 0 : 1); d.diag_src[c] = (int32_t)(dg.elem - lo); }
                else { d.diag_kind[c] = SRC_HALO; d.diag_row[c] = 0; d.diag_src[c] = halo_slot({dg.elem, row}); is_dependent[(size_t)le] = 1; }
            }
            // summation order: ascending global point id (== the oracle's order)
            int order[4], n = 0;
            for (int k = 0; k < 4; ++k)
                if (T.corner_gp[c][k] >= 0) order[n++] = k;
            for (int a = 1; a < n; ++a) // insertion sort of <= 4 entries
                for (int b = a; b > 0 && T.corner_gp[c][order[b]] < T.corner_gp[c][order[b - 1]]; --b) std::swap(order[b], order[b - 1]);
            uint8_t perm = 0;
            for (int k = 0; k < n; ++k) perm |= (uint8_t)(order[k] << (2 * k));
            d.corner_n[c] = (uint8_t)n;
            d.corner_perm[c] = perm;
        }
        pl.dss[(size_t)le] = d;
    }
    for (int p = 0; p < nranks; ++p)
        for (const RowRef &r : pl.send[(size_t)p]) is_boundary[(size_t)(r.elem - lo)] = 1;
    for (int64_t le = 0; le < pl.count; ++le) {
        (is_boundary[(size_t)le] ? pl.boundary_list : pl.interior_list).push_back((int32_t)le);
        (is_dependent[(size_t)le] ? pl.dependent_list : pl.independent_list).push_back((int32_t)le);
    }

    // ---- sweep-2 strips: runs of consecutive local elements linked E->W with equal orientation
    auto linked = [&](int32_t a, int32_t b) { // is b == a+1 the E neighbour of a, facing with its W edge, not reversed?
        const ElemDss &da = pl.dss[(size_t)a];
        return b == a + 1 && da.edge_src[EDGE_E] == b && da.edge_kind[EDGE_E] == SRC_TMPCOL && da.edge_row[EDGE_E] == 0 && !da.edge_rev[EDGE_E];
    };
    // corner_perm of a face-interior element of a row-major patch (must match the kernel's PERM_*)
    const int std_perm[4] = {3 | (1 << 2) | (2 << 4) | (0 << 6), 1 | (3 << 2) | (0 << 4) | (2 << 6), 2 | (0 << 2) | (3 << 4) | (1 << 6),
                             0 | (2 << 2) | (1 << 4) | (3 << 6)};
    auto regular_strip = [&](int32_t first, int n) {
        StripReg r;
        std::memset(&r, 0, sizeof r);
        bool all_std = true;
        for (int w = 0; w < n; ++w) {
            const ElemDss &d = pl.dss[(size_t)(first + w)];
            const ElemTopo &T = topo[(size_t)(first + w)];
            // S and N: a natural j-row of a local slab, or a halo row, same orientation
            for (int ed = EDGE_S; ed <= EDGE_N; ++ed) {
                if (d.edge_src[ed] < 0 || d.edge_rev[ed]) return r;
                uint8_t c0;
                if (d.edge_kind[ed] == SRC_TMP) c0 = (uint8_t)(4 * d.edge_row[ed]);
                else if (d.edge_kind[ed] == SRC_HALO) c0 = (uint8_t)ROW_IN_HALO;
                else return r; // the neighbour faces us with an i-column: not a TMA row
                (ed == EDGE_S ? r.s_src : r.n_src)[w] = d.edge_src[ed];
                (ed == EDGE_S ? r.s_c0 : r.n_c0)[w] = c0;
            }
            // interior links are mate links by construction of the strip; re-check
            if (w > 0 && !(d.edge_src[EDGE_W] == first + w - 1 && d.edge_kind[EDGE_W] == SRC_TMPCOL && d.edge_row[EDGE_W] == 1 && !d.edge_rev[EDGE_W])) return r;
            if (w < n - 1 && !(d.edge_src[EDGE_E] == first + w + 1 && d.edge_kind[EDGE_E] == SRC_TMPCOL && d.edge_row[EDGE_E] == 0 && !d.edge_rev[EDGE_E])) return r;
            // interior corners: the diagonal contributor must be the facing end of the mate's S/N neighbour row
            // (corner_gp[c] = {own, S/N neighbour's point, W/E neighbour's point, diagonal point}, -1 = none)
            if (w > 0) {
                const ElemTopo &L = topo[(size_t)(first + w - 1)];
                if (T.corner_gp[0][3] >= 0 && T.corner_gp[0][3] != L.corner_gp[1][1]) return r; // SW diag == east end of the left mate's S row
                if (T.corner_gp[2][3] >= 0 && T.corner_gp[2][3] != L.corner_gp[3][1]) return r; // NW diag == east end of its N row
            }
            if (w < n - 1) {
                const ElemTopo &R = topo[(size_t)(first + w + 1)];
                if (T.corner_gp[1][3] >= 0 && T.corner_gp[1][3] != R.corner_gp[0][1]) return r; // SE diag == west end of the right mate's S row
                if (T.corner_gp[3][3] >= 0 && T.corner_gp[3][3] != R.corner_gp[2][1]) return r; // NE diag == west end of its N row
            }
            for (int c = 0; c < 4; ++c) {
                r.corner_n[w][c] = d.corner_n[c];
                r.corner_perm[w][c] = d.corner_perm[c];
                if (d.corner_n[c] != 4 || d.corner_perm[c] != std_perm[c]) all_std = false;
            }
        }
        // the two strip ends: any source
        const ElemDss &d0 = pl.dss[(size_t)first], &d1 = pl.dss[(size_t)(first + n - 1)];
        const ElemDss *de[2] = {&d0, &d1};
        const int end_edge[2] = {EDGE_W, EDGE_E};
        for (int k = 0; k < 2; ++k) {
            const ElemDss &d = *de[k];
            const int ed = end_edge[k];
            if (d.edge_src[ed] < 0) return r; // open boundary: generic path
            r.end_src[k] = d.edge_src[ed];
            r.end_kind[k] = d.edge_kind[ed];
            r.end_row[k] = d.edge_row[ed];
            r.end_rev[k] = d.edge_rev[ed];
        }
        const int dcorner[4] = {0, 2, 1, 3}; // SW, NW of the first member; SE, NE of the last
        for (int k = 0; k < 4; ++k) {
            const ElemDss &d = k < 2 ? d0 : d1;
            const int c = dcorner[k];
            r.diag_src[k] = d.diag_src[c];
            if (d.diag_src[c] < 0) continue;
            r.diag_kind[k] = d.diag_kind[c];
            r.diag_pt[k] = d.diag_kind[c] == SRC_HALO ? d.diag_pos[c] : (uint8_t)((d.diag_row[c] ? 3 : 0) + 4 * d.diag_pos[c]);
        }
        r.all_std = all_std ? 1 : 0;
        r.regular = 1;
        return r;
    };
    auto make_strips = [&](const std::vector<int32_t> &list, std::vector<int32_t> *out, std::vector<StripReg> *reg) {
        out->clear();
        reg->clear();
        size_t i = 0;
        while (i < list.size()) {
            int n = 1;
            while (n < STRIP_MAX && i + n < list.size() && linked(list[i + n - 1], list[i + n])) ++n;
            out->push_back(list[i] * 8 + (n - 1));
            reg->push_back(regular_strip(list[i], n));
            reg->back().first = list[i];
            reg->back().count = n;
            i += (size_t)n;
        }
    };
    {
        std::vector<int32_t> all((size_t)pl.count);
        for (int64_t le = 0; le < pl.count; ++le) all[(size_t)le] = (int32_t)le;
        make_strips(all, &pl.strips_all, &pl.reg_all);
        // A strip is dependent (runs after the halo exchange) as soon as one member reads a received row.  The strips stay
        // whole: cutting the dependent elements out (round 1) left 176 one-element strips per rank at 8 ranks -- the
        // elements at both ends of every row whose face-edge neighbour lives on another rank -- each costing as much as
        // a full strip, and shortened their rows' last strips.
        for (size_t i = 0; i < pl.strips_all.size(); ++i) {
            const int f = strip_first(pl.strips_all[i]), n = strip_count(pl.strips_all[i]);
            bool dep = false;
            for (int e = f; e < f + n; ++e) dep = dep || is_dependent[(size_t)e];
            (dep ? pl.strips_dependent : pl.strips_independent).push_back(pl.strips_all[i]);
            (dep ? pl.reg_dependent : pl.reg_independent).push_back(pl.reg_all[i]);
        }
    }

    // ---- node groups of the local points (rspheremp, tests)
    pl.pt_start.assign((size_t)(16 * pl.count) + 1, 0);
    for (int64_t lp = 0; lp < 16 * pl.count; ++lp) {
        const int64_t gp = 16 * lo + lp;
        pl.pt_start[(size_t)lp + 1] = pl.pt_start[(size_t)lp] + (G.end(gp) - G.begin(gp));
    }
    pl.pt_member.resize((size_t)pl.pt_start.back());
    for (int64_t lp = 0; lp < 16 * pl.count; ++lp) {
        const int64_t gp = 16 * lo + lp;
        std::copy(G.begin(gp), G.end(gp), pl.pt_member.begin() + pl.pt_start[(size_t)lp]);
    }
    return pl;
}

// Host restatement of exactly what the sweep-2 kernel's gather does with an ElemDss.
bool strip_is_fast(const StripReg &r, int32_t strip)
{
    if (!r.regular || strip_count(strip) != STRIP_MAX) return false;
    for (int m = 1; m < STRIP_MAX; ++m)
        if (r.s_src[m] != r.s_src[0] + m || r.s_c0[m] != r.s_c0[0] || r.n_src[m] != r.n_src[0] + m || r.n_c0[m] != r.n_c0[0]) return false;
    return true;
}

bool strips_chain(const StripReg &a, const StripReg &b)
{
    if (b.first != a.first + STRIP_MAX) return false;
    if (b.s_src[0] != a.s_src[0] + STRIP_MAX || b.s_c0[0] != a.s_c0[0] || b.n_src[0] != a.n_src[0] + STRIP_MAX || b.n_c0[0] != a.n_c0[0]) return false;
    if (a.end_kind[1] != SRC_TMPCOL || a.end_src[1] != b.first || a.end_row[1] != 0 || a.end_rev[1]) return false;
    if (b.end_kind[0] != SRC_TMPCOL || b.end_src[0] != a.first + STRIP_MAX - 1 || b.end_row[0] != 1 || b.end_rev[0]) return false;
    auto diag_is = [](const StripReg &r, int d, int32_t src, int c0, int pos) { // diagonal d of r == value `pos` of the row (src, c0)
        if (r.diag_src[d] != src) return false;
        if (c0 == ROW_IN_HALO) return r.diag_kind[d] == SRC_HALO && r.diag_pt[d] == pos;
        return r.diag_kind[d] == SRC_TMPCOL && r.diag_pt[d] == c0 + pos;
    };
    return diag_is(a, 2, b.s_src[0], b.s_c0[0], 0) && diag_is(a, 3, b.n_src[0], b.n_c0[0], 0) &&
           diag_is(b, 0, a.s_src[STRIP_MAX - 1], a.s_c0[STRIP_MAX - 1], 3) && diag_is(b, 1, a.n_src[STRIP_MAX - 1], a.n_c0[STRIP_MAX - 1], 3);
}

StripGroups group_strips(const std::vector<int32_t> &S, const std::vector<StripReg> &R, int max_members)
{
    StripGroups G;
    G.strips.reserve(S.size());
    G.regs.reserve(S.size());
    for (int pass = 0; pass < 2; ++pass) // pass 0 collects the fast strips, pass 1 the rest
        for (size_t i = 0; i < S.size(); ++i) {
            const bool fast = strip_is_fast(R[i], S[i]);
            if (fast == (pass == 0)) { G.strips.push_back(S[i]); G.regs.push_back(R[i]); G.nfast += pass == 0; }
        }
    for (int64_t i = 0; i < G.nfast;) {
        int64_t run = 1;
        while (run < 4 && i + run < G.nfast && strips_chain(G.regs[(size_t)(i + run - 1)], G.regs[(size_t)(i + run)])) ++run;
        const int take = (run >= 4 && max_members >= 16) ? 4 : (run >= 2 && max_members >= 8) ? 2 : 1;
        const int gi = take == 4 ? 2 : take == 2 ? 1 : 0;
        for (int k = 0; k < take; ++k) G.grp[gi].push_back(G.regs[(size_t)(i + k)]);
        i += take;
    }
    return G;
}

void gather_elem_host(const Plan &pl, int64_t le, const double *field, double *out)
{
    const ElemDss &d = pl.dss[(size_t)le];
    const double *own = field + 16 * (pl.first + le);
    for (int t = 0; t < 16; ++t) out[t] = own[t];
    auto fetch_row = [&](int kind, int32_t src, int rowcoord, double *r4) {
        int64_t e;
        int row;
        if (kind == SRC_HALO) {
            int p = 0;
            while (src >= pl.recv_offset[(size_t)p + 1]) ++p;
            const RowRef &rr = pl.recv[(size_t)p][(size_t)(src - pl.recv_offset[(size_t)p])];
            e = rr.elem;
            row = rr.row;
        } else {
            e = pl.first + src;
            row = kind == SRC_TMP ? rowcoord : (rowcoord == 0 ? ROW_W : ROW_E);
        }
        const double *s = field + 16 * e;
        for (int t = 0; t < 4; ++t) r4[t] = row < 4 ? s[t + 4 * row] : s[(row == ROW_W ? 0 : 3) + 4 * t];
    };
    double nb[4][4] = {{0}};
    for (int ed = 0; ed < 4; ++ed) {
        if (d.edge_src[ed] < 0) continue;
        double r4[4];
        fetch_row(d.edge_kind[ed], d.edge_src[ed], d.edge_row[ed], r4);
        for (int t = 0; t < 4; ++t) nb[ed][t] = d.edge_rev[ed] ? r4[3 - t] : r4[t];
        out[edge_pt(ed, 1)] = own[edge_pt(ed, 1)] + nb[ed][1];
        out[edge_pt(ed, 2)] = own[edge_pt(ed, 2)] + nb[ed][2];
    }
    for (int c = 0; c < 4; ++c) {
        double v[4] = {own[kCornerPt[c]], nb[kCornerSN[c]][kCornerSNt[c]], nb[kCornerWE[c]][kCornerWEt[c]], 0.0};
        if (d.diag_src[c] >= 0) {
            double r4[4];
            fetch_row(d.diag_kind[c], d.diag_src[c], d.diag_row[c], r4);
            v[3] = r4[d.diag_pos[c]];
        }
        double s = v[d.corner_perm[c] & 3];
        for (int k = 1; k < d.corner_n[c]; ++k) s = s + v[(d.corner_perm[c] >> (2 * k)) & 3];
        out[kCornerPt[c]] = s;
    }
}

} // namespace bwk
