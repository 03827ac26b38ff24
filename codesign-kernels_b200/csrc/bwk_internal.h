// bwk_internal.h -- shared declarations of libbwk_b200 (not part of the public ABI).
#pragma once
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

namespace bwk {

// Row ids of the edge rows a neighbour can ask for.  Rows 0..3 are the natural j-rows of a
// slab (4 contiguous doubles); 4/5 are the packed i=1 / i=4 columns ("W"/"E" edge pack).
enum : int { ROW_S = 0, ROW_N = 3, ROW_W = 4, ROW_E = 5 };
// Edge order used everywhere: S (j=1), N (j=4), W (i=1), E (i=4).
enum : int { EDGE_S = 0, EDGE_N = 1, EDGE_W = 2, EDGE_E = 3 };
// Where a source row lives on the device: a natural j-row of a slab in the sweep-1 output (4
// contiguous doubles), an i-column of such a slab (4 doubles, 32-byte stride), or a packed row of
// the halo buffer received from another rank.
enum : int { SRC_TMP = 0, SRC_TMPCOL = 1, SRC_HALO = 2 };

// Per-element DSS gather descriptor consumed by the sweep-2 kernel (64 bytes).
// Replaces HOMME's edge putmap/getmap for this path (upstream edge_mod_base.F90; the
// reference itself has no DSS, only the comment F90:518-519).
struct ElemDss {
    int32_t edge_src[4];    // S,N,W,E: local element index (SRC_TMP/SRC_TMPCOL) or halo row slot; -1 = no neighbour
    int32_t diag_src[4];    // corners (1,1),(4,1),(1,4),(4,4): same encoding; -1 = no diagonal contributor
    uint8_t edge_kind[4];   // SRC_*
    uint8_t edge_row[4];    // natural row 0/3 (SRC_TMP), side 0=W/1=E (SRC_TMPCOL), 0 (SRC_HALO)
    uint8_t edge_rev[4];    // 1 = neighbour's edge runs in the opposite direction
    uint8_t diag_kind[4];   // SRC_TMPCOL or SRC_HALO: corner values are the ends of the W/E edge columns
    uint8_t diag_row[4];    // side 0=W/1=E (SRC_TMPCOL) or 0 (SRC_HALO)
    uint8_t diag_pos[4];    // position (0 or 3) of the corner value inside that row
    uint8_t corner_n[4];    // number of contributors to the corner node (own included)
    uint8_t corner_perm[4]; // summation order, 2 bits per slot: 0 own, 1 S/N-edge nbr, 2 W/E-edge nbr, 3 diagonal
};
static_assert(sizeof(ElemDss) == 64, "ElemDss must be 64 bytes");

// Gather sources of a REGULAR strip, the one the fully asynchronous sweep-2 path can run: every member's
// S and N neighbour faces it with a natural j-row of a local slab (or with a received halo row) in the
// same orientation, so the row can be staged by TMA; members are linked W-E through shared memory; the
// diagonal contributors of interior corners are the ends of the strip mates' S/N neighbour rows.  The two
// strip ends are free: any source (column or row of a local slab, halo row, either orientation) and any
// corner multiplicity / summation order, so strips along face edges and partition cuts qualify too.
// Any other strip runs the generic descriptor-driven path.  Element indices are local.
struct StripReg {
    int32_t regular;          // 0: use the generic path (other fields unused)
    int32_t s_src[4];         // S neighbour of member w: local element, or halo slot when s_c0 == ROW_IN_HALO
    int32_t n_src[4];         // N neighbour of member w
    uint8_t s_c0[4];          // first double of the facing row inside the neighbour's slab (0 or 12), or ROW_IN_HALO
    uint8_t n_c0[4];
    uint8_t corner_n[4][4];   // [member][corner SW,SE,NW,NE]: number of contributors (own included)
    uint8_t corner_perm[4][4];// summation order, as ElemDss::corner_perm
    int32_t end_src[2];       // W neighbour of the first member / E neighbour of the last: local element or halo slot
    uint8_t end_kind[2];      // SRC_TMP (a j-row), SRC_TMPCOL (an i-column) or SRC_HALO
    uint8_t end_row[2];       // row 0/3 (SRC_TMP), side 0=W/1=E (SRC_TMPCOL)
    uint8_t end_rev[2];       // 1 = runs the other way
    uint8_t all_std;          // 1 = every corner has four contributors in the row-major interior order (fast sums)
    uint8_t pad0;
    int32_t diag_src[4];      // SW, NW of the first member, SE, NE of the last: local element / halo slot, -1 = none
    uint8_t diag_kind[4];     // SRC_TMPCOL (local slab, diag_pt = point 0..15) or SRC_HALO (diag_pt = position in the row)
    uint8_t diag_pt[4];
    int32_t first;            // first local element of the strip (the persistent sweep reads the strip from this record alone)
    int32_t count;            // members (1..STRIP_MAX)
    uint8_t pad1[4];
};
static_assert(sizeof(StripReg) == 128, "StripReg must be 128 bytes");
constexpr int ROW_IN_HALO = 255;

struct RowRef {
    int64_t elem; // global element id
    int32_t row;  // ROW_*
    bool operator<(const RowRef &o) const { return elem != o.elem ? elem < o.elem : row < o.row; }
    bool operator==(const RowRef &o) const { return elem == o.elem && row == o.row; }
};

struct Plan {
    int64_t nelem_global = 0, first = 0, count = 0;
    int rank = 0, nranks = 1;
    int64_t unique_nodes = 0;
    std::vector<ElemDss> dss;                // [count]
    std::vector<std::vector<RowRef>> send;   // [nranks] rows this rank sends, wire order
    std::vector<std::vector<RowRef>> recv;   // [nranks] rows this rank receives, wire order
    std::vector<int64_t> recv_offset;        // [nranks+1] first halo slot of each peer
    std::vector<int32_t> boundary_list, interior_list;     // sweep-1 split (feeds a peer / does not)
    std::vector<int32_t> dependent_list, independent_list; // sweep-2 split (needs a halo / does not)
    // sweep-2 work units: W-E strips of up to STRIP_MAX consecutive local elements whose facing
    // edges coincide (e's E edge == (e+1)'s W edge, same orientation), packed first*8 + (count-1),
    // ascending.  A strip's interior W/E edges are exchanged through shared memory.  Strips are a
    // performance grouping only: the kernel re-checks the link and falls back to global loads.
    std::vector<int32_t> strips_all, strips_dependent, strips_independent;
    // one StripReg per strip, same order: the gather sources of a "regular" strip (see StripReg)
    std::vector<StripReg> reg_all, reg_dependent, reg_independent;
    // node groups of the local points: members of point p are pt_member[pt_start[p] .. pt_start[p+1]),
    // global point ids 16*e+pt in ascending order (own point included)
    std::vector<int64_t> pt_start, pt_member;
    std::string error; // non-empty if the mesh is not supported
};

constexpr int STRIP_MAX = 4;
inline int strip_first(int32_t s) { return s >> 3; }
inline int strip_count(int32_t s) { return (s & 7) + 1; }

// ---- grouping of the strip lists for the warp-strip sweep (bwk_plan.cpp; used by bwk_api.cu and by the CPU tests)
// fast = full regular strip whose members' S (and N) neighbours are consecutive elements (or consecutive halo slots) facing with
// the same row: one {4, SL slabs, NM elements} TMA box fetches the members' rows
bool strip_is_fast(const StripReg &r, int32_t strip);
// b continues a: same row of elements (plain W-E link between a's last and b's first member), the members' S and N neighbours
// continue too, and the diagonal contributors at the joint are the facing ends of the other strip's S/N rows
bool strips_chain(const StripReg &a, const StripReg &b);
struct StripGroups {
    std::vector<int32_t> strips;  // the list re-ordered: fast strips first (element order kept in both parts)
    std::vector<StripReg> regs;   // same order
    int64_t nfast = 0;
    std::vector<StripReg> grp[3]; // groups of 1 / 2 / 4 chained fast strips (4 / 8 / 16 members per warp), records back to back
};
StripGroups group_strips(const std::vector<int32_t> &strips, const std::vector<StripReg> &regs, int max_members);

// partition: rank r owns [first, first+count)
void partition(int64_t nelem_global, int rank, int nranks, int64_t *first, int64_t *count);
int owner_of(int64_t nelem_global, int nranks, int64_t e);
void cubed_sphere_gdof(int ne, int64_t *gdof);
// Builds the plan; on unsupported meshes plan.error is set.
Plan build_plan(int64_t nelem_global, const int64_t *gdof, int rank, int nranks);
// Host restatement of the sweep-2 gather for one local element (test hook).
void gather_elem_host(const Plan &pl, int64_t local_e, const double *field_global, double *out16);

// point index of position t (0..3) along edge `edge`
inline int edge_pt(int edge, int t)
{
    switch (edge) {
    case EDGE_S: return t;
    case EDGE_N: return 12 + t;
    case EDGE_W: return 4 * t;
    default: return 3 + 4 * t;
    }
}
inline int edge_rowid(int edge)
{
    switch (edge) {
    case EDGE_S: return ROW_S;
    case EDGE_N: return ROW_N;
    case EDGE_W: return ROW_W;
    default: return ROW_E;
    }
}
inline int rowid_edge(int row) { return row == ROW_S ? EDGE_S : row == ROW_N ? EDGE_N : row == ROW_W ? EDGE_W : EDGE_E; }

} // namespace bwk
