// world_host.h — host side of the World: geometry registry, OBB records, broad-phase tables.
// Product code (no oracle dependency).  Mirrors the reference World's construction path:
//   include/World.h:20-33, src/World.cpp:13-78, src/Object.cpp:11-85, src/OBB.cpp:93-124.
#pragma once

#include <cstdint>
#include <map>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/epp_c_api.h"

namespace epp {

// ---- device-visible layouts ------------------------------------------------------------

// Stored AABB of one R-tree value (src/World.cpp:60-65): 6 doubles, 48 B.
struct AabbRec {
    double mn[3];
    double mx[3];
};

// Narrow-phase record, 18 doubles = 144 B.
struct ObbRec {
    double center[3];
    double rot[9];  // row-major R(r,c)
    double half[3];
    double inflate; // radius of the owning kind (gate / obstacle), src/World.cpp:89-90
    uint32_t flags; // bit0 filling, bit1 gate
    uint32_t pad;
    uint64_t cell_min; // grid broad phase: 21-bit cell index per axis of aabb min (x | y<<21 | z<<42)
};
static_assert(sizeof(AabbRec) == 48, "AabbRec layout");
static_assert(sizeof(ObbRec) == 144, "ObbRec layout");

constexpr int kCellRec = 19;
constexpr uint32_t kFlagFilling = 1u;
constexpr uint32_t kFlagGate = 2u;

// Monotone binning shared by host table construction and device queries.  Monotonicity in x
// (rounded subtract, multiply by a positive constant, floor, clamp) is what makes the broad
// phases conservative: aabb_min < p < aabb_max  =>  bin(aabb_min) <= bin(p) <= bin(aabb_max).
struct AxisBins {
    double lo[3];
    double inv[3];
    int n[3];
};

int host_bin(double x, double lo, double inv, int n);

// Binning of the per-axis bitmask tables (EPP_BP_BITS): one rounded add moves the world's extent into a
// single binade [2^e, 2^(e+1)), where the high word of the double grows linearly with x; the bin is that
// word's offset from the binade's start, shifted down and clamped to [0, n-1].  Integer work after one
// FP64 add, monotone in x for every double (negative sums and NaNs clamp), identical on host and device.
struct KeyBins {
    double off[3];  // t = x + off
    int32_t h0[3];  // high word of 2^e
    int32_t shift[3];
    int32_t n[3];   // rows per axis = inner bins (a power of two) + 1 row beyond the binade
};

int host_key_bin(double x, double off, int32_t h0, int32_t shift, int32_t n);

// Per-axis bin bitmasks (EPP_BP_BITS).  Word w of bin b of axis a at [(a * nb + b) * Ws + w].
struct BitsTables {
    int nb = 0; // bins per axis
    int W = 0;  // 32-bit words per bin = ceil(M / 32)
    int Ws = 0; // row stride in words: 1 for W = 1, else (W + 1) | 1 — W words, a summary word at index W, odd stride
    KeyBins bins{};
    // W > 1: coarse 3-D cells (coarse bin = bin >> cshift) -> which WORDS hold an OBB whose bin range overlaps the cell;
    // appended to `inside` at word offset cmask_off (a multiple of 4) so that one copy brings both on chip.  A query
    // ORs the masks of the coarse cells it covers and looks at those words only (records are Morton-sorted at commit,
    // so the OBBs of a word are neighbours and a cell names one or two words).
    int cmask_off = 0;
    int cmask_words = 0;
    int cn[3] = {1, 1, 1};
    int cshift[3] = {0, 0, 0};
    std::vector<uint32_t> inside; // k set iff bin(min_k) <= b <= bin(max_k)
    // Ray queries: rows indexed by (first bin b0, bin span d < range_d) of the ray box on the x and y axes:
    // k set iff bin(min_k) <= b0 + d and bin(max_k) >= b0 — what max_ge[b0] & min_le[b0 + d] gives, in ONE row whose
    // summary word is selective (the cumulative rows' summaries are all ones).  Row of axis a at
    // ((a * nb + b0) * range_d + d) * Ws.  Boxes spanning more bins take the cumulative rows.
    int range_d = 0;
    std::vector<uint32_t> range;
    std::vector<uint32_t> min_le; // k set iff bin(min_k) <= b
    std::vector<uint32_t> max_ge; // k set iff bin(max_k) >= b
};

// Uniform grid with CSR candidate lists (EPP_BP_GRID).
struct GridTables {
    AxisBins bins{};
    std::vector<uint32_t> cell_start; // ncell + 1
    std::vector<uint32_t> cell_items;
    // Records in CELL order (z-rotated worlds): kCellRec doubles per entry of cell_items, so that a query reads a cell's
    // candidates from consecutive memory right behind the cell_start lookup instead of chasing item -> AabbRec -> ObbRec:
    // [0..5] stored AABB, [6..8] centre, [9..12] R(0,0) R(0,1) R(1,0) R(1,1), [13..15] half sizes, [16] inflate,
    // [17] bits: flags (2) | min cell of the OBB's AABB x, y, z (20 bits each, from bit 2), [18] bits: the cell this
    // copy is listed in, x, y, z (20 bits each).
    std::vector<double> cell_recs;
    // Occupancy bitmap: one bit per COARSE cell (fine cell index >> occ_shift per axis), set iff any fine cell
    // under it has an item.  Sized to live in shared memory (occ_budget_bytes): a unit whose coarse cells are
    // all clear has no candidate and never touches cell_start / cell_items.
    std::vector<uint32_t> occ;
    int occ_shift[3] = {0, 0, 0};
    int occ_n[3] = {1, 1, 1}; // coarse cells per axis; bit index = (z * occ_n[1] + y) * occ_n[0] + x
};

// ---- host world ------------------------------------------------------------------------

struct HostObb {
    double center[3];
    double half[3];
    double rot[9];
    bool filling;
};

// Key "{gate|obstacle}_{id}_obb_{i}" of src/World.cpp:63 as a tuple.
using ObbKey = std::tuple<int /*is_gate*/, int /*id*/, int /*obb index*/>;

struct TreeValue { // pair<box, string>, include/Types.h:13-16
    double mn[3];
    double mx[3];
    ObbKey key;
};

class HostWorld {
public:
    HostWorld(double inflate_gate, double inflate_obstacle)
        : inflate_gate_(inflate_gate), inflate_obstacle_(inflate_obstacle) {}

    int set_gate_geometry(int type_id, const epp_obb_desc *d, int n);
    int set_obstacle_geometry(const epp_obb_desc *d, int n);
    void reset();
    int add_gate(int gate_id, const double pose7[7], bool is_update);
    int add_obstacle(int obstacle_id, const double pose6[6]);

    int size() const { return static_cast<int>(index_.size()); }
    int get(int i, epp_obb *out) const;

    // Flatten the R-tree values (each resolved through the key map, src/World.cpp:87) into
    // device records.
    int flatten(std::vector<AabbRec> &aabbs, std::vector<ObbRec> &obbs) const;

    double inflate_gate() const { return inflate_gate_; }
    double inflate_obstacle() const { return inflate_obstacle_; }

private:
    int build_object(const std::vector<epp_obb_desc> &g, const double pos[3], const double rot[3],
                     std::vector<HostObb> &out) const;
    void add_object(const std::vector<HostObb> &objs, int id, bool is_gate, double inflate);
    int remove_object(int id, bool is_gate, int n_obbs, double inflate);

    double inflate_gate_, inflate_obstacle_;
    std::map<int, std::vector<epp_obb_desc>> gate_geom_;
    std::vector<epp_obb_desc> obstacle_geom_;
    std::map<ObbKey, HostObb> obbs_;  // std::map<std::string, OBB> obbs, include/World.h:112
    std::vector<TreeValue> index_;    // rtree index, include/World.h:109
};

void obb_aabb(const HostObb &o, double inflate, double mn[3], double mx[3]);

// Broad-phase construction over flattened records.
void build_bits(const std::vector<AabbRec> &aabbs, int bins_per_axis, BitsTables &out, int range_d = 5);
// Reorders the records along a Morton curve over the stored-AABB centres (x, y): neighbours in space become
// neighbours in index.  Results are order-independent (any hit invalidates), the broad phases get spatially coherent
// words / cell lists.
void morton_sort(std::vector<AabbRec> &aabbs, std::vector<ObbRec> &obbs);
void build_grid(const std::vector<AabbRec> &aabbs, std::vector<ObbRec> &obbs, long max_cells,
                GridTables &out, long occ_budget_bytes = 72 * 1024);

const char *status_message(int status);

} // namespace epp
