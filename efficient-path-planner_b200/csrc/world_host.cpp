// world_host.cpp — host-side World construction and broad-phase tables (product code).
// Compiled with -ffp-contract=off: every operation is an individually rounded binary64 op,
// 3-term dot products are summed left to right (SURVEY.md §8a note E).
#include "world_host.h"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace epp {

namespace {

struct V3 {
    double v[3];
};

// Eigen fixed-size product M * v, coefficient based: out_r = (M(r,0) v0 + M(r,1) v1) + M(r,2) v2.
V3 mul(const double m[9], const V3 &x)
{
    V3 o;
    for (int r = 0; r < 3; ++r) {
        double s = m[3 * r] * x.v[0];
        s = s + m[3 * r + 1] * x.v[1];
        s = s + m[3 * r + 2] * x.v[2];
        o.v[r] = s;
    }
    return o;
}

void mul33(const double a[9], const double b[9], double out[9])
{
    double t[9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
            double s = a[3 * r] * b[c];
            s = s + a[3 * r + 1] * b[3 + c];
            s = s + a[3 * r + 2] * b[6 + c];
            t[3 * r + c] = s;
        }
    std::memcpy(out, t, sizeof t);
}

} // namespace

const char *status_message(int status)
{
    switch (status) {
    case EPP_OK: return "ok";
    case EPP_ERR_ROT_X: return "Rotation around x axis is not supported";   // src/Object.cpp:41
    case EPP_ERR_ROT_Y: return "Rotation around y axis is not supported";   // src/Object.cpp:46
    case EPP_ERR_CENTER_Z: return "Center z position must be zero";         // src/Object.cpp:19
    case EPP_ERR_NO_GEOMETRY: return "no geometry registered for this component / gate type";
    case EPP_ERR_NO_SUCH_OBB: return "map::at";                             // src/World.cpp:74
    case EPP_ERR_INVALID_ARG: return "invalid argument";
    case EPP_ERR_NOT_COMMITTED: return "world has uncommitted changes or was never committed";
    case EPP_ERR_CUDA: return "CUDA error";
    case EPP_ERR_NO_DEVICE: return "no CUDA device available (there is no CPU fallback)";
    default: return "unknown status";
    }
}

int HostWorld::set_gate_geometry(int type_id, const epp_obb_desc *d, int n)
{
    if (n < 0 || (n > 0 && !d)) return EPP_ERR_INVALID_ARG;
    gate_geom_[type_id].assign(d, d + n);
    return EPP_OK;
}

int HostWorld::set_obstacle_geometry(const epp_obb_desc *d, int n)
{
    if (n < 0 || (n > 0 && !d)) return EPP_ERR_INVALID_ARG;
    obstacle_geom_.assign(d, d + n);
    return EPP_OK;
}

void HostWorld::reset() // include/World.h:29-33
{
    index_.clear();
    obbs_.clear();
}

// Object::createFromDescription -> Object::Object -> translate -> rotateZ (src/Object.cpp:11-85).
int HostWorld::build_object(const std::vector<epp_obb_desc> &g, const double pos[3],
                            const double rot[3], std::vector<HostObb> &out) const
{
    if (std::fabs(rot[0]) > 1e-6) return EPP_ERR_ROT_X;
    if (std::fabs(rot[1]) > 1e-6) return EPP_ERR_ROT_Y;
    if (pos[2] > 1e-6) return EPP_ERR_CENTER_Z;

    V3 global_center{{0.0, 0.0, 0.0}};
    for (int r = 0; r < 3; ++r) global_center.v[r] += pos[r];

    const double c = std::cos(rot[2]);
    const double s = std::sin(rot[2]);
    const double rz[9] = {c, -s, 0.0, s, c, 0.0, 0.0, 0.0, 1.0};

    out.clear();
    out.reserve(g.size());
    for (const epp_obb_desc &d : g) {
        HostObb o;
        for (int r = 0; r < 3; ++r) {
            o.center[r] = d.position[r];
            o.half[r] = d.size[r] / 2; // src/ConfigParserYAML.cpp:63
        }
        const double ident[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; // include/OBB.h:49
        std::memcpy(o.rot, ident, sizeof ident);
        o.filling = d.filling != 0;

        for (int r = 0; r < 3; ++r) o.center[r] += pos[r];
        V3 rel;
        for (int r = 0; r < 3; ++r) rel.v[r] = o.center[r] - global_center.v[r];
        V3 turned = mul(rz, rel);
        for (int r = 0; r < 3; ++r) o.center[r] = turned.v[r] + global_center.v[r];
        mul33(rz, o.rot, o.rot);
        out.push_back(o);
    }
    return EPP_OK;
}

// OBB::getAABB (src/OBB.cpp:93-124): extent of the 8 rotated corners of the UN-inflated box,
// then +- inflate for collision boxes only.
void obb_aabb(const HostObb &o, double inflate, double mn[3], double mx[3])
{
    for (int c = 0; c < 8; ++c) {
        const double sx = ((c & 3) == 1 || (c & 3) == 2) ? 1.0 : -1.0; // -1 1 1 -1 | -1 1 1 -1
        const double sy = (c & 2) ? 1.0 : -1.0;                         // -1 -1 1 1 | ...
        const double sz = (c & 4) ? 1.0 : -1.0;
        V3 corner{{sx * o.half[0], sy * o.half[1], sz * o.half[2]}};
        V3 g = mul(o.rot, corner);
        for (int r = 0; r < 3; ++r) {
            const double v = g.v[r] + o.center[r];
            if (c == 0) {
                mn[r] = v;
                mx[r] = v;
            } else {
                mn[r] = std::min(mn[r], v);
                mx[r] = std::max(mx[r], v);
            }
        }
    }
    if (!o.filling) {
        for (int r = 0; r < 3; ++r) {
            mn[r] -= inflate;
            mx[r] += inflate;
        }
    }
}

// World::addObject (src/World.cpp:57-67): map::insert keeps an existing key's OBB, the
// R-tree always receives the new (aabb, key) value.
void HostWorld::add_object(const std::vector<HostObb> &objs, int id, bool is_gate, double inflate)
{
    for (size_t i = 0; i < objs.size(); ++i) {
        TreeValue tv;
        obb_aabb(objs[i], inflate, tv.mn, tv.mx);
        tv.key = ObbKey(is_gate ? 1 : 0, id, static_cast<int>(i));
        obbs_.insert({tv.key, objs[i]});
        index_.push_back(tv);
    }
}

// World::removeObject (src/World.cpp:69-78).
int HostWorld::remove_object(int id, bool is_gate, int n_obbs, double inflate)
{
    for (int i = 0; i < n_obbs; ++i) {
        ObbKey key(is_gate ? 1 : 0, id, i);
        auto it = obbs_.find(key);
        if (it == obbs_.end()) return EPP_ERR_NO_SUCH_OBB;
        double mn[3], mx[3];
        obb_aabb(it->second, inflate, mn, mx);
        for (size_t t = 0; t < index_.size(); ++t) {
            const TreeValue &tv = index_[t];
            if (tv.key == key && std::equal(mn, mn + 3, tv.mn) && std::equal(mx, mx + 3, tv.mx)) {
                index_.erase(index_.begin() + static_cast<long>(t));
                break;
            }
        }
        obbs_.erase(it);
    }
    return EPP_OK;
}

// World::addGatePrivateOperation (src/World.cpp:13-35).
int HostWorld::add_gate(int gate_id, const double pose7[7], bool is_update)
{
    const double pos[3] = {pose7[0], pose7[1], 0.0};
    const double rot[3] = {pose7[3], pose7[4], pose7[5]};
    const int type = static_cast<int>(pose7[6]);
    auto g = gate_geom_.find(type);
    if (g == gate_geom_.end() || g->second.empty()) return EPP_ERR_NO_GEOMETRY;
    std::vector<HostObb> objs;
    int rc = build_object(g->second, pos, rot, objs);
    if (rc != EPP_OK) return rc;
    if (is_update) {
        rc = remove_object(gate_id, true, static_cast<int>(objs.size()), inflate_gate_);
        if (rc != EPP_OK) return rc;
    }
    add_object(objs, gate_id, true, inflate_gate_);
    return EPP_OK;
}

// World::addObstacle (src/World.cpp:47-55).
int HostWorld::add_obstacle(int obstacle_id, const double pose6[6])
{
    const double pos[3] = {pose6[0], pose6[1], pose6[2]};
    const double rot[3] = {pose6[3], pose6[4], pose6[5]};
    if (obstacle_geom_.empty()) return EPP_ERR_NO_GEOMETRY;
    std::vector<HostObb> objs;
    int rc = build_object(obstacle_geom_, pos, rot, objs);
    if (rc != EPP_OK) return rc;
    add_object(objs, obstacle_id, false, inflate_obstacle_);
    return EPP_OK;
}

int HostWorld::get(int i, epp_obb *out) const
{
    if (i < 0 || i >= size() || !out) return EPP_ERR_INVALID_ARG;
    const TreeValue &tv = index_[static_cast<size_t>(i)];
    auto it = obbs_.find(tv.key);
    if (it == obbs_.end()) return EPP_ERR_NO_SUCH_OBB;
    const HostObb &o = it->second;
    std::memcpy(out->center, o.center, sizeof o.center);
    std::memcpy(out->half, o.half, sizeof o.half);
    std::memcpy(out->rot, o.rot, sizeof o.rot);
    std::memcpy(out->aabb_min, tv.mn, sizeof tv.mn);
    std::memcpy(out->aabb_max, tv.mx, sizeof tv.mx);
    out->is_gate = std::get<0>(tv.key);
    out->object_id = std::get<1>(tv.key);
    out->obb_index = std::get<2>(tv.key);
    out->inflate = out->is_gate ? inflate_gate_ : inflate_obstacle_;
    out->filling = o.filling ? 1 : 0;
    return EPP_OK;
}

int HostWorld::flatten(std::vector<AabbRec> &aabbs, std::vector<ObbRec> &obbs) const
{
    aabbs.resize(index_.size());
    obbs.resize(index_.size());
    for (size_t i = 0; i < index_.size(); ++i) {
        epp_obb o;
        int rc = get(static_cast<int>(i), &o);
        if (rc != EPP_OK) return rc;
        std::memcpy(aabbs[i].mn, o.aabb_min, sizeof o.aabb_min);
        std::memcpy(aabbs[i].mx, o.aabb_max, sizeof o.aabb_max);
        ObbRec &r = obbs[i];
        std::memcpy(r.center, o.center, sizeof o.center);
        std::memcpy(r.rot, o.rot, sizeof o.rot);
        std::memcpy(r.half, o.half, sizeof o.half);
        r.inflate = o.inflate;
        r.flags = (o.filling ? kFlagFilling : 0u) | (o.is_gate ? kFlagGate : 0u);
        r.pad = 0;
        r.cell_min = 0;
    }
    return EPP_OK;
}

// ---- broad phase -----------------------------------------------------------------------

int host_bin(double x, double lo, double inv, int n)
{
    const double d = x - lo;
    const double t = d * inv;
    const double f = std::floor(t);
    if (!(f > 0.0)) return 0; // negative, zero and NaN
    if (f >= static_cast<double>(n - 1)) return n - 1;
    return static_cast<int>(f);
}

int host_key_bin(double x, double off, int32_t h0, int32_t shift, int32_t n)
{
    const double t = x + off;
    uint64_t bits;
    std::memcpy(&bits, &t, sizeof bits);
    int32_t h = static_cast<int32_t>(bits >> 32);
    if (h < h0) h = h0;
    const int32_t b = (h - h0) >> shift;
    return b > n - 1 ? n - 1 : b;
}

namespace {

void world_extent(const std::vector<AabbRec> &aabbs, double lo[3], double hi[3])
{
    for (int a = 0; a < 3; ++a) {
        lo[a] = 0.0;
        hi[a] = 1.0;
    }
    for (size_t k = 0; k < aabbs.size(); ++k)
        for (int a = 0; a < 3; ++a) {
            if (k == 0 || aabbs[k].mn[a] < lo[a]) lo[a] = aabbs[k].mn[a];
            if (k == 0 || aabbs[k].mx[a] > hi[a]) hi[a] = aabbs[k].mx[a];
        }
}

// n bins: bin 0 = (-inf, lo), bins 1..n-2 tile [lo, hi], bin n-1 = beyond hi.
void make_axis(double lo, double hi, int n, double &out_lo, double &out_inv)
{
    double span = hi - lo;
    if (!(span > 0.0) || !std::isfinite(span)) span = 1.0;
    const int inner = std::max(1, n - 2);
    const double h = span / inner;
    out_lo = lo - h;
    out_inv = 1.0 / h;
}

// Key binning of one axis: `inner` (power of two) bins over the binade that holds [lo, hi] after the shift.
void make_key_axis(double lo, double hi, int inner, double &off, int32_t &h0, int32_t &shift, int32_t &n)
{
    double span = hi - lo;
    if (!(span > 0.0) || !std::isfinite(span)) span = 1.0;
    int ex = 0;
    const double m = std::frexp(span, &ex); // span = m * 2^ex, m in [0.5, 1)
    int e = (m == 0.5) ? ex - 1 : ex;       // smallest e with 2^e >= span
    e = std::min(std::max(e, -200), 200);
    if (!std::isfinite(lo)) lo = 0.0;
    off = std::ldexp(1.0, e) - lo;
    h0 = static_cast<int32_t>((1023 + e) << 20);
    int lg = 0;
    while ((1 << (lg + 1)) <= inner && lg < 19) ++lg;
    shift = 20 - lg;
    n = (1 << lg) + 1;
}

} // namespace

void build_bits(const std::vector<AabbRec> &aabbs, int bins_per_axis, BitsTables &out, int range_d)
{
    const int M = static_cast<int>(aabbs.size());
    out.W = std::max(1, (M + 31) / 32);
    // Row = W words + (W > 1) one SUMMARY word at index W (bit w set iff word w of the row is non-zero: point queries AND
    // three summaries instead of scanning W words), padded to an odd stride (rows of unrelated bins in different banks).
    out.Ws = out.W == 1 ? 1 : ((out.W + 1) | 1);
    double lo[3], hi[3];
    world_extent(aabbs, lo, hi);
    for (int a = 0; a < 3; ++a)
        make_key_axis(lo[a], hi[a], std::max(2, bins_per_axis), out.bins.off[a], out.bins.h0[a], out.bins.shift[a],
                      out.bins.n[a]);
    out.nb = out.bins.n[0]; // the same power of two + 1 on every axis
    const size_t words = static_cast<size_t>(3) * out.nb * out.Ws;
    out.inside.assign(words, 0u);
    out.min_le.assign(words, 0u);
    out.max_ge.assign(words, 0u);
    for (int k = 0; k < M; ++k)
        for (int a = 0; a < 3; ++a) {
            const int b0 = host_key_bin(aabbs[k].mn[a], out.bins.off[a], out.bins.h0[a], out.bins.shift[a], out.nb);
            const int b1 = host_key_bin(aabbs[k].mx[a], out.bins.off[a], out.bins.h0[a], out.bins.shift[a], out.nb);
            const uint32_t bit = 1u << (k & 31);
            const size_t w = static_cast<size_t>(k >> 5);
            for (int b = 0; b < out.nb; ++b) {
                const size_t at = (static_cast<size_t>(a) * out.nb + b) * out.Ws + w;
                if (b0 <= b && b <= b1) out.inside[at] |= bit;
                if (b0 <= b) out.min_le[at] |= bit;
                if (b1 >= b) out.max_ge[at] |= bit;
            }
        }
    if (out.W > 1)
        for (size_t row = 0; row < static_cast<size_t>(3) * out.nb; ++row)
            for (std::vector<uint32_t> *t : {&out.inside, &out.min_le, &out.max_ge}) {
                uint32_t summary = 0u;
                for (int w = 0; w < out.W; ++w)
                    if ((*t)[row * out.Ws + w]) summary |= 1u << w;
                (*t)[row * out.Ws + out.W] = summary;
            }
    // (first bin, span) rows of the x and y axes for ray boxes
    out.range_d = range_d > 0 ? std::min(range_d, 16) : 0;
    out.range.clear();
    if (out.range_d > 0) {
        const int D = out.range_d;
        out.range.assign(static_cast<size_t>(2) * out.nb * D * out.Ws + 4, 0u);
        for (int k = 0; k < M; ++k)
            for (int a = 0; a < 2; ++a) {
                const int k0 = host_key_bin(aabbs[k].mn[a], out.bins.off[a], out.bins.h0[a], out.bins.shift[a], out.nb);
                const int k1 = host_key_bin(aabbs[k].mx[a], out.bins.off[a], out.bins.h0[a], out.bins.shift[a], out.nb);
                for (int b0 = 0; b0 <= k1 && b0 < out.nb; ++b0)
                    for (int d = 0; d < D; ++d) {
                        if (k0 > b0 + d) continue;
                        out.range[((static_cast<size_t>(a) * out.nb + b0) * D + d) * out.Ws + (k >> 5)] |= 1u << (k & 31);
                    }
            }
        if (out.W > 1)
            for (size_t row = 0; row < static_cast<size_t>(2) * out.nb * D; ++row) {
                uint32_t summary = 0u;
                for (int w = 0; w < out.W; ++w)
                    if (out.range[row * out.Ws + w]) summary |= 1u << w;
                out.range[row * out.Ws + out.W] = summary;
            }
    }
    // coarse cell -> word mask (W > 1 only), appended to `inside`
    out.cmask_off = out.cmask_words = 0;
    if (out.W > 1) {
        for (int a = 0; a < 3; ++a) {
            const int target = a == 2 ? 4 : 32; // the planner's worlds are flat
            out.cshift[a] = 0;
            while (((out.nb - 1) >> out.cshift[a]) + 1 > target + 1) ++out.cshift[a];
            out.cn[a] = ((out.nb - 1) >> out.cshift[a]) + 1;
        }
        out.cmask_off = static_cast<int>((words + 3) / 4 * 4);
        out.cmask_words = out.cn[0] * out.cn[1] * out.cn[2];
        out.inside.resize(static_cast<size_t>(out.cmask_off) + out.cmask_words + 4, 0u);
        uint32_t *cm = out.inside.data() + out.cmask_off;
        for (int k = 0; k < M; ++k) {
            int c0[3], c1[3];
            for (int a = 0; a < 3; ++a) {
                c0[a] = host_key_bin(aabbs[k].mn[a], out.bins.off[a], out.bins.h0[a], out.bins.shift[a], out.nb) >> out.cshift[a];
                c1[a] = host_key_bin(aabbs[k].mx[a], out.bins.off[a], out.bins.h0[a], out.bins.shift[a], out.nb) >> out.cshift[a];
            }
            for (int z = c0[2]; z <= c1[2]; ++z)
                for (int y = c0[1]; y <= c1[1]; ++y)
                    for (int x = c0[0]; x <= c1[0]; ++x)
                        cm[(static_cast<size_t>(z) * out.cn[1] + y) * out.cn[0] + x] |= 1u << (k >> 5);
        }
    }
}

void morton_sort(std::vector<AabbRec> &aabbs, std::vector<ObbRec> &obbs)
{
    const size_t M = aabbs.size();
    if (M < 2) return;
    double lo[3], hi[3];
    world_extent(aabbs, lo, hi);
    auto spread = [](uint32_t v) { // 16 bits -> every second bit
        uint64_t x = v & 0xffffu;
        x = (x | (x << 8)) & 0x00ff00ffull;
        x = (x | (x << 4)) & 0x0f0f0f0full;
        x = (x | (x << 2)) & 0x33333333ull;
        x = (x | (x << 1)) & 0x55555555ull;
        return x;
    };
    std::vector<std::pair<uint64_t, uint32_t>> key(M);
    for (size_t k = 0; k < M; ++k) {
        uint32_t q[2];
        for (int a = 0; a < 2; ++a) {
            const double span = hi[a] - lo[a];
            double u = span > 0.0 ? (0.5 * (aabbs[k].mn[a] + aabbs[k].mx[a]) - lo[a]) / span : 0.0;
            if (!(u >= 0.0)) u = 0.0;
            if (u > 1.0) u = 1.0;
            q[a] = static_cast<uint32_t>(u * 65535.0);
        }
        key[k] = {spread(q[0]) | (spread(q[1]) << 1), static_cast<uint32_t>(k)};
    }
    std::stable_sort(key.begin(), key.end());
    std::vector<AabbRec> a2(M);
    std::vector<ObbRec> o2(M);
    for (size_t k = 0; k < M; ++k) {
        a2[k] = aabbs[key[k].second];
        o2[k] = obbs[key[k].second];
    }
    aabbs.swap(a2);
    obbs.swap(o2);
}

void build_grid(const std::vector<AabbRec> &aabbs, std::vector<ObbRec> &obbs, long max_cells,
                GridTables &out, long occ_budget_bytes)
{
    const size_t M = aabbs.size();
    double lo[3], hi[3];
    world_extent(aabbs, lo, hi);

    // Cell edge ~ the mean stored-AABB extent (largest axis mean), so a box spans ~2 cells/axis.
    double mean_ext[3] = {0, 0, 0};
    for (size_t k = 0; k < M; ++k)
        for (int a = 0; a < 3; ++a) mean_ext[a] += aabbs[k].mx[a] - aabbs[k].mn[a];
    double edge = 0.0;
    for (int a = 0; a < 3; ++a) edge = std::max(edge, M ? mean_ext[a] / static_cast<double>(M) : 1.0);
    if (!(edge > 0.0) || !std::isfinite(edge)) edge = 1.0;

    int n[3];
    for (;;) {
        double cells = 1.0;
        for (int a = 0; a < 3; ++a) {
            double span = hi[a] - lo[a];
            if (!(span > 0.0) || !std::isfinite(span)) span = edge;
            double inner = std::ceil(span / edge);
            inner = std::min(std::max(inner, 1.0), 1048574.0);
            n[a] = static_cast<int>(inner) + 2;
            cells *= n[a];
        }
        if (cells <= static_cast<double>(max_cells)) break;
        edge *= 1.26; // ~ halves the cell count per step
    }
    for (int a = 0; a < 3; ++a) {
        make_axis(lo[a], hi[a], n[a], out.bins.lo[a], out.bins.inv[a]);
        out.bins.n[a] = n[a];
    }
    const size_t ncell = static_cast<size_t>(n[0]) * n[1] * n[2];
    out.cell_start.assign(ncell + 1, 0u);

    std::vector<int> b0(3 * M), b1(3 * M);
    for (size_t k = 0; k < M; ++k) {
        for (int a = 0; a < 3; ++a) {
            b0[3 * k + a] = host_bin(aabbs[k].mn[a], out.bins.lo[a], out.bins.inv[a], n[a]);
            b1[3 * k + a] = host_bin(aabbs[k].mx[a], out.bins.lo[a], out.bins.inv[a], n[a]);
        }
        obbs[k].cell_min = static_cast<uint64_t>(b0[3 * k]) | (static_cast<uint64_t>(b0[3 * k + 1]) << 21) |
                           (static_cast<uint64_t>(b0[3 * k + 2]) << 42);
    }
    auto cell_id = [&](int x, int y, int z) {
        return (static_cast<size_t>(z) * n[1] + y) * n[0] + x;
    };
    for (size_t k = 0; k < M; ++k)
        for (int z = b0[3 * k + 2]; z <= b1[3 * k + 2]; ++z)
            for (int y = b0[3 * k + 1]; y <= b1[3 * k + 1]; ++y)
                for (int x = b0[3 * k]; x <= b1[3 * k]; ++x) out.cell_start[cell_id(x, y, z) + 1]++;
    for (size_t c = 0; c < ncell; ++c) out.cell_start[c + 1] += out.cell_start[c];
    out.cell_items.assign(out.cell_start[ncell], 0u);
    std::vector<uint32_t> cursor(out.cell_start.begin(), out.cell_start.end() - 1);
    for (size_t k = 0; k < M; ++k)
        for (int z = b0[3 * k + 2]; z <= b1[3 * k + 2]; ++z)
            for (int y = b0[3 * k + 1]; y <= b1[3 * k + 1]; ++y)
                for (int x = b0[3 * k]; x <= b1[3 * k]; ++x)
                    out.cell_items[cursor[cell_id(x, y, z)]++] = static_cast<uint32_t>(k);

    // Records in cell order (only meaningful when every rotation is about z and cells fit 20 bits per axis)
    out.cell_recs.clear();
    bool zrot = n[0] < (1 << 20) && n[1] < (1 << 20) && n[2] < (1 << 20);
    for (size_t k = 0; k < M && zrot; ++k) {
        const double *r = obbs[k].rot;
        if (r[2] != 0.0 || r[5] != 0.0 || r[6] != 0.0 || r[7] != 0.0 || r[8] != 1.0) zrot = false;
    }
    if (zrot) {
        out.cell_recs.assign(out.cell_items.size() * static_cast<size_t>(kCellRec), 0.0);
        for (int z = 0; z < n[2]; ++z)
            for (int y = 0; y < n[1]; ++y)
                for (int x = 0; x < n[0]; ++x) {
                    const size_t c = cell_id(x, y, z);
                    for (uint32_t j = out.cell_start[c]; j < out.cell_start[c + 1]; ++j) {
                        const size_t k = out.cell_items[j];
                        double *rec = out.cell_recs.data() + static_cast<size_t>(j) * kCellRec;
                        for (int a = 0; a < 3; ++a) {
                            rec[a] = aabbs[k].mn[a];
                            rec[3 + a] = aabbs[k].mx[a];
                            rec[6 + a] = obbs[k].center[a];
                            rec[13 + a] = obbs[k].half[a];
                        }
                        rec[9] = obbs[k].rot[0];
                        rec[10] = obbs[k].rot[1];
                        rec[11] = obbs[k].rot[3];
                        rec[12] = obbs[k].rot[4];
                        rec[16] = obbs[k].inflate;
                        const uint64_t own = static_cast<uint64_t>(obbs[k].flags & 3u) | (static_cast<uint64_t>(b0[3 * k]) << 2) |
                                             (static_cast<uint64_t>(b0[3 * k + 1]) << 22) | (static_cast<uint64_t>(b0[3 * k + 2]) << 42);
                        const uint64_t here = static_cast<uint64_t>(x) | (static_cast<uint64_t>(y) << 20) | (static_cast<uint64_t>(z) << 40);
                        std::memcpy(rec + 17, &own, 8);
                        std::memcpy(rec + 18, &here, 8);
                    }
                }
    }

    // Occupancy bitmap over coarsened cells: fold z first (worlds of this planner are flat: 400 x 400 x 4 m), then
    // x and y together, until the map fits its shared-memory budget.
    for (int a = 0; a < 3; ++a) {
        out.occ_shift[a] = 0;
        out.occ_n[a] = n[a];
    }
    auto occ_bits = [&]() { return static_cast<double>(out.occ_n[0]) * out.occ_n[1] * out.occ_n[2]; };
    auto coarsen = [&](int a) {
        ++out.occ_shift[a];
        out.occ_n[a] = ((n[a] - 1) >> out.occ_shift[a]) + 1;
    };
    const double budget_bits = 8.0 * static_cast<double>(std::max(16l, occ_budget_bytes));
    while (occ_bits() > budget_bits) {
        if (out.occ_n[2] > 2) coarsen(2);
        else if (out.occ_n[0] > 1 || out.occ_n[1] > 1) {
            coarsen(0);
            coarsen(1);
        } else coarsen(2);
    }
    const size_t bits = static_cast<size_t>(occ_bits());
    out.occ.assign((bits + 31) / 32 + 4, 0u); // + slack: range queries read one word past the last bit
    for (int z = 0; z < n[2]; ++z)
        for (int y = 0; y < n[1]; ++y)
            for (int x = 0; x < n[0]; ++x) {
                const size_t c = cell_id(x, y, z);
                if (out.cell_start[c + 1] == out.cell_start[c]) continue;
                const size_t b = (static_cast<size_t>(z >> out.occ_shift[2]) * out.occ_n[1] + (y >> out.occ_shift[1])) *
                                     out.occ_n[0] + (x >> out.occ_shift[0]);
                out.occ[b >> 5] |= 1u << (b & 31);
            }
}

} // namespace epp
