// device_checks.cuh — narrow-phase predicates and broad-phase queries on the device.
//
// Arithmetic contract: binary64 throughout, every operation
// individually rounded through __dadd_rn/__dsub_rn/__dmul_rn/__ddiv_rn (never contracted to
// FMA, independent of compiler flags), 3-term dot products summed left to right,
// std::min/std::max in their comparison form, no fmin/fmax, no transcendentals.
#pragma once

#include <cstdint>

#include "world_host.h"

namespace epp {

enum CheckMode : int {
    kModeBool = 0,       // World::checkPointValidity(point, bool)   src/World.cpp:80-104
    kModeMinDist = 1     // World::checkPointValidity(point, double) src/World.cpp:106-128
};

// Everything a kernel needs to see of one committed world epoch; passed by value.
struct WorldView {
    const AabbRec *aabb;
    const ObbRec *obb;
    int M;
    // EPP_BP_BITS
    const uint32_t *bits_inside;
    const uint32_t *bits_min_le;
    const uint32_t *bits_max_ge;
    // worlds of <= 256 z-rotated OBBs: K2's folded 21-double records (kernels.cu kRayRec) for canPassGate = false and,
    // M records further, for canPassGate = true, built at commit; null: the kernels fold them per block
    const double *ray_recs;
    const uint32_t *bits_range; // BitsTables::range ((first bin, span) rows of x and y); range_d == 0: absent
    int range_d;
    int nb;
    int W;
    int Ws; // row stride of the bits tables in words (W = 1: 1; else (W + 1) | 1 with a summary word at index W)
    KeyBins kb;
    // W > 1: coarse cell -> word mask table behind bits_inside (BitsTables::cmask_*); 0 words: absent
    int cmask_off, cmask_words, cn0, cn1;
    int cshift[3];
    int zrot; // every OBB rotation is a rotation about z (rot[2,5,6,7] == 0, rot[8] == 1)
    // EPP_BP_GRID
    const uint32_t *cell_start;
    const uint32_t *cell_items;
    AxisBins gbins;
    const double *cell_recs; // GridTables::cell_recs (kCellRec doubles per entry of cell_items); null: not built
    // occupancy bitmap over coarse cells (GridTables::occ): bit (z' * occ_n1 + y') * occ_n0 + x', c' = c >> occ_shift
    const uint32_t *occ;
    uint32_t occ_bytes; // multiple of 16
    int occ_n0, occ_n1;
    int occ_shift[3];
};

struct P3 {
    double x, y, z;
};

__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }

// Same function as host_bin() in world_host.cpp (monotone in x).
__device__ __forceinline__ int dev_bin(double x, double lo, double inv, int n)
{
    int b = __double2int_rd(dmul(dsub(x, lo), inv)); // saturating, NaN -> 0
    b = b < 0 ? 0 : b;
    return b > n - 1 ? n - 1 : b;
}

// Same function as host_key_bin() in world_host.cpp: row of the per-axis bitmask tables for coordinate x.
__device__ __forceinline__ int dev_key_bin(double x, double off, int h0, int shift, int n)
{
    int h = __double2hiint(dadd(x, off));
    h = h < h0 ? h0 : h;
    const int b = (h - h0) >> shift;
    return b > n - 1 ? n - 1 : b;
}

// rotation.transpose() * d : local_i = (R(0,i) d0 + R(1,i) d1) + R(2,i) d2   (src/OBB.cpp:66)
__device__ __forceinline__ P3 to_local(const double *__restrict__ rot, double d0, double d1, double d2)
{
    P3 l;
    l.x = dadd(dadd(dmul(rot[0], d0), dmul(rot[3], d1)), dmul(rot[6], d2));
    l.y = dadd(dadd(dmul(rot[1], d0), dmul(rot[4], d1)), dmul(rot[7], d2));
    l.z = dadd(dadd(dmul(rot[2], d0), dmul(rot[5], d1)), dmul(rot[8], d2));
    return l;
}

// Boost within(point, box): strict on all axes (bgi::contains, src/World.cpp:83,109).
__device__ __forceinline__ bool within_strict(const AabbRec *__restrict__ a, const P3 &p)
{
    return p.x > a->mn[0] && p.x < a->mx[0] && p.y > a->mn[1] && p.y < a->mx[1] && p.z > a->mn[2] &&
           p.z < a->mx[2];
}

// OBB::checkCollisionWithPoint (src/OBB.cpp:63-91).  `inflate` is added to the half sizes
// only for collision boxes (shouldBeInflated, include/OBB.h:54-57).
__device__ __forceinline__ bool point_hit(const ObbRec *__restrict__ o, const P3 &p, double inflate)
{
    const P3 l = to_local(o->rot, dsub(p.x, o->center[0]), dsub(p.y, o->center[1]),
                          dsub(p.z, o->center[2]));
    double hx = o->half[0], hy = o->half[1], hz = o->half[2];
    if (!(o->flags & kFlagFilling)) {
        hx = dadd(hx, inflate);
        hy = dadd(hy, inflate);
        hz = dadd(hz, inflate);
    }
    return fabs(l.x) <= hx && fabs(l.y) <= hy && fabs(l.z) <= hz;
}

// One axis of the slab loop (src/OBB.cpp:32-59).  Returns false when the reference returns
// "no collision" from inside the loop.
__device__ __forceinline__ bool slab_axis(double dir, double ls, double h, double &t_min, double &t_max)
{
    const double bmin = -h, bmax = h;
    if (fabs(dir) < 1e-6) {
        if (ls < bmin || ls > bmax) return false;
    } else {
        const double inv = __ddiv_rn(1.0, dir);
        const double t1 = dmul(dsub(bmin, ls), inv);
        const double t2 = dmul(dsub(bmax, ls), inv);
        const double t_entry = (t2 < t1) ? t2 : t1;       // std::min(t1, t2)
        const double t_exit = (t1 < t2) ? t2 : t1;        // std::max(t1, t2)
        t_min = (t_min < t_entry) ? t_entry : t_min;      // std::max(tMin, tEntry)
        t_max = (t_exit < t_max) ? t_exit : t_max;        // std::min(tMax, tExit)
        if (t_min > t_max) return false;
    }
    return true;
}

// OBB::checkCollisionWithRay (src/OBB.cpp:10-61).
__device__ __forceinline__ bool ray_hit(const ObbRec *__restrict__ o, const P3 &a, const P3 &b, double inflate)
{
    if (point_hit(o, a, inflate) || point_hit(o, b, inflate)) return true; // :13-18
    const P3 ls = to_local(o->rot, dsub(a.x, o->center[0]), dsub(a.y, o->center[1]),
                           dsub(a.z, o->center[2])); // :21
    const P3 le = to_local(o->rot, dsub(b.x, o->center[0]), dsub(b.y, o->center[1]),
                           dsub(b.z, o->center[2])); // :22
    double t_min = 0.0, t_max = 1.0;
    // half sizes inflated for every type (:28)
    if (!slab_axis(dsub(le.x, ls.x), ls.x, dadd(o->half[0], inflate), t_min, t_max)) return false;
    if (!slab_axis(dsub(le.y, ls.y), ls.y, dadd(o->half[1], inflate), t_min, t_max)) return false;
    if (!slab_axis(dsub(le.z, ls.z), ls.z, dadd(o->half[2], inflate), t_min, t_max)) return false;
    return 0.0 <= t_min && t_min <= 1.0 && 0.0 <= t_max && t_max <= 1.0; // :60
}

// Full per-candidate point test of World::checkPointValidity (both overloads), after the
// R-tree query: returns true when candidate k makes the point INVALID.
__device__ __forceinline__ bool point_candidate_invalid(const AabbRec *__restrict__ aabb,
                                                        const ObbRec *__restrict__ obb, int k, const P3 &p,
                                                        int mode, bool can_pass_gate, double min_distance)
{
    if (!within_strict(aabb + k, p)) return false;
    const ObbRec *o = obb + k;
    const bool filling = o->flags & kFlagFilling;
    if (mode == kModeBool) {
        if (filling && can_pass_gate) return false;           // src/World.cpp:92-95
        return point_hit(o, p, o->inflate);                   // :97
    }
    if (filling) return false;                                // :116-119
    return point_hit(o, p, min_distance);                     // :121
}

// Closed box/box intersection of Boost intersects(box, box) (src/World.cpp:143).
__device__ __forceinline__ bool boxes_touch(const AabbRec *__restrict__ a, const P3 &mn, const P3 &mx)
{
    return !(mx.x < a->mn[0] || mn.x > a->mx[0] || mx.y < a->mn[1] || mn.y > a->mx[1] ||
             mx.z < a->mn[2] || mn.z > a->mx[2]);
}

__device__ __forceinline__ bool ray_candidate_invalid(const AabbRec *__restrict__ aabb,
                                                      const ObbRec *__restrict__ obb, int k, const P3 &a,
                                                      const P3 &b, const P3 &mn, const P3 &mx,
                                                      bool can_pass_gate)
{
    if (!boxes_touch(aabb + k, mn, mx)) return false;
    const ObbRec *o = obb + k;
    if ((o->flags & kFlagFilling) && can_pass_gate) return false; // src/World.cpp:150-153
    return ray_hit(o, a, b, o->inflate);                          // :157
}

// rowwise().minCoeff()/maxCoeff() over the two endpoints (src/World.cpp:134-139).
__device__ __forceinline__ void ray_box(const P3 &a, const P3 &b, P3 &mn, P3 &mx)
{
    mn.x = (b.x < a.x) ? b.x : a.x;
    mx.x = (b.x > a.x) ? b.x : a.x;
    mn.y = (b.y < a.y) ? b.y : a.y;
    mx.y = (b.y > a.y) ? b.y : a.y;
    mn.z = (b.z < a.z) ? b.z : a.z;
    mx.z = (b.z > a.z) ? b.z : a.z;
}

} // namespace epp
