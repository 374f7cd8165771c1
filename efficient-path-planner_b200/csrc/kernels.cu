// kernels.cu — sm_100a kernels of the validity-check hot path and their launchers.
//
// K1  k_states_warp   World::checkPointValidity (both overloads) over an n x 3 AoS batch (k_states: small batches)
// K2  k_segments      World::checkRayValid over two n x 3 AoS batches
// K3a k_traj_samples  PathPlanner::checkTrajectoryValidity over column-major matrices
// K3b k_poly_traj     evaluateRange + Horner + checkTrajectoryValidity fused
// plus generators / reductions used by the multi-GPU and benchmark plumbing.
//
// One thread per unit; a warp's 32 verdicts are packed with __ballot_sync into one mask
// word.  Broad phase is a template parameter (brute / per-axis bitmasks / grid); all three
// produce identical masks because every candidate still takes the reference's exact
// stored-AABB test before the OBB test (DESIGN.md "Broad phase").
#include <cuda_runtime.h>

#include <climits>
#include <cstdint>

#include "device_checks.cuh"
#include "kernels.h"

namespace epp {

namespace {

constexpr int kBlock = 256;
constexpr int kWarpsPerBlock = kBlock / 32; // direct kernels

extern __shared__ __align__(16) unsigned char g_smem[];

__device__ __forceinline__ void block_copy_words(uint32_t *dst, const uint32_t *__restrict__ src, int words)
{
    for (int i = threadIdx.x; i < words; i += blockDim.x) dst[i] = __ldg(src + i);
}

// ---- point queries -----------------------------------------------------------------------

template <int BP>
struct PointWorld;

// Brute force: stored AABBs staged in shared memory when they fit (smem_aabb != nullptr).
template <>
struct PointWorld<EPP_BP_BRUTE> {
    const AabbRec *aabb;
    __device__ void init(const WorldView &wv, bool aabb_in_smem)
    {
        aabb = wv.aabb;
        if (aabb_in_smem) {
            uint32_t *dst = reinterpret_cast<uint32_t *>(g_smem);
            block_copy_words(dst, reinterpret_cast<const uint32_t *>(wv.aabb), wv.M * (int)(sizeof(AabbRec) / 4));
            __syncthreads();
            aabb = reinterpret_cast<const AabbRec *>(g_smem);
        }
    }
    // Streaming form: the table arrives by bulk copy; table() names its source, bind() points at the copy.
    __device__ static uint32_t table(const WorldView &wv, bool aabb_in_smem, const void *&src)
    {
        src = wv.aabb;
        return aabb_in_smem ? (uint32_t)(wv.M * sizeof(AabbRec)) : 0u;
    }
    __device__ void bind(const WorldView &wv, bool aabb_in_smem)
    {
        aabb = aabb_in_smem ? reinterpret_cast<const AabbRec *>(g_smem) : wv.aabb;
    }
    // any(): true iff test(k) holds for some broad-phase candidate k of p (early exit).
    template <class F>
    __device__ __forceinline__ bool any(const WorldView &wv, const P3 &p, F test) const
    {
        for (int k = 0; k < wv.M; ++k) {
            if (!within_strict(aabb + k, p)) continue;
            if (test(k)) return true;
        }
        return false;
    }
    __device__ __forceinline__ bool has_candidates(const WorldView &wv, const P3 &p) const
    {
        return any(wv, p, [](int) { return true; });
    }
    __device__ __forceinline__ bool invalid(const WorldView &wv, const P3 &p, int mode, bool cpg, double md) const
    {
        return any(wv, p, [&](int k) { return point_candidate_invalid(wv.aabb, wv.obb, k, p, mode, cpg, md); });
    }
};

// Per-axis bin bitmasks: candidates = inside_x[bin(x)] & inside_y[bin(y)] & inside_z[bin(z)].
// W > 1: the words worth a look come from the coarse-cell table behind the rows (one read: the cell of p names the
// words that hold an OBB near it; records are Morton-sorted, so that is one word or two).
__device__ __forceinline__ int bits_words(const WorldView &wv) { return wv.cmask_words ? wv.cmask_off + wv.cmask_words : 3 * wv.nb * wv.Ws; }

template <>
struct PointWorld<EPP_BP_BITS> {
    const uint32_t *inside;
    __device__ void init(const WorldView &wv, bool)
    {
        uint32_t *dst = reinterpret_cast<uint32_t *>(g_smem);
        block_copy_words(dst, wv.bits_inside, bits_words(wv));
        __syncthreads();
        inside = dst;
    }
    __device__ static uint32_t table(const WorldView &wv, bool, const void *&src)
    {
        src = wv.bits_inside;
        return (uint32_t)(bits_words(wv) * 4);
    }
    __device__ void bind(const WorldView &, bool) { inside = reinterpret_cast<const uint32_t *>(g_smem); }
    // rows of p's bins and the set of words to look at
    __device__ __forceinline__ uint32_t rows(const WorldView &wv, const P3 &p, const uint32_t *&tx, const uint32_t *&ty,
                                             const uint32_t *&tz) const
    {
        const int nb = wv.nb, Ws = wv.Ws;
        const int bx = dev_key_bin(p.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb);
        const int by = dev_key_bin(p.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb);
        const int bz = dev_key_bin(p.z, wv.kb.off[2], wv.kb.h0[2], wv.kb.shift[2], nb);
        tx = inside + bx * Ws;
        ty = inside + (nb + by) * Ws;
        tz = inside + (2 * nb + bz) * Ws;
        if (wv.W == 1) return 1u;
        if (wv.cmask_words)
            return inside[wv.cmask_off + ((bz >> wv.cshift[2]) * wv.cn1 + (by >> wv.cshift[1])) * wv.cn0 + (bx >> wv.cshift[0])];
        return tx[wv.W] & ty[wv.W] & tz[wv.W]; // per-axis summaries (word W of a row)
    }
    template <class F>
    __device__ __forceinline__ bool any(const WorldView &wv, const P3 &p, F test) const
    {
        const uint32_t *tx, *ty, *tz;
        uint32_t nz = rows(wv, p, tx, ty, tz);
        while (nz) {
            const int w = __ffs(nz) - 1;
            nz &= nz - 1;
            uint32_t m = tx[w] & ty[w] & tz[w];
            while (m) {
                const int k = (w << 5) + __ffs(m) - 1;
                m &= m - 1;
                if (test(k)) return true;
            }
        }
        return false;
    }
    __device__ __forceinline__ bool has_candidates(const WorldView &wv, const P3 &p) const
    {
        const uint32_t *tx, *ty, *tz;
        uint32_t nz = rows(wv, p, tx, ty, tz);
        while (nz) {
            const int w = __ffs(nz) - 1;
            nz &= nz - 1;
            if (tx[w] & ty[w] & tz[w]) return true;
        }
        return false;
    }
    __device__ __forceinline__ bool invalid(const WorldView &wv, const P3 &p, int mode, bool cpg, double md) const
    {
        return any(wv, p, [&](int k) { return point_candidate_invalid(wv.aabb, wv.obb, k, p, mode, cpg, md); });
    }
};

// EPP_BP_BITS for worlds of up to 32 OBBs (the race track): one word per bin, so the candidate set of a
// point is a single 32-bit mask from three table reads; no loops over words.  Internal variant of the warp form.
constexpr int kBpBits1 = 4;
template <>
struct PointWorld<kBpBits1> {
    const uint32_t *inside;
    __device__ static uint32_t table(const WorldView &wv, bool, const void *&src)
    {
        src = wv.bits_inside;
        return (uint32_t)(3 * wv.nb * 4);
    }
    __device__ void bind(const WorldView &, bool) { inside = reinterpret_cast<const uint32_t *>(g_smem); }
    // Tried and dropped (round 2, profiles/r02_c_ab.log, r02_e): 8 copies of the table in separate bank groups (16 M
    // states: 109 instead of 92 us — the address arithmetic costs more than the ~3.5-deep bank conflicts of 32 random
    // bins) and leaving the nearly full z rows out of phase 1 (no change).
    __device__ __forceinline__ uint32_t candidates(const WorldView &wv, const P3 &p) const
    {
        const int nb = wv.nb;
        const int bx = dev_key_bin(p.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb);
        const int by = dev_key_bin(p.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb);
        const int bz = dev_key_bin(p.z, wv.kb.off[2], wv.kb.h0[2], wv.kb.shift[2], nb);
        return inside[bx] & inside[nb + by] & inside[2 * nb + bz];
    }
    template <class F>
    __device__ __forceinline__ bool any(const WorldView &wv, const P3 &p, F test) const
    {
        uint32_t m = candidates(wv, p);
        while (m) {
            const int k = __ffs(m) - 1;
            m &= m - 1;
            if (test(k)) return true;
        }
        return false;
    }
    __device__ __forceinline__ bool has_candidates(const WorldView &wv, const P3 &p) const
    {
        return candidates(wv, p) != 0u;
    }
};

// ---- records in cell order (GridTables::cell_recs): the whole per-candidate test from one kCellRec-double record.
// Same arithmetic as rec_hit / ray_rec_invalid below (z-rotation form of src/OBB.cpp:63-91, :10-61; inputs finite).

__device__ __forceinline__ bool cellrec_point_invalid(const double *__restrict__ r, const P3 &p, int mode, bool cpg,
                                                      double md)
{
    const bool in = (p.x > __ldg(r + 0)) & (p.x < __ldg(r + 3)) & (p.y > __ldg(r + 1)) & (p.y < __ldg(r + 4)) &
                    (p.z > __ldg(r + 2)) & (p.z < __ldg(r + 5));
    if (!in) return false;
    const uint32_t flags = (uint32_t)__double_as_longlong(__ldg(r + 17)) & 3u;
    const bool filling = (flags & kFlagFilling) != 0;
    if (mode == kModeBool ? (filling && cpg) : filling) return false; // src/World.cpp:92-95, :116-119
    const double infl = filling ? 0.0 : (mode == kModeBool ? __ldg(r + 16) : md);
    const double d0 = dsub(p.x, __ldg(r + 6)), d1 = dsub(p.y, __ldg(r + 7)), d2 = dsub(p.z, __ldg(r + 8));
    const double lx = dadd(dmul(__ldg(r + 9), d0), dmul(__ldg(r + 11), d1));
    const double ly = dadd(dmul(__ldg(r + 10), d0), dmul(__ldg(r + 12), d1));
    const double hx = filling ? __ldg(r + 13) : dadd(__ldg(r + 13), infl);
    const double hy = filling ? __ldg(r + 14) : dadd(__ldg(r + 14), infl);
    const double hz = filling ? __ldg(r + 15) : dadd(__ldg(r + 15), infl);
    return (fabs(lx) <= hx) & (fabs(ly) <= hy) & (fabs(d2) <= hz);
}

__device__ __forceinline__ bool cellrec_ray_invalid(const double *__restrict__ r, const P3 &a, const P3 &b, const P3 &mn,
                                                    const P3 &mx, bool cpg)
{
    if (mx.x < __ldg(r + 0) || mn.x > __ldg(r + 3) || mx.y < __ldg(r + 1) || mn.y > __ldg(r + 4) || mx.z < __ldg(r + 2) ||
        mn.z > __ldg(r + 5))
        return false;
    const uint32_t flags = (uint32_t)__double_as_longlong(__ldg(r + 17)) & 3u;
    const bool filling = (flags & kFlagFilling) != 0;
    if (filling && cpg) return false; // src/World.cpp:150-153
    const double infl = __ldg(r + 16);
    const double h0 = __ldg(r + 13), h1 = __ldg(r + 14), h2 = __ldg(r + 15);
    const double e0 = filling ? h0 : dadd(h0, infl), e1 = filling ? h1 : dadd(h1, infl), e2 = filling ? h2 : dadd(h2, infl);
    const double a0 = dsub(a.x, __ldg(r + 6)), a1 = dsub(a.y, __ldg(r + 7)), a2 = dsub(a.z, __ldg(r + 8));
    const double b0 = dsub(b.x, __ldg(r + 6)), b1 = dsub(b.y, __ldg(r + 7)), b2 = dsub(b.z, __ldg(r + 8));
    const double r00 = __ldg(r + 9), r01 = __ldg(r + 10), r10 = __ldg(r + 11), r11 = __ldg(r + 12);
    const double ax = dadd(dmul(r00, a0), dmul(r10, a1)), ay = dadd(dmul(r01, a0), dmul(r11, a1));
    const double bx = dadd(dmul(r00, b0), dmul(r10, b1)), by = dadd(dmul(r01, b0), dmul(r11, b1));
    if (fabs(ax) <= e0 && fabs(ay) <= e1 && fabs(a2) <= e2) return true; // end points first (src/OBB.cpp:13-18)
    if (fabs(bx) <= e0 && fabs(by) <= e1 && fabs(b2) <= e2) return true;
    double t_min = 0.0, t_max = 1.0;
    if (!slab_axis(dsub(bx, ax), ax, dadd(h0, infl), t_min, t_max)) return false; // every type inflated (:28)
    if (!slab_axis(dsub(by, ay), ay, dadd(h1, infl), t_min, t_max)) return false;
    if (!slab_axis(dsub(b2, a2), a2, dadd(h2, infl), t_min, t_max)) return false;
    return 0.0 <= t_min && t_min <= 1.0 && 0.0 <= t_max && t_max <= 1.0;
}

// Uniform grid: candidates = CSR list of the cell holding p.  The warp-pipelined form first asks an occupancy
// bitmap over (coarsened) cells that lives in shared memory: in a 10 000-OBB world spread over 400 x 400 m a few
// per cent of the states are in an occupied cell, the rest never touch global memory for the broad phase (with
// cell_start read per state the kernel was bound by L2 traffic: 2.3 sectors per 24-byte state).
__device__ __forceinline__ size_t occ_bit(const WorldView &wv, int bx, int by, int bz)
{
    return ((size_t)(bz >> wv.occ_shift[2]) * wv.occ_n1 + (by >> wv.occ_shift[1])) * wv.occ_n0 + (bx >> wv.occ_shift[0]);
}

template <>
struct PointWorld<EPP_BP_GRID> {
    const uint32_t *occ; // warp form: the bitmap's copy in shared memory
    __device__ void init(const WorldView &wv, bool) { occ = wv.occ; }
    __device__ static uint32_t table(const WorldView &wv, bool, const void *&src)
    {
        src = wv.occ;
        return wv.occ_bytes;
    }
    __device__ void bind(const WorldView &, bool) { occ = reinterpret_cast<const uint32_t *>(g_smem); }
    __device__ __forceinline__ static void cell_of(const WorldView &wv, const P3 &p, int &bx, int &by, int &bz)
    {
        bx = dev_bin(p.x, wv.gbins.lo[0], wv.gbins.inv[0], wv.gbins.n[0]);
        by = dev_bin(p.y, wv.gbins.lo[1], wv.gbins.inv[1], wv.gbins.n[1]);
        bz = dev_bin(p.z, wv.gbins.lo[2], wv.gbins.inv[2], wv.gbins.n[2]);
    }
    __device__ __forceinline__ static void cell_range(const WorldView &wv, const P3 &p, uint32_t &s, uint32_t &e)
    {
        int bx, by, bz;
        cell_of(wv, p, bx, by, bz);
        const size_t cell = ((size_t)bz * wv.gbins.n[1] + by) * wv.gbins.n[0] + bx;
        s = __ldg(wv.cell_start + cell);
        e = __ldg(wv.cell_start + cell + 1);
    }
    template <class F>
    __device__ __forceinline__ bool any(const WorldView &wv, const P3 &p, F test) const
    {
        uint32_t s, e;
        cell_range(wv, p, s, e);
        for (uint32_t j = s; j < e; ++j)
            if (test((int)__ldg(wv.cell_items + j))) return true;
        return false;
    }
    // bitmap only (a superset test: the coarse cell may be occupied where the point's own cell is empty)
    __device__ __forceinline__ bool has_candidates(const WorldView &wv, const P3 &p) const
    {
        int bx, by, bz;
        cell_of(wv, p, bx, by, bz);
        const size_t b = occ_bit(wv, bx, by, bz);
        return (occ[b >> 5] >> (b & 31)) & 1u;
    }
    // The cell's candidates from the records stored in cell order: one dependent round trip behind cell_start, two
    // records in flight at a time (item -> AabbRec -> ObbRec was 3 round trips per candidate, one candidate at a time:
    // ncu r02_b showed the warps of the 10 000-OBB run waiting on exactly those loads).
    __device__ __forceinline__ bool invalid(const WorldView &wv, const P3 &p, int mode, bool cpg, double md) const
    {
        if (!wv.cell_recs) return any(wv, p, [&](int k) { return point_candidate_invalid(wv.aabb, wv.obb, k, p, mode, cpg, md); });
        uint32_t s, e;
        cell_range(wv, p, s, e);
        for (uint32_t j = s; j < e; j += 2) {
            const double *r0 = wv.cell_recs + (size_t)j * kCellRec;
            const bool two = j + 1 < e;
            const double *r1 = two ? r0 + kCellRec : r0;
            const bool h0 = cellrec_point_invalid(r0, p, mode, cpg, md);
            const bool h1 = cellrec_point_invalid(r1, p, mode, cpg, md);
            if (h0 | h1) return true;
        }
        return false;
    }
};

// PointWorld<GRID>::invalid as a free function (callable from code that is compiled, but never run, for other BP)
__device__ __forceinline__ bool world_invalid_grid(const WorldView &wv, const P3 &p, int mode, bool cpg, double md)
{
    PointWorld<EPP_BP_GRID> w;
    w.occ = nullptr;
    return w.invalid(wv, p, mode, cpg, md);
}

// ---- segment queries ---------------------------------------------------------------------

template <int BP>
struct RayWorld;

// Every RayWorld offers any(wv, mn, mx, f): calls f(k) for the broad-phase candidates k of the ray box
// [mn, mx] (a superset of the OBBs whose stored AABB the box touches) until one returns true.
template <>
struct RayWorld<EPP_BP_BRUTE> {
    const AabbRec *aabb;
    __device__ void init(const WorldView &wv, bool aabb_in_smem)
    {
        aabb = wv.aabb;
        if (aabb_in_smem) {
            uint32_t *dst = reinterpret_cast<uint32_t *>(g_smem);
            block_copy_words(dst, reinterpret_cast<const uint32_t *>(wv.aabb), wv.M * (int)(sizeof(AabbRec) / 4));
            __syncthreads();
            aabb = reinterpret_cast<const AabbRec *>(g_smem);
        }
    }
    __device__ void bind(const WorldView &) {}
    template <typename F>
    __device__ __forceinline__ bool any(const WorldView &wv, const P3 &, const P3 &, F f) const
    {
        for (int k = 0; k < wv.M; ++k)
            if (f(k)) return true;
        return false;
    }
    __device__ __forceinline__ bool invalid(const WorldView &wv, const P3 &a, const P3 &b, bool cpg) const
    {
        P3 mn, mx;
        ray_box(a, b, mn, mx);
        for (int k = 0; k < wv.M; ++k) {
            if (!boxes_touch(aabb + k, mn, mx)) continue;
            if (ray_candidate_invalid(wv.aabb, wv.obb, k, a, b, mn, mx, cpg)) return true;
        }
        return false;
    }
};

// candidates = AND over axes of max_ge[bin(rayMin)] & min_le[bin(rayMax)].  Rows have an ODD stride of
// Ws words (W + a summary word, padded to odd; built that way on the host): lanes look up unrelated bins, and with the natural stride of W = 8 words every row
// starts in one of four banks (ncu: 10.5 M bank-conflict cycles per 1 M segments, 40 % of the kernel's cycles).
template <>
struct RayWorld<EPP_BP_BITS> {
    const uint32_t *min_le; // cumulative rows: shared memory when the world has no range rows, else global (rare path)
    const uint32_t *max_ge;
    const uint32_t *range;  // (first bin, span) rows of x and y in shared memory
    __device__ void init(const WorldView &wv, bool)
    {
        uint32_t *dst = reinterpret_cast<uint32_t *>(g_smem);
        if (wv.range_d > 0) {
            block_copy_words(dst, wv.bits_range, 2 * wv.nb * wv.range_d * wv.Ws);
            range = dst;
            min_le = wv.bits_min_le;
            max_ge = wv.bits_max_ge;
        } else {
            const int words = 3 * wv.nb * wv.Ws;
            block_copy_words(dst, wv.bits_min_le, words);
            block_copy_words(dst + words, wv.bits_max_ge, words);
            range = nullptr;
            min_le = dst;
            max_ge = dst + words;
        }
        __syncthreads();
    }
    // The two range rows of a ray box; false: the box spans too many bins (cumulative rows, any()).
    __device__ __forceinline__ bool rows(const WorldView &wv, const P3 &mn, const P3 &mx, const uint32_t *&rx,
                                         const uint32_t *&ry) const
    {
        const int nb = wv.nb, Ws = wv.Ws, D = wv.range_d;
        const int gx_ = dev_key_bin(mn.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb), lx_ = dev_key_bin(mx.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb);
        const int gy_ = dev_key_bin(mn.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb), ly_ = dev_key_bin(mx.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb);
        if (!(lx_ - gx_ < D && ly_ - gy_ < D)) return false;
        rx = range + ((gx_ * D) + (lx_ - gx_)) * Ws;
        ry = range + (((nb + gy_) * D) + (ly_ - gy_)) * Ws;
        return true;
    }
    // the range rows arrived in shared memory by bulk copy (k_segments_warp)
    __device__ void bind(const WorldView &wv)
    {
        range = reinterpret_cast<const uint32_t *>(g_smem);
        min_le = wv.bits_min_le;
        max_ge = wv.bits_max_ge;
    }
    template <typename F>
    __device__ __forceinline__ bool any(const WorldView &wv, const P3 &mn, const P3 &mx, F f) const
    {
        const int nb = wv.nb, W = wv.W, Ws = wv.Ws;
        const int gx_ = dev_key_bin(mn.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb), lx_ = dev_key_bin(mx.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb);
        const int gy_ = dev_key_bin(mn.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb), ly_ = dev_key_bin(mx.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb);
        const int D = wv.range_d;
        if (lx_ - gx_ < D && ly_ - gy_ < D) {
            // One row per axis: the OBBs whose bin range meets [bin(rayMin), bin(rayMax)] on x and on y (z is left to
            // the exact stored-AABB test every candidate takes anyway).  Two reads name the words, two more per word.
            const uint32_t *rx = range + ((gx_ * D) + (lx_ - gx_)) * Ws, *ry = range + (((nb + gy_) * D) + (ly_ - gy_)) * Ws;
            uint32_t nz = W == 1 ? 1u : (rx[W] & ry[W]);
            while (nz) {
                const int w = __ffs(nz) - 1;
                nz &= nz - 1;
                uint32_t m = rx[w] & ry[w];
                while (m) {
                    const int k = (w << 5) + __ffs(m) - 1;
                    m &= m - 1;
                    if (f(k)) return true;
                }
            }
            return false;
        }
        // Boxes spanning many bins (or worlds without range rows): cumulative rows, all three axes.
        const int gz_ = dev_key_bin(mn.z, wv.kb.off[2], wv.kb.h0[2], wv.kb.shift[2], nb), lz_ = dev_key_bin(mx.z, wv.kb.off[2], wv.kb.h0[2], wv.kb.shift[2], nb);
        const uint32_t *gx = max_ge + gx_ * Ws, *gy = max_ge + (nb + gy_) * Ws, *gz = max_ge + (2 * nb + gz_) * Ws;
        const uint32_t *lx = min_le + lx_ * Ws, *ly = min_le + (nb + ly_) * Ws, *lz = min_le + (2 * nb + lz_) * Ws;
        // Pass 1, converged: which words of the candidate set are non-zero at all.  Walking the candidates inside this
        // loop made the lanes of a warp drift through it separately (ncu: the row AND ran 2.6 x per warp).
        uint32_t nz = 0u;
        for (int w = 0; w < W; ++w)
            if (gx[w] & gy[w] & gz[w] & lx[w] & ly[w] & lz[w]) nz |= 1u << w;
        // Pass 2: the few non-zero words again, in order, with the exact tests
        while (nz) {
            const int w = __ffs(nz) - 1;
            nz &= nz - 1;
            uint32_t m = gx[w] & gy[w] & gz[w] & lx[w] & ly[w] & lz[w];
            while (m) {
                const int k = (w << 5) + __ffs(m) - 1;
                m &= m - 1;
                if (f(k)) return true;
            }
        }
        return false;
    }
    __device__ __forceinline__ bool invalid(const WorldView &wv, const P3 &a, const P3 &b, bool cpg) const
    {
        P3 mn, mx;
        ray_box(a, b, mn, mx);
        return any(wv, mn, mx, [&](int k) -> bool { return ray_candidate_invalid(wv.aabb, wv.obb, k, a, b, mn, mx, cpg); });
    }
};

// Cells overlapped by the ray box; a candidate listed in several of them is evaluated only
// in the first cell of (ray cells) x (candidate cells).  has_candidates() asks the occupancy bitmap (PointWorld<GRID>
// above) about the coarse cells under the ray box first: rows of x-bits, at most two words per (y, z) row.
template <>
struct RayWorld<EPP_BP_GRID> {
    const uint32_t *occ;
    __device__ void init(const WorldView &wv, bool) { occ = wv.occ; }
    __device__ void bind(const WorldView &) { occ = reinterpret_cast<const uint32_t *>(g_smem); }
    __device__ __forceinline__ static void cells_of(const WorldView &wv, const P3 &mn, const P3 &mx, int &x0, int &x1,
                                                    int &y0, int &y1, int &z0, int &z1)
    {
        const AxisBins &g = wv.gbins;
        x0 = dev_bin(mn.x, g.lo[0], g.inv[0], g.n[0]), x1 = dev_bin(mx.x, g.lo[0], g.inv[0], g.n[0]);
        y0 = dev_bin(mn.y, g.lo[1], g.inv[1], g.n[1]), y1 = dev_bin(mx.y, g.lo[1], g.inv[1], g.n[1]);
        z0 = dev_bin(mn.z, g.lo[2], g.inv[2], g.n[2]), z1 = dev_bin(mx.z, g.lo[2], g.inv[2], g.n[2]);
    }
    template <bool GLOBAL>
    __device__ __forceinline__ bool has_candidates_t(const WorldView &wv, const P3 &mn, const P3 &mx) const
    {
        int x0, x1, y0, y1, z0, z1;
        cells_of(wv, mn, mx, x0, x1, y0, y1, z0, z1);
        x0 >>= wv.occ_shift[0], x1 >>= wv.occ_shift[0];
        y0 >>= wv.occ_shift[1], y1 >>= wv.occ_shift[1];
        z0 >>= wv.occ_shift[2], z1 >>= wv.occ_shift[2];
        const int len = x1 - x0 + 1;
        for (int z = z0; z <= z1; ++z)
            for (int y = y0; y <= y1; ++y) {
                size_t s = ((size_t)z * wv.occ_n1 + y) * wv.occ_n0 + x0;
                if (len <= 32) { // the usual case: the row's bits lie in two consecutive words (the map has slack words)
                    const uint32_t lo = GLOBAL ? __ldg(occ + (s >> 5)) : occ[s >> 5];
                    const uint32_t hi = GLOBAL ? __ldg(occ + (s >> 5) + 1) : occ[(s >> 5) + 1];
                    const uint32_t bits = __funnelshift_r(lo, hi, (uint32_t)(s & 31));
                    if (bits & (len == 32 ? 0xffffffffu : ((1u << len) - 1u))) return true;
                    continue;
                }
                int left = len;
                while (left > 0) {
                    const int off = (int)(s & 31), take = (32 - off) < left ? (32 - off) : left;
                    const uint32_t m = (take == 32 ? 0xffffffffu : ((1u << take) - 1u)) << off;
                    const uint32_t w = GLOBAL ? __ldg(occ + (s >> 5)) : occ[s >> 5];
                    if (w & m) return true;
                    s += take;
                    left -= take;
                }
            }
        return false;
    }
    __device__ __forceinline__ bool has_candidates(const WorldView &wv, const P3 &mn, const P3 &mx) const
    {
        return has_candidates_t<false>(wv, mn, mx);
    }
    template <typename F>
    __device__ __forceinline__ bool any(const WorldView &wv, const P3 &mn, const P3 &mx, F f) const
    {
        const AxisBins &g = wv.gbins;
        int x0, x1, y0, y1, z0, z1;
        cells_of(wv, mn, mx, x0, x1, y0, y1, z0, z1);
        for (int z = z0; z <= z1; ++z)
            for (int y = y0; y <= y1; ++y) {
                const size_t row = ((size_t)z * g.n[1] + y) * g.n[0];
                for (int x = x0; x <= x1; ++x) {
                    const uint32_t s = __ldg(wv.cell_start + row + x), e = __ldg(wv.cell_start + row + x + 1);
                    for (uint32_t j = s; j < e; ++j) {
                        const int k = (int)__ldg(wv.cell_items + j);
                        const uint64_t cm = wv.obb[k].cell_min;
                        const int kx = (int)(cm & 0x1FFFFF), ky = (int)((cm >> 21) & 0x1FFFFF), kz = (int)(cm >> 42);
                        if (x != (kx > x0 ? kx : x0) || y != (ky > y0 ? ky : y0) || z != (kz > z0 ? kz : z0)) continue;
                        if (f(k)) return true;
                    }
                }
            }
        return false;
    }
    // Candidates from the records stored in cell order.  The cells of one (y, z) row under the ray box are consecutive,
    // so their CSR ranges are too: one pair of cell_start reads per row, then the row's records back to back.  A copy
    // is evaluated only in the first cell of (ray cells) x (the OBB's cells) — both are in the record.
    __device__ __forceinline__ bool walk_recs(const WorldView &wv, const P3 &a, const P3 &b, const P3 &mn, const P3 &mx, bool cpg) const
    {
        const AxisBins &g = wv.gbins;
        int x0, x1, y0, y1, z0, z1;
        cells_of(wv, mn, mx, x0, x1, y0, y1, z0, z1);
        for (int z = z0; z <= z1; ++z)
            for (int y = y0; y <= y1; ++y) {
                const size_t row = ((size_t)z * g.n[1] + y) * g.n[0];
                const uint32_t s = __ldg(wv.cell_start + row + x0), e = __ldg(wv.cell_start + row + x1 + 1);
                for (uint32_t j = s; j < e; ++j) {
                    const double *r = wv.cell_recs + (size_t)j * kCellRec;
                    const uint64_t own = (uint64_t)__double_as_longlong(__ldg(r + 17)), here = (uint64_t)__double_as_longlong(__ldg(r + 18));
                    const int kx = (int)((own >> 2) & 0xFFFFF), ky = (int)((own >> 22) & 0xFFFFF), kz = (int)((own >> 42) & 0xFFFFF);
                    const int cx = (int)(here & 0xFFFFF), cy = (int)((here >> 20) & 0xFFFFF), cz = (int)((here >> 40) & 0xFFFFF);
                    if (cx != (kx > x0 ? kx : x0) || cy != (ky > y0 ? ky : y0) || cz != (kz > z0 ? kz : z0)) continue;
                    if (cellrec_ray_invalid(r, a, b, mn, mx, cpg)) return true;
                }
            }
        return false;
    }
    __device__ __forceinline__ bool invalid(const WorldView &wv, const P3 &a, const P3 &b, bool cpg) const
    {
        P3 mn, mx;
        ray_box(a, b, mn, mx);
        if (!has_candidates_t<true>(wv, mn, mx)) return false; // direct form: the bitmap is read from global memory
        if (wv.cell_recs) return walk_recs(wv, a, b, mn, mx, cpg);
        return any(wv, mn, mx, [&](int k) -> bool { return ray_candidate_invalid(wv.aabb, wv.obb, k, a, b, mn, mx, cpg); });
    }
};

// ---- K1: states ----------------------------------------------------------------------------

// Tile = kBlock consecutive states; block t, t + gridDim, ... ; the next tile's coordinates are
// fetched before the current tile is evaluated so two tiles per thread are in flight.
template <int BP>
__global__ void __launch_bounds__(kBlock) k_states(WorldView wv, const double *__restrict__ xyz, size_t n, int mode,
                                                   int can_pass_gate, double min_distance, int aabb_in_smem,
                                                   uint32_t *__restrict__ mask)
{
    PointWorld<BP> world;
    world.init(wv, aabb_in_smem != 0);
    const size_t tiles = (n + kBlock - 1) / kBlock;
    size_t tile = blockIdx.x;
    if (tile >= tiles) return;

    size_t i = tile * kBlock + threadIdx.x;
    P3 cur = {0.0, 0.0, 0.0};
    if (i < n) {
        cur.x = __ldg(xyz + 3 * i);
        cur.y = __ldg(xyz + 3 * i + 1);
        cur.z = __ldg(xyz + 3 * i + 2);
    }
    for (;;) {
        const size_t next_tile = tile + gridDim.x;
        const size_t ni = next_tile * kBlock + threadIdx.x;
        P3 nxt = {0.0, 0.0, 0.0};
        if (next_tile < tiles && ni < n) {
            nxt.x = __ldg(xyz + 3 * ni);
            nxt.y = __ldg(xyz + 3 * ni + 1);
            nxt.z = __ldg(xyz + 3 * ni + 2);
        }
        bool valid = false;
        if (i < n) valid = !world.invalid(wv, cur, mode, can_pass_gate != 0, min_distance);
        const uint32_t word = __ballot_sync(0xffffffffu, valid);
        if ((threadIdx.x & 31) == 0 && i < n) mask[i >> 5] = word;
        if (next_tile >= tiles) break;
        tile = next_tile;
        i = ni;
        cur = nxt;
    }
}

// ---- TMA bulk-copy plumbing (cp.async.bulk -> SASS UBLKCP, completion on an mbarrier) ------------

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}

// Make freshly initialised barriers visible to the async (TMA) proxy.
__device__ __forceinline__ void mbar_init_fence()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_addr(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) {
    }
}

// global -> shared bulk copy; dst, src and bytes must be multiples of 16.
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}

// ---- K1 (warp-pipelined form): every warp streams its own contiguous range of states -----------------
//
// No block-wide barrier after the prologue.  Warp j of the grid owns a contiguous run of mask words and walks
// it in steps of S words (S*32 states, lane l takes state l of each word).  Each warp runs its own
// cp.async.bulk ring (`stages` slots of S*768 B, one mbarrier per slot, lane 0 issues), so a warp never waits
// on another warp.  A step has two dense phases:
//   phase 1  every lane bins its S states and asks the broad phase only WHETHER there is a candidate
//            (~10 % of the states of the race track); states with candidates are compacted into a per-warp
//            byte queue with one ballot per word, no atomics, no loops;
//   phase 2  lanes take queued states one each and walk their candidates through the reference's exact
//            test with early exit; hits are OR-ed into the warp's S-word invalid bitmap.
// For worlds of up to kRecSmemMax OBBs the prologue folds the per-call rules of World::checkPointValidity
// (src/World.cpp:92-97, :116-121) into one 22-double record per OBB in shared memory: stored AABB, centre,
// rotation and the EFFECTIVE half sizes (half + inflate for collision boxes, half for filling boxes, -1 for
// boxes the call skips, which no |local| can be <=).
// Record = 17 doubles (odd stride: lanes reading the same field of different records hit different banks):
// [0..2] aabb min, [3..5] aabb max, [6..8] centre, [9..12] R(0,0) R(0,1) R(1,0) R(1,1), [13..15] eff. half.
// Only worlds whose rotations are all about z get records (WorldView::zrot; the only kind World::addGate /
// addObstacle can build, src/Object.cpp:38-47).  There R(2,0) = R(2,1) = R(0,2) = R(1,2) = 0 and R(2,2) = 1, so the
// reference's 3-term sums are (a + b) + (+-0) and (+-0 + +-0) + d2: dropping the zero terms changes at most
// the sign of a zero, which fabs() and the comparisons cannot see.  The point is finite whenever the result
// matters because it passed the strict stored-AABB test.
#ifdef EPP_TRACE
// Profiling build only (make trace): per-warp timeline of the warp form, 16 globaltimer stamps per warp.
__device__ unsigned long long g_trace[1 << 20];
__device__ uint32_t g_trace_base; // 0 or 1 << 19: which half of g_trace the running launch writes
__device__ __forceinline__ void trace(uint32_t g, int ev)
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if ((threadIdx.x & 31) == 0 && ev < 16 && g * 16 + ev < (1u << 19)) g_trace[g_trace_base + g * 16 + ev] = t;
}
__global__ void k_trace_base(uint32_t base) { g_trace_base = base; }
#define EPP_TRACE_EV(g, ev) trace(g, ev)
#else
#define EPP_TRACE_EV(g, ev)
#endif


// ---- mask exchange over peer memory (GatherView, kernels.h) --------------------------------------------

// This call's number = calls completed + 1, read by ONE thread per block at its start (after the previous
// launch on the gather has completed).  The block that reads last closes the count for the next launch, so
// nothing is left to do at the tail of the kernel.  Requires what the launchers guarantee: the next launch on
// this gather does not read before this grid has completed (stream order / griddepcontrol.wait).
__device__ __forceinline__ uint32_t gather_seq(const GatherView &gv)
{
    uint32_t v;
    // an ACQUIRE load: it is performed before the atomic below counts this block (the block that is counted last
    // bumps ctl[1]; a full fence here cost 3 us per launch)
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(gv.ctl + 1) : "memory");
    if (atomicAdd(&gv.ctl[0], 1u) == gridDim.x - 1u) {
        gv.ctl[0] = 0u;
        asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(gv.ctl + 1), "r"(v + 1u) : "memory");
    }
    return v + 1u;
}

// Exchange units: a QUAD = three mask words + flags in one 16-byte store (layout at ll_store); unit u of a rank's
// slot holds words 3u .. 3u + 2.  Round 1 sent (word, number) PAIRS: 8 bytes per word over NVLink, 5.3 now.
// Staging row of (parity, source) = header quad + acknowledgement quad (deferred form) + ceil(slot_words / 3) data quads.

// quad index of unit 0 of (parity of seq, source rank) in a staging area
__device__ __forceinline__ uint64_t gather_row(const GatherView &gv, uint32_t seq, int src)
{
    return ((uint64_t)(seq & 1u) * (uint64_t)gv.world + (uint64_t)src) * gv.row_quads + 2u;
}

// A 16-byte vector access is NOT single-copy atomic (PTX treats it as separate element accesses, and a reader polling
// while the writer's store is in flight did see the flag of a quad whose other half was still stale: a header read
// that way zeroed one block's share of a peer's slot about once in a hundred first calls).  So each 8-byte HALF of a
// quad carries the low 16 bits of the sequence number next to 48 data bits, and a quad counts only when both agree:
//   half 0 = w0 | w1[15:0] << 32 | flag << 48        half 1 = w1[31:16] | w2 << 16 | flag << 48
// 16 bits suffice: a row is reused every second call, so a stale flag differs by 2 (and a fresh row holds 0, which
// no call's flag equals before 65536 calls have made every row non-zero).
__device__ __forceinline__ void ll_store(uint4 *p, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t seq)
{
    const unsigned long long f = (unsigned long long)(seq & 0xffffu) << 48;
    const unsigned long long h0 = (unsigned long long)w0 | ((unsigned long long)(w1 & 0xffffu) << 32) | f;
    const unsigned long long h1 = (unsigned long long)(w1 >> 16) | ((unsigned long long)w2 << 16) | f;
    asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(h0), "l"(h1) : "memory");
}

// Returns the quad as (w0, w1, w2, ok): ok = 1 iff both halves carry this call's flag.
__device__ __forceinline__ uint4 ll_load(const uint4 *p, uint32_t seq)
{
    unsigned long long h0, h1;
    asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(h0), "=l"(h1) : "l"(p) : "memory");
    const uint32_t f = seq & 0xffffu;
    uint4 q;
    q.x = (uint32_t)h0;
    q.y = (uint32_t)((h0 >> 32) & 0xffffu) | ((uint32_t)(h1 & 0xffffu) << 16);
    q.z = (uint32_t)(h1 >> 16);
    q.w = ((uint32_t)(h0 >> 48) == f && (uint32_t)(h1 >> 48) == f) ? 1u : 0u;
    return q;
}

// Wait until both halves of the quad carry this call's flag.  Bounded: a time-out is counted and reads as 0.
__device__ __forceinline__ uint4 ll_wait(const uint4 *p, uint32_t seq, const GatherView &gv, bool &dead)
{
    uint32_t spins = 0;
    unsigned long long t0 = 0, t;
    for (;;) {
        const uint4 q = ll_load(p, seq);
        if (q.w) return q;
        if (dead) return make_uint4(0u, 0u, 0u, 1u); // this thread has already waited one full time-out in this call
        if ((++spins & 63u) == 0u) {
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t0 == 0) t0 = t;
            else if (t - t0 > gv.timeout_ns) {
                atomicAdd(&gv.ctl[2], 1u);
                dead = true;
                return make_uint4(0u, 0u, 0u, 1u);
            }
        }
    }
}

// One thread per peer announces how many words this rank sends in this call.
__device__ __forceinline__ void gather_send_header(const GatherView &gv, uint32_t seq, int lane)
{
    if (lane < gv.world && lane != gv.rank && !(gv.debug & 2))
        ll_store(gv.peer_ll[lane] + gather_row(gv, seq, gv.rank) - 2u, (uint32_t)gv.my_words, 0u, 0u, seq);
}

// How many words each peer sends in this call (its header quad), one lane per peer, into cnt[] (shared memory).
// One warp per BLOCK does this: thousands of warps polling one address would serialise on its L2 slice.
__device__ __forceinline__ void gather_read_headers(const GatherView &gv, uint32_t seq, uint32_t *cnt, int lane)
{
    bool dead = false;
    if (lane < gv.world && lane != gv.rank) cnt[lane] = ll_wait(gv.my_ll + gather_row(gv, seq, lane) - 2u, seq, gv, dead).x;
}

// A warp unpacks units [u0, u1) of every peer's slot into the local gathered buffer (0 beyond what the peer sends).
// The quads of ALL peers for a lane's unit are requested before any is examined: one L2 round trip per unit
// instead of one per peer (7 dependent round trips at 8 ranks were ~2.5 us at the tail of every launch).
__device__ __forceinline__ void gather_unpack(const GatherView &gv, uint32_t seq, const uint32_t *cnt, uint64_t u0,
                                              uint64_t u1, int lane, uint32_t *out = nullptr, bool zero_own_tail = true)
{
    bool dead = false;
    if (!out) out = gv.out;
    const uint4 *base = gv.my_ll + (uint64_t)(seq & 1u) * (uint64_t)gv.world * gv.row_quads + 2u;
    uint32_t *own = out + (uint64_t)gv.rank * gv.slot_words; // own slot: the check wrote [0, my_words); the rest reads 0
    if (zero_own_tail) {
        const uint64_t w1 = 3u * u1 < gv.slot_words ? 3u * u1 : gv.slot_words;
        for (uint64_t w = (3u * u0 > gv.my_words ? 3u * u0 : gv.my_words) + (uint64_t)lane; w < w1; w += 32u) own[w] = 0u;
    }
    for (uint64_t u = u0 + (uint64_t)lane; u < u1; u += 32u) {
        uint4 q[kMaxRanks];
#pragma unroll
        for (int p = 0; p < kMaxRanks; ++p) {
            q[p] = make_uint4(0u, 0u, 0u, 1u);
            if (p < gv.world && p != gv.rank && 3u * u < cnt[p]) q[p] = ll_load(base + (uint64_t)p * gv.row_quads + u, seq);
        }
#pragma unroll
        for (int p = 0; p < kMaxRanks; ++p) {
            if (p >= gv.world || p == gv.rank) continue;
            if (!q[p].w) q[p] = ll_wait(base + (uint64_t)p * gv.row_quads + u, seq, gv, dead);
            uint32_t *dst = out + (uint64_t)p * gv.slot_words + 3u * u;
            const uint64_t w = 3u * u, c = cnt[p];
            if (w < gv.slot_words) dst[0] = w < c ? q[p].x : 0u;
            if (w + 1u < gv.slot_words) dst[1] = w + 1u < c ? q[p].y : 0u;
            if (w + 2u < gv.slot_words) dst[2] = w + 2u < c ? q[p].z : 0u;
        }
    }
}

// Generic form: any check kernel wrote this rank's words into its own slot of the local gathered buffer;
// send them to every peer as quads, then unpack the peers' slots.
__global__ void __launch_bounds__(256) k_gather_push(const __grid_constant__ GatherView gv)
{
    __shared__ uint32_t s_seq, s_cnt[kMaxRanks];
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0) s_seq = gather_seq(gv);
    __syncthreads();
    const uint32_t seq = s_seq;
    const int lane = threadIdx.x & 31;
    if (blockIdx.x == 0 && threadIdx.x < 32) gather_send_header(gv, seq, lane);
    const uint32_t *src = gv.out + (uint64_t)gv.rank * gv.slot_words;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x, t0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t my_units = (gv.my_words + 2u) / 3u, slot_units = (gv.slot_words + 2u) / 3u;
    if (!(gv.debug & 2))
        for (int p = 0; p < gv.world; ++p) {
            if (p == gv.rank) continue;
            uint4 *dst = gv.peer_ll[p] + gather_row(gv, seq, gv.rank);
            for (uint64_t i = t0; i < my_units; i += stride) {
                const uint64_t w = 3u * i;
                ll_store(dst + i, src[w], w + 1u < gv.my_words ? src[w + 1u] : 0u, w + 2u < gv.my_words ? src[w + 2u] : 0u, seq);
            }
        }
    // this block's share of the peers' slots, a warp at a time
    const uint64_t warps = (uint64_t)gridDim.x * (blockDim.x >> 5), wid = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const uint64_t per = (slot_units + warps - 1) / warps;
    const uint64_t u0 = wid * per < slot_units ? wid * per : slot_units;
    const uint64_t u1 = u0 + per < slot_units ? u0 + per : slot_units;
    if (!(gv.debug & 1)) {
        if (threadIdx.x < 32) gather_read_headers(gv, seq, s_cnt, lane);
        __syncthreads();
        gather_unpack(gv, seq, s_cnt, u0, u1, lane);
    }
}

// Deferred form of the exchange for SMALL steps (1 M states per rank: a 9 us check).  In the forms above a call ends
// only when the peers' words of THIS call have arrived: every step is a barrier across the ranks plus one NVLink
// latency, with the GPU idle meanwhile (8 ranks: 15.7 us per step around an 8.5 us check).  Here call s sends its
// words and unpacks the peers' words of call s - 1, which arrived a step ago: nothing waits.  Buffer s - 1 is complete
// when call s has ended; epp_gather_flush completes the last one.  A rank may run ahead of a peer by a step, so before
// it overwrites the rows of call s - 2 it needs the peers' ACKNOWLEDGEMENT that they have unpacked them (one quad per
// peer, stored by the block that finishes the unpack last).  flush = 1: no sends, unpack the last call.
__global__ void __launch_bounds__(256) k_gather_deferred(const __grid_constant__ GatherView gv, int flush)
{
    __shared__ uint32_t s_seq, s_last, s_cnt[kMaxRanks];
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (threadIdx.x == 0) {
        if (flush) asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(s_seq) : "l"(gv.ctl + 1) : "memory");
        else s_seq = gather_seq(gv);
    }
    __syncthreads();
    const uint32_t seq = s_seq; // flush: the number of the last call
    const int lane = threadIdx.x & 31;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x, t0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t my_units = (gv.my_words + 2u) / 3u, slot_units = (gv.slot_words + 2u) / 3u;
    if (!flush) {
        uint32_t *own = gv.out + (uint64_t)gv.rank * gv.slot_words;
        for (uint64_t w = gv.my_words + t0; w < gv.slot_words; w += stride) own[w] = 0u; // beyond this rank's units
        // the rows of this parity still hold call seq - 2 until every peer has said it is done with them
        if (seq > 2u && threadIdx.x < 32 && lane < gv.world && lane != gv.rank) {
            bool dead = false;
            ll_wait(gv.my_ll + gather_row(gv, seq, lane) - 1u, seq - 2u, gv, dead);
        }
        __syncthreads();
        if (blockIdx.x == 0 && threadIdx.x < 32) gather_send_header(gv, seq, lane);
        const uint32_t *src = gv.out + (uint64_t)gv.rank * gv.slot_words;
        for (int p = 0; p < gv.world; ++p) {
            if (p == gv.rank) continue;
            uint4 *dst = gv.peer_ll[p] + gather_row(gv, seq, gv.rank);
            for (uint64_t i = t0; i < my_units; i += stride) {
                const uint64_t w = 3u * i;
                ll_store(dst + i, src[w], w + 1u < gv.my_words ? src[w + 1u] : 0u, w + 2u < gv.my_words ? src[w + 2u] : 0u, seq);
            }
        }
    }
    // unpack: call seq - 1 into the previous call's buffer (flush: call seq into its own)
    const uint32_t useq = flush ? seq : seq - 1u;
    uint32_t *uout = flush ? gv.out : gv.prev_out;
    if (useq == 0u || !uout) return;
    const uint64_t warps = (uint64_t)gridDim.x * (blockDim.x >> 5), wid = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const uint64_t per = (slot_units + warps - 1) / warps;
    const uint64_t u0 = wid * per < slot_units ? wid * per : slot_units;
    const uint64_t u1 = u0 + per < slot_units ? u0 + per : slot_units;
    if (threadIdx.x < 32) gather_read_headers(gv, useq, s_cnt, lane);
    __syncthreads();
    gather_unpack(gv, useq, s_cnt, u0, u1, lane, uout, false);
    // the block that finishes last tells every peer that its rows of call useq may be overwritten
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&gv.ctl[3], 1u) == gridDim.x - 1u ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    if (threadIdx.x == 0) gv.ctl[3] = 0u;
    if (threadIdx.x < 32 && lane < gv.world && lane != gv.rank)
        ll_store(gv.peer_ll[lane] + gather_row(gv, useq, gv.rank) - 1u, useq, 0u, 0u, useq);
}

// Bulk form of the exchange for LARGE slots (config 5: 3.9 M mask words per rank at 8 ranks).  Flags in the data make
// a 9 us step cheap but cost a third more NVLink bytes, a staging area and an unpack pass; for megabytes the plain
// way wins: every block copies its share of the rank's slot straight into slot `rank` of every peer's gathered buffer
// (16-byte stores, 4 bytes per word on the wire, nothing to unpack), makes its stores visible system-wide and is
// counted; the block that is counted last tells the peers (their header quad) and waits for theirs.  When the kernel
// has ended, every peer's slot is in place.  Peers write into the gathered buffer itself, so consecutive calls must
// alternate buffers (checked on the host).  8 ranks, 1 B states: 0.2 ms instead of 0.7 ms behind the check kernel.
__global__ void __launch_bounds__(256) k_gather_bulk(const __grid_constant__ GatherView gv)
{
    __shared__ uint32_t s_seq, s_last, s_cnt[kMaxRanks];
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0) s_seq = gather_seq(gv);
    __syncthreads();
    const uint32_t seq = s_seq;
    uint32_t *own = gv.out + (uint64_t)gv.rank * gv.slot_words;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x, t0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t n4 = gv.slot_words / 4; // slot_words is a multiple of 4 (epp_gather_create)
    for (uint64_t i = t0; i < n4; i += stride) {
        const uint64_t w = 4u * i;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (w + 3u < gv.my_words) {
            v = *reinterpret_cast<const uint4 *>(own + w);
        } else {
            if (w < gv.my_words) v.x = own[w];
            if (w + 1u < gv.my_words) v.y = own[w + 1u];
            if (w + 2u < gv.my_words) v.z = own[w + 2u];
            *reinterpret_cast<uint4 *>(own + w) = v; // words beyond this rank's units read 0 everywhere
        }
        if (!(gv.debug & 2))
            for (int p = 0; p < gv.world; ++p)
                if (p != gv.rank)
                    asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(gv.peer_out[p] + (uint64_t)gv.rank * gv.slot_words + w),
                                 "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&gv.ctl[3], 1u) == gridDim.x - 1u ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    if (threadIdx.x == 0) gv.ctl[3] = 0u;
    __threadfence_system();
    const int lane = threadIdx.x & 31;
    if (threadIdx.x < 32) {
        gather_send_header(gv, seq, lane);
        if (!(gv.debug & 1)) gather_read_headers(gv, seq, s_cnt, lane);
        __threadfence_system();
    }
}

constexpr int kRecDoubles = 17;
constexpr int kRecSmemMax = 256; // default of the states_rec_smem_max knob; one prologue thread per record (<= 992)
constexpr int kWordBytes = 32 * 3 * 8; // one mask word of states = 768 B of coordinates

__device__ __forceinline__ bool rec_hit(const double *__restrict__ r, const P3 &p)
{
    const bool in = (p.x > r[0]) & (p.x < r[3]) & (p.y > r[1]) & (p.y < r[4]) & (p.z > r[2]) & (p.z < r[5]);
    const double d0 = dsub(p.x, r[6]), d1 = dsub(p.y, r[7]), d2 = dsub(p.z, r[8]);
    const double lx = dadd(dmul(r[9], d0), dmul(r[11], d1));  // R(0,0) d0 + R(1,0) d1   (src/OBB.cpp:66)
    const double ly = dadd(dmul(r[10], d0), dmul(r[12], d1)); // R(0,1) d0 + R(1,1) d1
    return in & (fabs(lx) <= r[13]) & (fabs(ly) <= r[14]) & (fabs(d2) <= r[15]);
}

// SOA: the slot holds three runs of S*32 doubles (x, y, z columns of a trajectory matrix) instead of xyz triples.
// Returns, in lanes < S, the INVALID bits of the lane's word (0 elsewhere); mask_out may be null.
template <int BP, int S, bool RS, bool GA, bool SOA = false>
__device__ __forceinline__ uint32_t warp_states_step(const PointWorld<BP> &world, const WorldView &wv,
                                                 const double *__restrict__ rec, const double *__restrict__ pts,
                                                 int units, int mode, bool cpg, double md, unsigned char *queue,
                                                 uint32_t *inv, uint32_t *__restrict__ mask_out, int lane,
                                                 const GatherView &gv, uint32_t seq, uint64_t ll_word0,
                                                 uint32_t *stage_words)
{
    // ---- phase 1: which of my S states have a broad-phase candidate at all
    const uint32_t lt = (1u << lane) - 1u;
    int queued = 0;
#ifdef EPP_TRACE
    const int dbg = (mode >> 8) & 0xff; // profiling build only (epp_set_tuning states_debug): 1 skips phase 2, 2 also phase 1
#else
    constexpr int dbg = 0;
#endif
    mode &= 0xff;
#pragma unroll
    for (int u = 0; u < S; ++u) {
        const int unit = u * 32 + lane;
        bool has = false;
        if (unit < units) {
            const P3 p = SOA ? P3{pts[unit], pts[S * 32 + unit], pts[2 * S * 32 + unit]}
                             : P3{pts[3 * unit], pts[3 * unit + 1], pts[3 * unit + 2]};
            has = dbg == 2 ? (p.x + p.y + p.z == 12345.0) : world.has_candidates(wv, p);
        }
        const uint32_t b = __ballot_sync(0xffffffffu, has);
        if (has) queue[queued + __popc(b & lt)] = (unsigned char)unit;
        queued += __popc(b);
    }
    if (queued != 0 && dbg == 0) {
        __syncwarp();
        // ---- phase 2: one queued state per lane, exact tests with early exit
        for (int q = lane; q < queued; q += 32) {
            const int unit = queue[q];
            const P3 p = SOA ? P3{pts[unit], pts[S * 32 + unit], pts[2 * S * 32 + unit]}
                             : P3{pts[3 * unit], pts[3 * unit + 1], pts[3 * unit + 2]};
            bool hit;
            if constexpr (BP == EPP_BP_GRID) {
                hit = world.invalid(wv, p, mode, cpg, md); // records in cell order when the world has them
            } else {
                hit = world.any(wv, p, [&](int k) -> bool {
                    if (RS) return rec_hit(rec + k * kRecDoubles, p);
                    return point_candidate_invalid(wv.aabb, wv.obb, k, p, mode, cpg, md);
                });
            }
            if (hit) atomicOr(&inv[unit >> 5], 1u << (unit & 31));
        }
        __syncwarp();
    }
    // ---- verdicts: bit = valid; bits of units beyond `units` stay 0
    uint32_t bad = 0u;
    if (lane < S && lane * 32 < units) {
        const int live = units - 32 * lane;
        const uint32_t active = live >= 32 ? 0xffffffffu : ((1u << live) - 1u);
        const uint32_t verdict = ~inv[lane] & active;
        bad = inv[lane] & active;
        if (mask_out) mask_out[lane] = verdict;
        if (GA && stage_words) stage_words[lane] = verdict; // sent with its group (k_states_warp)
        inv[lane] = 0u;
    }
    __syncwarp();
    return bad;
}

// Trajectory-matrix form of the warp-pipelined kernel (K3a): the units are the rows of n_traj column-major matrices,
// position = columns 0, 3, 6 (src/PathPlanner.cpp:272-273).  Every trajectory is cut into cpt chunks of S*32 rows;
// its mask words live at sample_mask + t * wpt.
struct TrajArgs {
    const double *mats;
    size_t rows, ld, traj_stride;
    uint32_t wpt, cpt;       // mask words and chunks per trajectory
    int aligned;             // columns start on 16 B and ld / traj_stride are even: cp.async.bulk can move them
    uint32_t *sample_mask;   // may be null
    long long *first_row;    // LLONG_MAX on entry; atomicMin of the first invalid row per trajectory
};

// Block = 1 producer warp + up to 31 consumer warps, one block per SM by default.  Block b owns words
// [b*wpb + min(b, rem), +wpb + (b < rem)) (the host passes wpb = words / blocks and rem = words % blocks, so no
// thread divides) and cuts them into chunks of S words.  Chunks flow through a block-wide ring of n_slots (<= 64)
// shared-memory slots:
//   producer   lane l issues chunks l, l+32, ... in order (cp.async.bulk, `full` barrier of the slot), waiting
//              on the slot's `empty` barrier once the ring has wrapped;
//   consumers  take the next chunk number from a shared counter, wait for its `full` barrier, run the two
//              phases of warp_states_step on it, write its S mask words and release the slot.
// Why a producer warp: an SM's share of HBM bandwidth is ~44 GB/s and its bulk-copy queue is shallow, so
// cp.async.bulk BLOCKS at issue once ~100 KB are queued (measured: a warp that issues two chunks up front
// stalls for microseconds before it can touch the first).  Why a shared counter: chunks arrive over several
// microseconds in issue order; with static per-warp ranges the warps whose data comes last start last and
// finish last while the others idle.  The broad-phase table and the records are requested before the first
// coordinate so they are not queued behind the coordinates.
constexpr int kMaxWarps = 32;
constexpr int kMaxSlots = 64;
// per consumer warp: byte queue (S * 32) + invalid words (32 B); grid worlds: room for the deferred list of up to
// 31 + S * 32 state indices instead (see k_states_warp)
__host__ __device__ constexpr uint32_t warp_area_bytes(int bp, int S) { return bp == EPP_BP_GRID ? (32u + S * 32u) * 4u : S * 32u + 32u; }
constexpr uint32_t kGaStageWords = 1024; // GA: blocks of up to this many mask words send them in groups

__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}

// TR: trajectory-matrix form (TrajArgs above); wpb / rem then count CHUNKS, xyz / n / mask are unused.
template <int BP, int S, bool RS, bool GA, bool TR = false>
__global__ void __launch_bounds__(1024, 1) k_states_warp(WorldView wv, const double *__restrict__ xyz, size_t n, int mode,
                                                      int can_pass_gate, double min_distance, int aabb_in_smem,
                                                      uint32_t n_slots, uint32_t table_bytes, uint32_t rec_bytes,
                                                      uint32_t wpb, uint32_t rem, uint32_t *__restrict__ mask,
                                                      const __grid_constant__ GatherView gv, uint32_t ga_group,
                                                      const __grid_constant__ TrajArgs ta, uint32_t gran)
{
    static_assert(!(TR && GA), "the trajectory form has no fused exchange");
    __shared__ uint32_t s_seq, s_hdr, s_unpacked, s_cnt[kMaxRanks]; // GA: call number; header state (0 none, 1 claimed, 2 read); warps done unpacking
    __shared__ __align__(8) uint64_t s_full[kMaxSlots];
    __shared__ __align__(8) uint64_t s_empty[kMaxSlots];
    __shared__ __align__(8) uint64_t s_table;
    __shared__ uint32_t s_next;
    __shared__ uint32_t s_rounds_done[kMaxSlots]; // per slot: how many rounds of it have been consumed
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t consumers = (blockDim.x >> 5) - 1u;
    const bool cpg = can_pass_gate != 0;
    constexpr uint32_t kSlotBytes = S * kWordBytes;
    const uint32_t slots = n_slots;
    // shared memory: [broad-phase table][records][ring][per consumer: byte queue (S*32) | invalid words (32 B)]
    double *rec = reinterpret_cast<double *>(g_smem + table_bytes);
    unsigned char *ring = g_smem + table_bytes + rec_bytes;
    unsigned char *areas = ring + slots * kSlotBytes;
    // GA, small blocks of words (ga_group > 0): verdict words are parked here and leave for the peers in groups of
    // ga_group chunks (one 8 * S * ga_group byte run per peer instead of ga_group runs of 8 * S bytes: at 8 ranks the
    // exchange is bound by the NUMBER of NVLink writes, not their bytes); behind them one counter per group.
    uint32_t *ga_words = reinterpret_cast<uint32_t *>(areas + ((blockDim.x >> 5) - 1u) * warp_area_bytes(BP, S));
    uint32_t *ga_count = ga_words + kGaStageWords;
    EPP_TRACE_EV(blockIdx.x * consumers + (warp == 0 ? 0 : warp - 1), warp == 0 ? 13 : 14);

    // ownership in units of `gran` words (1; 12 with the fused exchange: whole chunks AND whole 3-word exchange units)
    const uint32_t words_total = (uint32_t)((n + 31) / 32);
    const uint32_t b0 = (blockIdx.x * wpb + (blockIdx.x < rem ? blockIdx.x : rem)) * gran;
    uint32_t b1 = b0 + (wpb + (blockIdx.x < rem ? 1u : 0u)) * gran;
    if (!TR && b1 > words_total) b1 = words_total;
    const uint32_t chunks = TR ? (b1 - b0) : (b1 - b0 + S - 1) / S;
    const bool aligned = TR ? ta.aligned != 0 : (reinterpret_cast<uintptr_t>(xyz) & 15u) == 0;
    // TR: trajectory and first row of chunk i of this block
    auto traj_of = [&](uint32_t i, uint32_t &t, uint32_t &j) {
        const uint32_t gi = b0 + i;
        t = gi / ta.cpt;
        j = gi - t * ta.cpt;
    };
    // states in chunk i of this block
    auto chunk_units = [&](uint32_t i) -> int {
        if (TR) {
            uint32_t t, j;
            traj_of(i, t, j);
            const size_t left = ta.rows - (size_t)j * (S * 32);
            return left < (size_t)(S * 32) ? (int)left : S * 32;
        }
        const uint32_t cw = b0 + i * S;
        const uint32_t sw = (b1 - cw) < (uint32_t)S ? (b1 - cw) : (uint32_t)S;
        int units = (int)(sw * 32u);
        if (cw + sw == words_total) units = (int)(n - (size_t)cw * 32u); // the batch's ragged last word
        return units;
    };
    // cp.async.bulk moves multiples of 16 B from 16 B-aligned addresses: an odd number of states (only
    // possible in the last chunk of the batch) or an 8 B-aligned batch is loaded by the consumer's lanes.
    auto by_tma = [&](int units) -> bool { return aligned && (units & 1) == 0; };

    // thread 0 requests the table while threads 32.. initialise the ring's barriers (one slot each)
    if (threadIdx.x >= 32) {
        for (uint32_t s = threadIdx.x - 32; s < slots; s += blockDim.x - 32) {
            mbar_init(&s_full[s], 1);
            mbar_init(&s_empty[s], 1);
            s_rounds_done[s] = 0u;
        }
        if (GA && ga_group)
            for (uint32_t s = threadIdx.x - 32; s < kGaStageWords / S; s += blockDim.x - 32) ga_count[s] = 0u;
        if (threadIdx.x - 32 < slots) mbar_init_fence();
    }
    if (threadIdx.x == 0) {
        s_next = 0u;
        s_hdr = 0u;
        s_unpacked = 0u;
        mbar_init(&s_table, 1);
        mbar_init_fence();
        const void *tsrc;
        const uint32_t tbytes = (PointWorld<BP>::table(wv, aabb_in_smem != 0, tsrc) + 15u) & ~15u;
        if (tbytes) {
            mbar_expect_tx(&s_table, tbytes);
            bulk_g2s(g_smem, tsrc, tbytes, &s_table);
        }
    }
    EPP_TRACE_EV(blockIdx.x * consumers + (warp == 0 ? 0 : warp - 1), warp == 0 ? 11 : 12);
    // record fields into registers (threads 32 .. 32+M-1, i.e. consumer warps) before any coordinate is requested
    double rv[kRecDoubles];
    const int rk = (int)threadIdx.x - 32;
    const bool have_rec = RS && rk >= 0 && rk < wv.M;
    uint32_t rec_flags = 0u;
    if (have_rec) {
        const AabbRec *a = wv.aabb + rk;
        const ObbRec *o = wv.obb + rk;
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            rv[d] = __ldg(&a->mn[d]);
            rv[3 + d] = __ldg(&a->mx[d]);
            rv[6 + d] = __ldg(&o->center[d]);
            rv[13 + d] = __ldg(&o->half[d]);
        }
        rv[9] = __ldg(&o->rot[0]);
        rv[10] = __ldg(&o->rot[1]);
        rv[11] = __ldg(&o->rot[3]);
        rv[12] = __ldg(&o->rot[4]);
        rv[16] = __ldg(&o->inflate);
        rec_flags = __ldg(&o->flags);
    }
    __syncthreads(); // barriers initialised; the world's requests are ahead of every coordinate request
    // Programmatic dependent launch: everything above touched only this block's shared memory and the
    // immutable world epoch, so it may overlap the tail of the previous kernel in the stream; the batch and
    // the mask may not.  (Without the launch attribute both instructions are no-ops.)
    // late_wait (batches of one multi-batch call, known to be independent of each other and complete before the call):
    // the wait moves to the END of the kernel, so this launch streams its states while the previous one is still in
    // its tail, and still does not complete before it — ordering against later launches is kept transitively.
    const bool late_wait = !GA && ((mode >> 16) & 1) != 0;
    if (!late_wait) asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    if (warp == 0) {
        // ---- producer
        // Lane l owns ring slots l and l + 32 and arms them round after round, in program order, so a lane never
        // waits on a phase of `empty` that lies more than one behind the barrier's current phase (any slot count).
        for (uint32_t round = 0; round * slots + (uint32_t)lane < chunks; ++round)
        for (uint32_t slot = (uint32_t)lane; slot < slots; slot += 32u) {
            const uint32_t i = round * slots + slot;
            if (i >= chunks) break;
            if (round > 0) mbar_wait(&s_empty[slot], (round - 1u) & 1u);
            const int units = chunk_units(i);
            if (by_tma(units)) {
                mbar_expect_tx(&s_full[slot], (uint32_t)units * 24u);
                if (TR) {
                    uint32_t t, j;
                    traj_of(i, t, j);
                    const double *col = ta.mats + (size_t)t * ta.traj_stride + (size_t)j * (S * 32);
#pragma unroll
                    for (int d = 0; d < 3; ++d)
                        bulk_g2s(ring + slot * kSlotBytes + d * (S * 32 * 8), col + (size_t)(3 * d) * ta.ld,
                                 (uint32_t)units * 8u, &s_full[slot]);
                } else {
                    bulk_g2s(ring + slot * kSlotBytes, xyz + (size_t)(b0 + i * S) * 96, (uint32_t)units * 24u, &s_full[slot]);
                }
            } else {
                mbar_arrive(&s_full[slot]); // the consumer loads this chunk with its lanes
            }
        }
        return;
    }

    // ---- consumers
    const uint32_t c = (uint32_t)warp - 1u;
    uint32_t seq = 0;
    uint64_t ll_row = 0;
    if (GA && threadIdx.x == 32) s_seq = (gv.debug & 4) ? 1u : gather_seq(gv); // one read per block, published by the barrier below
    if (GA && gv.deferred && c == 0) {
        // deferred form: the rows of this parity hold call seq - 2 until every peer has acknowledged unpacking them (it
        // did so a step ago unless it lags); one warp per block checks, the barrier below holds the others back
        __syncwarp();
        const uint32_t sq = *(volatile uint32_t *)&s_seq;
        if (sq > 2u && lane < gv.world && lane != gv.rank) {
            bool dead = false;
            ll_wait(gv.my_ll + gather_row(gv, sq, lane) - 1u, sq - 2u, gv, dead);
        }
    }
#ifdef EPP_TRACE
    const uint32_t g = blockIdx.x * consumers + c;
    int tstep = 0;
#endif
    EPP_TRACE_EV(g, 0);
    unsigned char *queue = areas + c * warp_area_bytes(BP, S);
    uint32_t *inv = reinterpret_cast<uint32_t *>(queue + S * 32);
    // Grid worlds, plain state batches: phase 2 (the cell's records, dependent global loads) is DEFERRED — states whose
    // coarse cell is occupied (a few per cent) go on a per-warp list by global index, the chunk's mask words are
    // written as valid and the ring slot is released; whenever 32 states are waiting each lane takes one (coordinates
    // re-read from global memory, fresh in L2) and clears the bit of a state that collides.  Per chunk, phase 2 kept a
    // slot for its whole latency with ~8 of 32 lanes busy.  Not with the fused exchange (words leave for the peers when
    // the chunk ends) and not for trajectory matrices (the first invalid row is reduced per chunk).
    constexpr bool DEFER = BP == EPP_BP_GRID && !GA && !TR;
    uint32_t *defer = reinterpret_cast<uint32_t *>(queue);
    int fill = 0;
    const uint32_t lt = (1u << lane) - 1u;
    auto deferred_round = [&](int first, int count) {
        __syncwarp();
        if constexpr (DEFER) if (lane < count) {
            const uint32_t id = defer[first + lane];
            const double *gp = xyz + (size_t)id * 3;
            const P3 p = {__ldg(gp), __ldg(gp + 1), __ldg(gp + 2)};
            if (world_invalid_grid(wv, p, mode & 0xff, cpg, min_distance)) atomicAnd(mask + (id >> 5), ~(1u << (id & 31)));
        }
        __syncwarp();
    };
    if (!DEFER && lane < 8) inv[lane] = 0u;
    // fold the call's rules into the records (src/World.cpp:92-97, :116-121; src/OBB.cpp:67-72)
    if (have_rec) {
        double *r = rec + rk * kRecDoubles;
        const bool filling = (rec_flags & kFlagFilling) != 0;
        const bool skip = (mode & 0xff) == kModeBool ? (filling && cpg) : filling;
        const double infl = (mode & 0xff) == kModeBool ? rv[16] : min_distance;
#pragma unroll
        for (int d = 0; d < 13; ++d) r[d] = rv[d];
#pragma unroll
        for (int d = 0; d < 3; ++d) r[13 + d] = skip ? -1.0 : (filling ? rv[13 + d] : dadd(rv[13 + d], infl));
        r[16] = 0.0;
    }
    PointWorld<BP> world;
    world.bind(wv, aabb_in_smem != 0);
    EPP_TRACE_EV(g, 1);
    asm volatile("bar.sync 1, %0;" ::"r"(consumers * 32u) : "memory"); // consumers only: publishes the records
    if (GA) {
        seq = s_seq;
        ll_row = gather_row(gv, seq, gv.rank);
        if (blockIdx.x == 0 && c == 0) gather_send_header(gv, seq, lane);
    }
    EPP_TRACE_EV(g, 2);
    {
        const void *tsrc;
        if (PointWorld<BP>::table(wv, aabb_in_smem != 0, tsrc)) mbar_wait(&s_table, 0);
    }
    EPP_TRACE_EV(g, 3);

    for (;;) {
        uint32_t i = 0;
        if (lane == 0) i = atomicAdd(&s_next, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= chunks) break;
        const uint32_t round = i / slots, slot = i - round * slots;
        const int units = chunk_units(i);
        double *pts = reinterpret_cast<double *>(ring + slot * kSlotBytes);
        // A parity wait is only meaningful while the barrier is in phase `round` or `round + 1`.  Chunk numbers
        // are handed out in order but finish out of order, so a fast warp can be handed round r of a slot whose
        // round r-1 is still being loaded or worked on; the explicit count closes that window.
        if (round > 0) {
            uint32_t done;
            do {
                asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(done) : "r"(smem_addr(&s_rounds_done[slot])) : "memory");
            } while (done < round);
        }
        mbar_wait(&s_full[slot], round & 1u);
        uint32_t tr_t = 0, tr_j = 0;
        if (TR) traj_of(i, tr_t, tr_j);
        if (!by_tma(units)) {
            if (TR) {
                const double *col = ta.mats + (size_t)tr_t * ta.traj_stride + (size_t)tr_j * (S * 32);
#pragma unroll
                for (int d = 0; d < 3; ++d)
                    for (int u = lane; u < units; u += 32) pts[d * (S * 32) + u] = __ldg(col + (size_t)(3 * d) * ta.ld + u);
            } else {
                const double *src = xyz + (size_t)(b0 + i * S) * 96;
                for (int e = lane; e < 3 * units; e += 32) pts[e] = __ldg(src + e);
            }
            __syncwarp();
        }
        EPP_TRACE_EV(g, 4 + 2 * tstep);
        const bool staged = GA && ga_group != 0u;
        if (TR) {
            uint32_t *mo = ta.sample_mask ? ta.sample_mask + (size_t)tr_t * ta.wpt + tr_j * S : nullptr;
            const uint32_t bad = warp_states_step<BP, S, RS, false, true>(world, wv, rec, pts, units, mode, cpg, min_distance,
                                                                          queue, inv, mo, lane, gv, 0u, 0u, nullptr);
            // first invalid row of the chunk -> atomicMin into the trajectory's slot (rows < 2^32 - 1, host-checked)
            const uint32_t cand = bad ? tr_j * (S * 32) + (uint32_t)lane * 32u + (uint32_t)(__ffs(bad) - 1) : 0xffffffffu;
            const uint32_t first = __reduce_min_sync(0xffffffffu, cand);
            if (lane == 0 && first != 0xffffffffu) atomicMin(ta.first_row + tr_t, (long long)first);
        } else if constexpr (DEFER) {
#pragma unroll
            for (int u = 0; u < S; ++u) {
                const int unit = u * 32 + lane;
                bool has = false;
                if (unit < units) has = world.has_candidates(wv, P3{pts[3 * unit], pts[3 * unit + 1], pts[3 * unit + 2]});
                const uint32_t bal = __ballot_sync(0xffffffffu, has);
                if (has) defer[fill + __popc(bal & lt)] = (b0 + i * S) * 32u + (uint32_t)unit;
                fill += __popc(bal);
            }
            if (lane < S && lane * 32 < units) {
                const int live = units - 32 * lane;
                mask[b0 + i * S + lane] = live >= 32 ? 0xffffffffu : ((1u << live) - 1u);
            }
            __syncwarp();
        } else {
            warp_states_step<BP, S, RS, GA>(world, wv, rec, pts, units, mode, cpg, min_distance, queue, inv,
                                            mask + b0 + i * S, lane, gv, seq, ll_row + b0 + i * S,
                                            staged ? ga_words + i * S : nullptr);
        }
        if (staged) {
            // count the chunk into its group; the warp that completes a group sends it
            const uint32_t grp = i / ga_group, g0 = grp * ga_group;
            const uint32_t g_chunks = (chunks - g0) < ga_group ? (chunks - g0) : ga_group;
            uint32_t old = 0;
            if (lane == 0) {
                __threadfence_block();
                old = atomicAdd(&ga_count[grp], 1u);
            }
            old = __shfl_sync(0xffffffffu, old, 0);
            if (old + 1u == g_chunks && !(gv.debug & 2)) {
                __threadfence_block();
                // the group's words as quads: b0 and the group's first word are multiples of 12, so unit = word / 3
                const uint32_t w0 = g0 * S;
                const uint32_t w1 = (w0 + g_chunks * S) < (b1 - b0) ? (w0 + g_chunks * S) : (b1 - b0);
                const volatile uint32_t *gw = ga_words;
                for (int p = 0; p < gv.world; ++p) {
                    if (p == gv.rank) continue;
                    uint4 *dst = gv.peer_ll[p] + ll_row + (b0 + w0) / 3u;
                    for (uint32_t w = w0 + 3u * (uint32_t)lane; w < w1; w += 96u)
                        ll_store(dst + (w - w0) / 3u, gw[w], w + 1u < w1 ? gw[w + 1u] : 0u, w + 2u < w1 ? gw[w + 2u] : 0u, seq);
                }
            }
        }
        EPP_TRACE_EV(g, 5 + 2 * tstep);
#ifdef EPP_TRACE
        if (tstep < 3) ++tstep;
#endif
        // the step ended with __syncwarp: every lane's reads of the slot are done, the producer may refill it
        if (lane == 0) {
            mbar_arrive(&s_empty[slot]);
            asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(smem_addr(&s_rounds_done[slot])), "r"(round + 1u) : "memory");
        }
        if constexpr (DEFER) {
            while (fill >= 32) {
                fill -= 32;
                deferred_round(fill, 32);
            }
        }
    }
    if constexpr (DEFER) {
        if (fill > 0) deferred_round(0, fill);
    }
    if (late_wait) asm volatile("griddepcontrol.wait;" ::: "memory");
    EPP_TRACE_EV(g, 15);
    if (GA) {
        // Out of states: take this warp's share of the peers' slots.  Block b owns units
        // [b, b + 1) * slot_units / blocks of every peer's slot, its consumer warps split that range.
        const uint64_t slot_units = (gv.slot_words + 2u) / 3u; // exchange units (quads) per slot
        const uint64_t per_b = (slot_units + gridDim.x - 1) / gridDim.x;
        const uint64_t per_w = (per_b + consumers - 1) / consumers;
        uint64_t e = ((uint64_t)blockIdx.x + 1u) * per_b;
        e = e < slot_units ? e : slot_units;
        uint64_t u0 = (uint64_t)blockIdx.x * per_b + (uint64_t)c * per_w;
        u0 = u0 < e ? u0 : e;
        const uint64_t u1 = u0 + per_w < e ? u0 + per_w : e;
        // Deferred form (epp_gather_set_deferred): the peers' words of THIS call are not waited for; the words of the
        // previous call, which arrived a step ago, are unpacked into that call's buffer instead.
        const uint32_t useq = gv.deferred ? seq - 1u : seq;
        uint32_t *uout = gv.deferred ? gv.prev_out : gv.out;
        if (gv.deferred) { // own slot beyond this rank's units reads 0 (the immediate form does it inside the unpack)
            uint32_t *own = gv.out + (uint64_t)gv.rank * gv.slot_words;
            const uint64_t w1 = 3u * u1 < gv.slot_words ? 3u * u1 : gv.slot_words;
            for (uint64_t w = (3u * u0 > gv.my_words ? 3u * u0 : gv.my_words) + (uint64_t)lane; w < w1; w += 32u) own[w] = 0u;
        }
        if (!(gv.debug & 1) && useq != 0u && uout != nullptr) {
            // the first warp of the block to get here reads the peers' headers for all of them
            uint32_t st = 0;
            if (lane == 0) st = atomicCAS(&s_hdr, 0u, 1u);
            st = __shfl_sync(0xffffffffu, st, 0);
            if (st == 0u) {
                gather_read_headers(gv, useq, s_cnt, lane);
                __syncwarp();
                if (lane == 0) asm volatile("st.release.cta.shared.u32 [%0], 2;" ::"r"(smem_addr(&s_hdr)) : "memory");
            } else {
                while (st != 2u) asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(st) : "r"(smem_addr(&s_hdr)) : "memory");
            }
            __syncwarp();
            gather_unpack(gv, useq, s_cnt, u0, u1, lane, uout, !gv.deferred);
            if (gv.deferred) {
                // the warp that finishes last tells every peer that its rows of call useq may be overwritten
                // (counted per block in shared memory first: thousands of atomics on one word serialise in L2)
                __threadfence();
                uint32_t last = 0;
                if (lane == 0 && atomicAdd(&s_unpacked, 1u) == consumers - 1u) last = atomicAdd(&gv.ctl[3], 1u) == gridDim.x - 1u ? 1u : 0u;
                last = __shfl_sync(0xffffffffu, last, 0);
                if (last) {
                    if (lane == 0) gv.ctl[3] = 0u;
                    if (lane < gv.world && lane != gv.rank)
                        ll_store(gv.peer_ll[lane] + gather_row(gv, useq, gv.rank) - 1u, useq, 0u, 0u, useq);
                }
            }
        }
    }
}

// ---- K2: segments --------------------------------------------------------------------------

template <int BP>
__global__ void __launch_bounds__(kBlock) k_segments(WorldView wv, const double *__restrict__ pa,
                                                     const double *__restrict__ pb, size_t n, int can_pass_gate,
                                                     int aabb_in_smem, uint32_t *__restrict__ mask)
{
    RayWorld<BP> world;
    world.init(wv, aabb_in_smem != 0);
    const size_t tiles = (n + kBlock - 1) / kBlock;
    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const size_t i = tile * kBlock + threadIdx.x;
        bool valid = false;
        if (i < n) {
            const P3 a = {__ldg(pa + 3 * i), __ldg(pa + 3 * i + 1), __ldg(pa + 3 * i + 2)};
            const P3 b = {__ldg(pb + 3 * i), __ldg(pb + 3 * i + 1), __ldg(pb + 3 * i + 2)};
            valid = !world.invalid(wv, a, b, can_pass_gate != 0);
        }
        const uint32_t word = __ballot_sync(0xffffffffu, valid);
        if ((threadIdx.x & 31) == 0 && i < n) mask[i >> 5] = word;
    }
}

// ---- K2 with the world's records in shared memory ------------------------------------------------------------
//
// ncu on the form above (256-OBB world): FP64 pipe 5 % busy, 25 % of issue slots used, the rest is waiting on
// the per-lane record fetches of the candidate loop — every lane reads a DIFFERENT AabbRec / ObbRec, so one load
// instruction touches up to 32 cache lines.  For worlds of up to kRayRecMax OBBs whose rotations are all about z
// (the only kind World::addGate / addObstacle build, src/Object.cpp:38-47) the prologue folds the call's rules
// into one 21-double record per OBB in shared memory (odd stride: lanes on different records hit different
// banks) and the next tile's end points are fetched while the current tile is evaluated.
// Record: [0..5] stored AABB min / max (min.x = +inf for a box the call skips: filling && canPassGate,
// src/World.cpp:150-153), [6..8] centre, [9..12] R(0,0) R(0,1) R(1,0) R(1,1), [13..15] half sizes of the END
// POINT tests (half + inflate for collision boxes, half for filling ones, src/OBB.cpp:67-72), [16..18] half sizes
// of the SLAB test (half + inflate for every type, :28).  With R(2,0) = R(2,1) = R(0,2) = R(1,2) = 0 and
// R(2,2) = 1 the reference's 3-term sums are (a + b) + (+-0) and (+-0 + +-0) + d2: dropping the zero terms can
// only change the sign of a zero, which fabs() and the comparisons cannot see.
constexpr int kRayRec = 21;
constexpr int kRayRecMax = 256;

__device__ __forceinline__ bool ray_rec_invalid(const double *__restrict__ r, const P3 &a, const P3 &b, const P3 &mn,
                                                const P3 &mx)
{
    // z first: the bits rows only look at x and y, so z is where most candidates leave
    if (mx.z < r[2] || mn.z > r[5] || mx.x < r[0] || mn.x > r[3] || mx.y < r[1] || mn.y > r[4]) return false;
    const double a0 = dsub(a.x, r[6]), a1 = dsub(a.y, r[7]), a2 = dsub(a.z, r[8]);
    const double b0 = dsub(b.x, r[6]), b1 = dsub(b.y, r[7]), b2 = dsub(b.z, r[8]);
    const double ax = dadd(dmul(r[9], a0), dmul(r[11], a1)), ay = dadd(dmul(r[10], a0), dmul(r[12], a1));
    const double bx = dadd(dmul(r[9], b0), dmul(r[11], b1)), by = dadd(dmul(r[10], b0), dmul(r[12], b1));
    // end points first (src/OBB.cpp:13-18)
    if (fabs(ax) <= r[13] && fabs(ay) <= r[14] && fabs(a2) <= r[15]) return true;
    if (fabs(bx) <= r[13] && fabs(by) <= r[14] && fabs(b2) <= r[15]) return true;
    double t_min = 0.0, t_max = 1.0;
    if (!slab_axis(dsub(bx, ax), ax, r[16], t_min, t_max)) return false;
    if (!slab_axis(dsub(by, ay), ay, r[17], t_min, t_max)) return false;
    if (!slab_axis(dsub(b2, a2), a2, r[18], t_min, t_max)) return false;
    return 0.0 <= t_min && t_min <= 1.0 && 0.0 <= t_max && t_max <= 1.0;
}

// The call's rules folded into one kRayRec-double record per OBB (layout above), by all threads of the block.
__device__ __forceinline__ void fold_ray_records(const WorldView &wv, double *rec, bool can_pass_gate)
{
    for (int k = threadIdx.x; k < wv.M; k += blockDim.x) {
        const AabbRec *ab = wv.aabb + k;
        const ObbRec *o = wv.obb + k;
        double *r = rec + k * kRayRec;
        const uint32_t flags = __ldg(&o->flags);
        const bool filling = (flags & kFlagFilling) != 0;
        const double infl = __ldg(&o->inflate);
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            r[d] = __ldg(&ab->mn[d]);
            r[3 + d] = __ldg(&ab->mx[d]);
            r[6 + d] = __ldg(&o->center[d]);
            const double h = __ldg(&o->half[d]);
            r[13 + d] = filling ? h : dadd(h, infl);
            r[16 + d] = dadd(h, infl);
        }
        if (filling && can_pass_gate) r[0] = __longlong_as_double(0x7ff0000000000000ll);
        r[9] = __ldg(&o->rot[0]);
        r[10] = __ldg(&o->rot[1]);
        r[11] = __ldg(&o->rot[3]);
        r[12] = __ldg(&o->rot[4]);
        r[19] = r[20] = 0.0;
    }
}

template <int BP, int TB>
__global__ void __launch_bounds__(TB, TB >= 1024 ? 1 : (TB >= 512 ? 2 : 4)) k_segments_rs(WorldView wv, const double *__restrict__ pa,
                                                            const double *__restrict__ pb, size_t n, int can_pass_gate,
                                                            uint32_t table_bytes, uint32_t *__restrict__ mask)
{
    RayWorld<BP> world;
    world.init(wv, false);
    double *rec = reinterpret_cast<double *>(g_smem + table_bytes);
    fold_ray_records(wv, rec, can_pass_gate != 0);
    __syncthreads();
    const size_t tiles = (n + TB - 1) / TB;
    size_t tile = blockIdx.x;
    if (tile >= tiles) return;
    size_t i = tile * TB + threadIdx.x;
    P3 a = {0.0, 0.0, 0.0}, b = {0.0, 0.0, 0.0};
    if (i < n) {
        a = {__ldg(pa + 3 * i), __ldg(pa + 3 * i + 1), __ldg(pa + 3 * i + 2)};
        b = {__ldg(pb + 3 * i), __ldg(pb + 3 * i + 1), __ldg(pb + 3 * i + 2)};
    }
    for (;;) {
        const size_t next_tile = tile + gridDim.x;
        const size_t ni = next_tile * TB + threadIdx.x;
        P3 na = {0.0, 0.0, 0.0}, nb = {0.0, 0.0, 0.0};
        if (next_tile < tiles && ni < n) {
            na = {__ldg(pa + 3 * ni), __ldg(pa + 3 * ni + 1), __ldg(pa + 3 * ni + 2)};
            nb = {__ldg(pb + 3 * ni), __ldg(pb + 3 * ni + 1), __ldg(pb + 3 * ni + 2)};
        }
        bool valid = false;
        if (i < n) {
            P3 mn, mx;
            ray_box(a, b, mn, mx);
            valid = !world.any(wv, mn, mx, [&](int k) -> bool { return ray_rec_invalid(rec + k * kRayRec, a, b, mn, mx); });
        }
        const uint32_t word = __ballot_sync(0xffffffffu, valid);
        if ((threadIdx.x & 31) == 0 && i < n) mask[i >> 5] = word;
        if (next_tile >= tiles) break;
        tile = next_tile;
        i = ni;
        a = na;
        b = nb;
    }
}

// ---- K2 (warp-pipelined form): end points through a cp.async.bulk ring ---------------------------------------
//
// The direct forms above fetch a segment's six coordinates with scalar loads at a 24-byte stride (ncu: 3.5 L1 sectors
// per algorithmic sector) and, on the grid broad phase, walk cell lists of very different lengths lane by lane.  This
// form has K1's structure: block = 1 producer warp + up to 31 consumer warps per SM; the block's mask words are cut into
// chunks of kSegS words (64 segments); the producer brings a chunk's two end-point runs (2 x 1536 B) into a ring slot
// with two bulk copies on one mbarrier; consumers take chunk numbers from a shared counter.
//   bits / brute: every lane evaluates its kSegS segments from shared memory (records folded as in k_segments_rs for
//                 worlds of <= 256 z-rotated OBBs).  Compacting (segment, OBB) pairs into dense warp-wide steps was
//                 built twice (round 1; round 2 with survivors deferred across chunks) and was 4 % SLOWER both times:
//                 lane utilisation rose from 13 to 24 of 32 but the queue bookkeeping cost more than it saved;
//   grid:         phase 1 asks the occupancy bitmap (shared memory) whether any coarse cell under the ray box is
//                 occupied and compacts those segments (~1 in 8 in the 10 000-OBB world) into a per-warp queue,
//                 phase 2 walks their cell lists one queued segment per lane.
constexpr int kSegS = 2;
constexpr int kSegSlotBytes = kSegS * 2 * kWordBytes; // both end points of kSegS * 32 segments
constexpr int kSegMaxSlots = 64;
constexpr int kSegWarpBytes = 96 * 4; // per consumer warp: deferred list of up to 31 + kSegS * 32 segment indices (grid)

template <int BP, bool RS>
__global__ void __launch_bounds__(1024, 1) k_segments_warp(WorldView wv, const double *__restrict__ pa,
                                                          const double *__restrict__ pb, size_t n, int can_pass_gate,
                                                          int aabb_in_smem, uint32_t slots, uint32_t table_bytes,
                                                          uint32_t rec_bytes, uint32_t wpb, uint32_t rem,
                                                          int late_wait, uint32_t *__restrict__ mask)
{
    constexpr int S = kSegS;
    __shared__ __align__(8) uint64_t s_full[kSegMaxSlots];
    __shared__ __align__(8) uint64_t s_empty[kSegMaxSlots];
    __shared__ __align__(8) uint64_t s_table;
    __shared__ uint32_t s_next;
    __shared__ uint32_t s_rounds_done[kSegMaxSlots];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t consumers = (blockDim.x >> 5) - 1u;
    const bool cpg = can_pass_gate != 0;
    double *rec = reinterpret_cast<double *>(g_smem + table_bytes);
    unsigned char *ring = g_smem + table_bytes + rec_bytes;
    unsigned char *areas = ring + slots * kSegSlotBytes;

    const uint32_t b0 = blockIdx.x * wpb + (blockIdx.x < rem ? blockIdx.x : rem);
    const uint32_t b1 = b0 + wpb + (blockIdx.x < rem ? 1u : 0u);
    const uint32_t chunks = (b1 - b0 + S - 1) / S;
    const uint32_t words_total = (uint32_t)((n + 31) / 32);
    const bool aligned = ((reinterpret_cast<uintptr_t>(pa) | reinterpret_cast<uintptr_t>(pb)) & 15u) == 0;
    auto chunk_units = [&](uint32_t i) -> int {
        const uint32_t cw = b0 + i * S;
        const uint32_t sw = (b1 - cw) < (uint32_t)S ? (b1 - cw) : (uint32_t)S;
        int units = (int)(sw * 32u);
        if (cw + sw == words_total) units = (int)(n - (size_t)cw * 32u);
        return units;
    };
    auto by_tma = [&](int units) -> bool { return aligned && (units & 1) == 0; };

    // The broad-phase table and the folded records come by bulk copy (one thread, two instructions): a block-wide
    // copy loop + record fold behind a barrier was ~3 us of every 30 us launch (ncu r02_e: the copy loop had the most
    // stall samples of the kernel), and it cannot overlap the previous launch because two blocks do not fit one SM.
    RayWorld<BP> world;
    const void *tsrc = nullptr;
    uint32_t tbytes = 0;
    if (BP == EPP_BP_GRID) {
        tsrc = wv.occ;
        tbytes = wv.occ_bytes;
    } else if (BP == EPP_BP_BITS && wv.range_d > 0) {
        tsrc = wv.bits_range;
        tbytes = (uint32_t)(2 * wv.nb * wv.range_d * wv.Ws * 4 + 15) & ~15u;
    }
    const bool recs_by_tma = RS && wv.ray_recs != nullptr;
    if (threadIdx.x == 0) {
        s_next = 0u;
        mbar_init(&s_table, 1);
        mbar_init_fence();
        const uint32_t rbytes = recs_by_tma ? (uint32_t)(wv.M * kRayRec * 8 + 15) & ~15u : 0u;
        if (tbytes + rbytes) mbar_expect_tx(&s_table, tbytes + rbytes);
        if (tbytes) bulk_g2s(g_smem, tsrc, tbytes, &s_table);
        if (rbytes) bulk_g2s(rec, wv.ray_recs + (size_t)(cpg ? 1 : 0) * wv.M * kRayRec, rbytes, &s_table);
    }
    if (BP != EPP_BP_GRID && !tbytes) world.init(wv, aabb_in_smem != 0); // brute / bits without range rows: block-wide copy
    if (RS && !recs_by_tma) fold_ray_records(wv, rec, cpg);
    if (threadIdx.x >= 32) {
        for (uint32_t s = threadIdx.x - 32; s < slots; s += blockDim.x - 32) {
            mbar_init(&s_full[s], 1);
            mbar_init(&s_empty[s], 1);
            s_rounds_done[s] = 0u;
        }
        if (threadIdx.x - 32 < slots) mbar_init_fence();
    }
    __syncthreads();
    // late_wait: batches of one multi-batch call (independent of each other, inputs complete before the call) — the
    // dependency wait moves to the END of the kernel, see k_states_warp
    if (!late_wait) asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    if (warp == 0) {
        // ---- producer: lane l owns ring slots l and l + 32 and arms them round after round, in program order (a
        // parity wait on `empty` is only meaningful one phase behind the barrier, so a slot never changes hands)
        for (uint32_t round = 0; round * slots + (uint32_t)lane < chunks; ++round) {
            for (uint32_t slot = (uint32_t)lane; slot < slots; slot += 32u) {
                const uint32_t i = round * slots + slot;
                if (i >= chunks) break;
                if (round > 0) mbar_wait(&s_empty[slot], (round - 1u) & 1u);
                const int units = chunk_units(i);
                if (by_tma(units)) {
                    unsigned char *dst = ring + slot * kSegSlotBytes;
                    const size_t off = (size_t)(b0 + i * S) * 96;
                    mbar_expect_tx(&s_full[slot], (uint32_t)units * 48u);
                    bulk_g2s(dst, pa + off, (uint32_t)units * 24u, &s_full[slot]);
                    bulk_g2s(dst + S * kWordBytes, pb + off, (uint32_t)units * 24u, &s_full[slot]);
                } else {
                    mbar_arrive(&s_full[slot]);
                }
            }
        }
        return;
    }

    // ---- consumers
    const uint32_t c = (uint32_t)warp - 1u;
    if (tbytes) world.bind(wv);
    if (tbytes || recs_by_tma) mbar_wait(&s_table, 0);
    __syncwarp();
    const uint32_t lt = (1u << lane) - 1u;
    // Grid: phase 2 (the walk through the cell lists, a chain of dependent global loads) is DEFERRED.  Segments with an
    // occupied cell under their box (~1 in 8) are listed by global index; when 32 are waiting, each lane takes one
    // (end points re-read from global memory: they were streamed through L2 a moment ago) and clears the mask bit of
    // a segment that collides.  Doing phase 2 per 64-segment chunk kept a ring slot for its whole latency with ~8 of
    // 32 lanes busy (ncu r02_e: 13.6 active threads per instruction, consumers waiting for slots).
    uint32_t *defer = reinterpret_cast<uint32_t *>(areas + c * kSegWarpBytes);
    int fill = 0;
    auto deferred_round = [&](int first, int count) {
        __syncwarp();
        if constexpr (BP == EPP_BP_GRID) if (lane < count) {
            const uint32_t id = defer[first + lane];
            const double *ga = pa + (size_t)id * 3, *gb = pb + (size_t)id * 3;
            const P3 a = {__ldg(ga), __ldg(ga + 1), __ldg(ga + 2)};
            const P3 b = {__ldg(gb), __ldg(gb + 1), __ldg(gb + 2)};
            P3 mn, mx;
            ray_box(a, b, mn, mx);
            const bool hit = wv.cell_recs ? world.walk_recs(wv, a, b, mn, mx, cpg)
                                          : world.any(wv, mn, mx, [&](int k) -> bool {
                                                return ray_candidate_invalid(wv.aabb, wv.obb, k, a, b, mn, mx, cpg);
                                            });
            if (hit) atomicAnd(mask + (id >> 5), ~(1u << (id & 31)));
        }
        __syncwarp();
    };
    for (;;) {
        uint32_t i = 0;
        if (lane == 0) i = atomicAdd(&s_next, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= chunks) break;
        const uint32_t slot = i % slots, round = i / slots;
        const int units = chunk_units(i);
        double *A = reinterpret_cast<double *>(ring + slot * kSegSlotBytes);
        double *B = A + S * 96;
        if (round > 0) { // see k_states_warp: the parity wait below is only meaningful once round - 1 has been consumed
            uint32_t done;
            do {
                asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(done) : "r"(smem_addr(&s_rounds_done[slot])) : "memory");
            } while (done < round);
        }
        mbar_wait(&s_full[slot], round & 1u);
        if (!by_tma(units)) {
            const size_t off = (size_t)(b0 + i * S) * 96;
            for (int e = lane; e < 3 * units; e += 32) {
                A[e] = __ldg(pa + off + e);
                B[e] = __ldg(pb + off + e);
            }
            __syncwarp();
        }
        uint32_t *mo = mask + b0 + i * S;
        if constexpr (BP == EPP_BP_GRID) {
            // phase 1: which of my segments have an occupied coarse cell under their box; their global indices go on
            // the warp's deferred list, the mask words are written as if they were valid, the slot is released
#pragma unroll
            for (int u = 0; u < S; ++u) {
                const int unit = u * 32 + lane;
                bool has = false;
                if (unit < units) {
                    const P3 a = {A[3 * unit], A[3 * unit + 1], A[3 * unit + 2]};
                    const P3 b = {B[3 * unit], B[3 * unit + 1], B[3 * unit + 2]};
                    P3 mn, mx;
                    ray_box(a, b, mn, mx);
                    has = world.has_candidates(wv, mn, mx);
                }
                const uint32_t bal = __ballot_sync(0xffffffffu, has);
                if (has) defer[fill + __popc(bal & lt)] = (b0 + i * S) * 32u + (uint32_t)unit;
                fill += __popc(bal);
            }
            if (lane < S && lane * 32 < units) {
                const int live = units - 32 * lane;
                mo[lane] = live >= 32 ? 0xffffffffu : ((1u << live) - 1u);
            }
            __syncwarp();
        } else {
            uint32_t mine = 0u;
#pragma unroll
            for (int u = 0; u < S; ++u) {
                const int unit = u * 32 + lane;
                bool valid = false;
                if (unit < units) {
                    const P3 a = {A[3 * unit], A[3 * unit + 1], A[3 * unit + 2]};
                    const P3 b = {B[3 * unit], B[3 * unit + 1], B[3 * unit + 2]};
                    P3 mn, mx;
                    ray_box(a, b, mn, mx);
                    valid = !world.any(wv, mn, mx, [&](int k) -> bool {
                        if (RS) return ray_rec_invalid(rec + k * kRayRec, a, b, mn, mx);
                        return ray_candidate_invalid(wv.aabb, wv.obb, k, a, b, mn, mx, cpg);
                    });
                }
                const uint32_t w = __ballot_sync(0xffffffffu, valid);
                if (lane == u) mine = w;
            }
            if (lane < S && lane * 32 < units) mo[lane] = mine;
            __syncwarp();
        }
        if (lane == 0) {
            mbar_arrive(&s_empty[slot]);
            asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(smem_addr(&s_rounds_done[slot])), "r"(round + 1u) : "memory");
        }
        if constexpr (BP == EPP_BP_GRID) {
            while (fill >= 32) { // phase 2 on full rounds only: 32 busy lanes, no ring slot held meanwhile
                fill -= 32;
                deferred_round(fill, 32);
            }
        }
    }
    if constexpr (BP == EPP_BP_GRID) {
        if (fill > 0) deferred_round(0, fill);
    }
    if (late_wait) asm volatile("griddepcontrol.wait;" ::: "memory");
}

// ---- K3a: pre-sampled trajectory matrices ----------------------------------------------------

__global__ void k_fill_i64(int64_t *p, size_t n, int64_t v)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// blockIdx.x = trajectory, blockIdx.y = tile of kBlock rows.  Position = columns 0, 3, 6
// (src/PathPlanner.cpp:272-273).  first_row holds LLONG_MAX on entry.
template <int BP>
__global__ void __launch_bounds__(kBlock) k_traj_samples(WorldView wv, const double *__restrict__ mats, size_t rows,
                                                         size_t ld, size_t traj_stride, double min_distance,
                                                         int aabb_in_smem, uint32_t *__restrict__ sample_mask,
                                                         long long *__restrict__ first_row)
{
    PointWorld<BP> world;
    world.init(wv, aabb_in_smem != 0);
    const size_t t = blockIdx.x;
    const size_t r = (size_t)blockIdx.y * kBlock + threadIdx.x;
    const double *m = mats + t * traj_stride;
    bool valid = false;
    if (r < rows) {
        const P3 p = {__ldg(m + r), __ldg(m + 3 * ld + r), __ldg(m + 6 * ld + r)};
        valid = !world.invalid(wv, p, kModeMinDist, false, min_distance);
    }
    const uint32_t word = __ballot_sync(0xffffffffu, valid);
    const uint32_t active = __ballot_sync(0xffffffffu, r < rows);
    if ((threadIdx.x & 31) == 0 && active) {
        const size_t words_per_traj = (rows + 31) / 32;
        if (sample_mask) sample_mask[t * words_per_traj + (r >> 5)] = word;
        const uint32_t bad = active & ~word;
        if (bad) atomicMin(first_row + t, (long long)(r + (size_t)(__ffs(bad) - 1)));
    }
}

// first_row: LLONG_MAX -> -1 (valid); one validity bit per trajectory.
__global__ void k_traj_finalize(long long *first_row, int64_t *first_out, size_t n_traj,
                                uint32_t *__restrict__ traj_valid)
{
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = false;
    if (t < n_traj) {
        const long long f = first_row[t];
        valid = (f == LLONG_MAX);
        if (first_out) first_out[t] = valid ? -1 : (int64_t)f;
    }
    const uint32_t word = __ballot_sync(0xffffffffu, valid);
    if ((threadIdx.x & 31) == 0 && (t - (threadIdx.x & 31)) < n_traj) traj_valid[t >> 5] = word;
}

// ---- K3b: fused polynomial sweep --------------------------------------------------------------

constexpr int kPolyThreads = 256;  // 1 producer warp + 7 consumer warps
constexpr int kPolyRingMax = 8;    // rounds of parked samples per block: at most (LaunchCfg::poly_ring, default 4)
constexpr int kPolyBars = 16;      // "round published" barriers, indexed by round % 16 (more than the consumers' spread)
constexpr int kPolyPitch = 33;     // odd pitch: conflict-free column writes and row reads

// Polynomial::evaluate(t, 0) (polynomial.h:136-150): Horner from the top coefficient.
__device__ __forceinline__ double horner(const double *__restrict__ c, int n, double t)
{
    double r = __ldg(c + n - 1);
    for (int j = n - 2; j >= 0; --j) r = dadd(dmul(r, t), __ldg(c + j));
    return r;
}

// N = 10 (trajectory_generator.cpp:17), the only degree the reference builds: fully unrolled, coefficients fetched
// two at a time (a segment's 30 coefficients start on a 16-byte boundary when the batch does).
__device__ __forceinline__ double horner10(const double *__restrict__ c, double t)
{
    const double2 *c2 = reinterpret_cast<const double2 *>(c);
    const double2 c89 = __ldg(c2 + 4), c67 = __ldg(c2 + 3), c45 = __ldg(c2 + 2), c23 = __ldg(c2 + 1), c01 = __ldg(c2);
    double r = c89.y;
    r = dadd(dmul(r, t), c89.x);
    r = dadd(dmul(r, t), c67.y);
    r = dadd(dmul(r, t), c67.x);
    r = dadd(dmul(r, t), c45.y);
    r = dadd(dmul(r, t), c45.x);
    r = dadd(dmul(r, t), c23.y);
    r = dadd(dmul(r, t), c23.x);
    r = dadd(dmul(r, t), c01.y);
    r = dadd(dmul(r, t), c01.x);
    return r;
}

// A block owns G <= 32 trajectories.  Its PRODUCER warp runs the reference's sequential sampler
// (trajectory.cpp:121-140: running sums tin += dt, acc += dt — k * dt would change bits), lane l on trajectory l, 32
// samples per trajectory and round, and parks (segment, time in segment) in a ring of `ring` rounds in shared memory.
// The CONSUMER warps take (round, trajectory) items from a shared counter and evaluate the 32 parked samples of the
// item, one per lane (3 Horner chains + the world check), into one mask word.  Round 1 had the sampler and the
// evaluation in the same warp: with the G = 2 trajectories per warp that filled the GPU, 2 of 32 lanes worked in the
// sampler and it was a third of all issued instructions (ncu r02_a); with G = 32 there were 313 warps in all.
// A trajectory with more than max_samples samples is reported INVALID with first row max_samples (the first row that
// was not checked) unless an earlier row is invalid — never silently valid.
template <int BP, bool N10, int TB, int MB = (TB >= 512 ? 2 : 5)>
__global__ void __launch_bounds__(TB, MB) k_poly_traj(WorldView wv, const double *__restrict__ coeffs,
                                                            const double *__restrict__ durations,
                                                            const double *__restrict__ dts, size_t n_traj,
                                                            int n_segments, int n_coeffs, double min_distance,
                                                            size_t max_samples, int aabb_in_smem, int G,
                                                            uint32_t table_bytes, int ring, int sleep_wait,
                                                            uint32_t *__restrict__ sample_mask,
                                                            uint32_t *__restrict__ n_samples_out,
                                                            long long *__restrict__ first_row)
{
    __shared__ uint32_t s_next, s_produced, s_total, s_done[kPolyRingMax];
    __shared__ uint64_t s_pub[kPolyBars];
    PointWorld<BP> world;
    if (threadIdx.x == 0) {
        s_next = 0u;
        s_produced = 0u;
        s_total = 0xffffffffu;
        for (int r = 0; r < kPolyRingMax; ++r) s_done[r] = 0u;
        for (int r = 0; r < kPolyBars; ++r) mbar_init(&s_pub[r], 1);
    }
    world.init(wv, aabb_in_smem != 0);
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t traj0 = (size_t)blockIdx.x * (size_t)G;
    const uint32_t Gb = (uint32_t)((n_traj - traj0) < (size_t)G ? (n_traj - traj0) : (size_t)G); // trajectories of this block
    const size_t words_per_traj = max_samples / 32;
    // ring behind the broad-phase table: per round G rows of parked times, G rows of segment numbers, G counts
    const size_t seg_bytes = ((size_t)G * kPolyPitch * 2 + 7) & ~(size_t)7;
    const size_t round_bytes = ((size_t)G * kPolyPitch * 8 + seg_bytes + (size_t)G * 4 + 15) & ~(size_t)15;
    auto ring_t = [&](uint32_t slot) { return reinterpret_cast<double *>(g_smem + table_bytes + slot * round_bytes); };
    auto ring_seg = [&](uint32_t slot) {
        return reinterpret_cast<unsigned short *>(g_smem + table_bytes + slot * round_bytes + (size_t)G * kPolyPitch * 8);
    };
    auto ring_cnt = [&](uint32_t slot) {
        return reinterpret_cast<uint32_t *>(g_smem + table_bytes + slot * round_bytes + (size_t)G * kPolyPitch * 8 + seg_bytes);
    };

    if (warp == 0) {
        // ---- producer: sampler state of my trajectory, evaluateRange(0, maxTime, dt), trajectory.cpp:95-119
        const size_t my = traj0 + lane;
        const bool have = (uint32_t)lane < Gb;
        const double *dur = durations + my * n_segments;
        double t_end = 0.0, dt = 0.0, acc = 0.0, tin = 0.0, seg_T = 0.0;
        int seg = 0;
        bool done = !have;
        if (have) {
            dt = __ldg(dts + my);
            for (int k = 0; k < n_segments; ++k) t_end = dadd(t_end, __ldg(dur + k)); // trajectory.h:64-71
            // start-segment search with t_start = 0 (:99-114)
            double a = 0.0;
            int k = 0;
            for (k = 0; k < n_segments; ++k) {
                a = dadd(a, __ldg(dur + k));
                if (a > 0.0) break;
            }
            if (0.0 > a || k >= n_segments) {
                done = true;
            } else {
                seg = k;
                seg_T = __ldg(dur + seg);
                acc = dsub(a, seg_T);      // :117
                tin = dsub(0.0, acc);      // :118
            }
        }
        uint32_t emitted_total = 0;
        // One step of the sampler's loop up to (not including) the emission: returns true when a sample is due.
        auto advance = [&]() -> bool {
            while (!done) {
                if (!(acc < t_end)) { // :121
                    done = true;
                    break;
                }
                if (tin > seg_T) { // :122-130
                    tin = dsub(tin, seg_T);
                    ++seg;
                    if (seg >= n_segments) {
                        done = true;
                        break;
                    }
                    seg_T = __ldg(dur + seg);
                    continue;
                }
                return true;
            }
            return false;
        };
        uint32_t round = 0;
        for (; round < (uint32_t)words_per_traj; ++round) {
            const uint32_t slot = round % (uint32_t)ring;
            if (round >= (uint32_t)ring) { // all Gb items of the slot's previous round have been evaluated
                const uint32_t need = Gb * (round / (uint32_t)ring);
                for (;;) {
                    uint32_t d;
                    asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(d) : "r"(smem_addr(&s_done[slot])) : "memory");
                    if (d >= need) break;
                    if (sleep_wait) __nanosleep(100); // the ring is full: nothing is waiting for this warp
                }
            }
            double *st = ring_t(slot) + lane * kPolyPitch;
            unsigned short *sg = ring_seg(slot) + lane * kPolyPitch;
            int cnt = 0;
            // The sampler's critical path is compare -> branch -> add per sample, and the producer warp shares its
            // scheduler with consumers (ncu r02_e: consumers spent a quarter of the kernel waiting for it).  Between two
            // events (segment change, end of the trajectory: ~9 per 1000 samples) the loop only adds, so: advance()
            // settles the exact state machine at the current sample, then a run of n samples is emitted with no
            // compares, n chosen so that NO event can fall inside it — a bound in real arithmetic with a margin that
            // covers the rounding of up to 2^20 running-sum steps.  Same sequence of rounded adds, same values.
            while (cnt < 32 && advance()) {
                int n = 1;
                if (dt > 0.0) {
                    const double room_t = dsub(dsub(seg_T, tin), dmul(1e-9, dadd(fabs(seg_T), 1.0)));
                    const double room_a = dsub(dsub(t_end, acc), dmul(1e-9, dadd(fabs(t_end), 1.0)));
                    const double steps = __ddiv_rn(room_t < room_a ? room_t : room_a, dt);
                    if (steps > 3.0) n = steps > 64.0 ? 32 : (int)steps - 2; // states 1 .. n-1 emit for sure
                }
                n = n < 32 - cnt ? n : 32 - cnt;
                for (int u = 0; u < n; ++u) {
                    st[cnt + u] = tin; // :132
                    sg[cnt + u] = (unsigned short)seg;
                    tin = dadd(tin, dt); // :138-139
                    acc = dadd(acc, dt);
                }
                cnt += n;
            }
            if (have) ring_cnt(slot)[lane] = (uint32_t)cnt;
            emitted_total += (uint32_t)cnt;
            if (__ballot_sync(0xffffffffu, cnt > 0) == 0u) break; // every trajectory of the block has ended
            __syncwarp();
            if (lane == 0) {
                asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(smem_addr(&s_produced)), "r"(round + 1u) : "memory");
                if (sleep_wait) mbar_arrive(&s_pub[round % kPolyBars]); // wakes the consumers that sleep on this round
            }
        }
        if (lane == 0) {
            asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(smem_addr(&s_total)), "r"(round) : "memory");
            if (sleep_wait)
                for (int b = 0; b < kPolyBars; ++b) mbar_arrive(&s_pub[b]); // and those that sleep on rounds that will not come
        }
        if (have) {
            if (advance()) atomicMin(first_row + my, (long long)max_samples); // truncated: rows beyond max_samples unchecked
            if (n_samples_out) n_samples_out[my] = emitted_total;
        }
        return;
    }

    // ---- consumers: a work item is one round of TWO trajectories (j0, j0 + 1): their Horner chains are independent,
    // so a lane has six dependent DMUL/DADD chains in flight instead of three (ncu r02_e: the FP64 pipe was 27 % busy
    // with the warps waiting on its latency), and the per-item bookkeeping is halved.
    const uint32_t Hb = (Gb + 1u) >> 1;
    const uint32_t hb_magic = Hb > 1u ? (uint32_t)(0x100000000ull / Hb) + ((0x100000000ull % Hb) ? 1u : 0u) : 0u;
    for (;;) {
        uint32_t i = 0;
        if (lane == 0) i = atomicAdd(&s_next, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        const uint32_t round = Hb == 1u ? i : __umulhi(i, hb_magic), j0 = (i - round * Hb) * 2u; // i / Hb (exact for i < 2^27)
        bool gone = false;
        // Wait for the round, or learn that it will never come.  s_produced / s_total decide; the barrier of the round
        // is only where the warp SLEEPS meanwhile (ncu r02_z: with a 200 ns nanosleep poll the waiting consumers were
        // 32 % of all issued instructions, on the schedulers the producer they wait for needs).  A try_wait that
        // returns early (time limit, or a barrier already two phases on) just polls again.
        for (;;) {
            uint32_t p, t;
            asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(p) : "r"(smem_addr(&s_produced)) : "memory");
            if (p > round) break;
            asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(t) : "r"(smem_addr(&s_total)) : "memory");
            if (t <= round) {
                gone = true;
                break;
            }
            if (sleep_wait) (void)mbar_try_wait(&s_pub[round % kPolyBars], (round / kPolyBars) & 1u);
            else __nanosleep(200);
        }
        if (gone) break;
        const uint32_t slot = round % (uint32_t)ring;
        const bool two = j0 + 1u < Gb;
        const int cnt0 = (int)ring_cnt(slot)[j0], cnt1 = two ? (int)ring_cnt(slot)[j0 + 1u] : 0;
        P3 p0 = {0.0, 0.0, 0.0}, p1 = {0.0, 0.0, 0.0};
        {
            // both samples' coefficients are requested before either chain starts
            const size_t tj0 = traj0 + j0, tj1 = tj0 + (two ? 1u : 0u);
            const uint32_t at0 = j0 * kPolyPitch + lane, at1 = (j0 + (two ? 1u : 0u)) * kPolyPitch + lane;
            const bool a0 = lane < cnt0, a1 = lane < cnt1;
            const double t0 = a0 ? ring_t(slot)[at0] : 0.0, t1 = a1 ? ring_t(slot)[at1] : 0.0;
            const int sg0 = a0 ? ring_seg(slot)[at0] : 0, sg1 = a1 ? ring_seg(slot)[at1] : 0;
            const double *c0 = coeffs + ((tj0 * n_segments + sg0) * 3) * (size_t)n_coeffs; // segment.cpp:51-58
            const double *c1 = coeffs + ((tj1 * n_segments + sg1) * 3) * (size_t)n_coeffs;
            if (N10) {
                if (a0 | a1) {
                    p0 = {horner10(c0, t0), horner10(c0 + 10, t0), horner10(c0 + 20, t0)};
                    p1 = {horner10(c1, t1), horner10(c1 + 10, t1), horner10(c1 + 20, t1)};
                }
            } else {
                if (a0) p0 = {horner(c0, n_coeffs, t0), horner(c0 + n_coeffs, n_coeffs, t0), horner(c0 + 2 * n_coeffs, n_coeffs, t0)};
                if (a1) p1 = {horner(c1, n_coeffs, t1), horner(c1 + n_coeffs, n_coeffs, t1), horner(c1 + 2 * n_coeffs, n_coeffs, t1)};
            }
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int cj = k ? cnt1 : cnt0;
            if (cj == 0) continue;
            const size_t tj = traj0 + j0 + k;
            bool valid = false;
            if (lane < cj) valid = !world.invalid(wv, k ? p1 : p0, kModeMinDist, false, min_distance);
            const uint32_t word = __ballot_sync(0xffffffffu, valid);
            if (lane == 0) {
                if (sample_mask) sample_mask[tj * words_per_traj + round] = word;
                const uint32_t active = cj >= 32 ? 0xffffffffu : ((1u << cj) - 1u);
                const uint32_t bad = active & ~word;
                if (bad) atomicMin(first_row + tj, (long long)((size_t)round * 32 + (size_t)(__ffs(bad) - 1)));
            }
        }
        __syncwarp();
        if (lane == 0) {
            __threadfence_block();
            atomicAdd(&s_done[slot], two ? 2u : 1u);
        }
    }
}

__global__ void k_fill_u32(uint32_t *p, size_t n, uint32_t v)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// ---- generators (bit-identical to the oracle's host generators) ------------------------------

__device__ __forceinline__ uint64_t mix64(uint64_t seed, uint64_t counter)
{
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (counter + 1);
    z ^= z >> 30;
    z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27;
    z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

__device__ __forceinline__ double u01(uint64_t seed, uint64_t index, unsigned lane)
{
    return dmul((double)(mix64(seed, index * 8 + lane) >> 11), 0x1.0p-53);
}

struct Box3 {
    double lo[3], hi[3];
};

__global__ void k_gen_states(uint64_t seed, uint64_t first, size_t n, Box3 box, double *__restrict__ xyz)
{
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; // one thread per coordinate
    if (e >= 3 * n) return;
    const size_t k = e / 3;
    const unsigned d = (unsigned)(e - 3 * k);
    xyz[e] = dadd(box.lo[d], dmul(u01(seed, first + k, d), dsub(box.hi[d], box.lo[d])));
}

__global__ void k_gen_segments(uint64_t seed, uint64_t first, size_t n, Box3 box, double max_len,
                               double *__restrict__ pa, double *__restrict__ pb)
{
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint64_t idx = first + k;
    double a[3], v[3];
    for (unsigned d = 0; d < 3; ++d) {
        a[d] = dadd(box.lo[d], dmul(u01(seed, idx, d), dsub(box.hi[d], box.lo[d])));
        v[d] = dsub(dmul(2.0, u01(seed, idx, 3 + d)), 1.0);
    }
    const double len2 = dadd(dadd(dmul(v[0], v[0]), dmul(v[1], v[1])), dmul(v[2], v[2]));
    const double norm = __dsqrt_rn(len2);
    const double L = dmul(max_len, dsub(1.0, u01(seed, idx, 6)));
    for (unsigned d = 0; d < 3; ++d) {
        const double u = norm > 0.0 ? __ddiv_rn(v[d], norm) : 0.0;
        double e = dadd(a[d], dmul(L, u));
        if (e < box.lo[d]) e = box.lo[d];
        if (e > box.hi[d]) e = box.hi[d];
        pa[3 * k + d] = a[d];
        pb[3 * k + d] = e;
    }
}

// out[0] += popcount, out[1] ^= XOR over words of mix64(0x5EED, (global word index << 32) | word).
__global__ void k_mask_summary(const uint32_t *__restrict__ mask, size_t n_words, uint64_t first_word,
                               unsigned long long *out)
{
    unsigned long long pop = 0, x = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t w = mask[i];
        pop += (unsigned)__popc(w);
        x ^= mix64(0x5EEDull, ((first_word + i) << 32) | w);
    }
    for (int o = 16; o; o >>= 1) {
        pop += __shfl_xor_sync(0xffffffffu, pop, o);
        x ^= __shfl_xor_sync(0xffffffffu, x, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(out, pop);
        atomicXor(out + 1, x);
    }
}

// ---- single-unit mailbox: StateValidator::isValid / MotionValidator::checkMotion one call at a time --------------
//
// OMPL's planners ask one state or one edge at a time (src/PathPlanner.cpp:106-129).  A kernel launch plus a stream
// synchronisation per question costs ~23 us; the reference answers in ~0.13 us.  k_mailbox is ONE resident warp per
// host thread and world epoch that polls a request line in mapped pinned host memory (MailboxSlot, kernels.h): the host
// writes the unit and bumps a sequence number, the warp evaluates it — lane k takes OBBs k, k + 32, ... with the
// same per-candidate functions as the batch kernels (grid worlds: the cell list) — and stores (sequence number,
// verdict) back as one 8-byte store; the host spins on that line.  No launch and no synchronisation per question.
// The warp leaves when no request has come for idle_ns (so that device-wide synchronisations elsewhere in the
// process are never held up for longer than that) or on a QUIT request; the host relaunches it on demand.
template <int BP>
__global__ void __launch_bounds__(32) k_mailbox(WorldView wv, MailboxSlot *slot, uint32_t last_seq,
                                                unsigned long long idle_ns)
{
    const int lane = threadIdx.x;
    PointWorld<BP> pw;
    RayWorld<BP> rw;
    pw.init(wv, false);
    rw.init(wv, false);
    const unsigned long long *req = const_cast<const unsigned long long *>(reinterpret_cast<const volatile unsigned long long *>(slot->req));
    unsigned long long t_last;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_last));
    uint32_t polls = 0;
    bool leaving = false;
    for (;;) {
        // one 64-byte read of the request line: lane l < 8 takes 8 bytes of it
        unsigned long long v = 0ull;
        if (lane < 8) asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(req + lane) : "memory");
        const unsigned long long head = __shfl_sync(0xffffffffu, v, 0), tail = __shfl_sync(0xffffffffu, v, 7);
        const uint32_t seq_a = (uint32_t)head, flags = (uint32_t)(head >> 32), seq_b = (uint32_t)tail;
        if (seq_a == last_seq || seq_a != seq_b) { // nothing new (or the two halves of the line are of different age)
            if (leaving) break;
            if ((++polls & 15u) == 0u) {
                unsigned long long t;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                if (t - t_last > idle_ns || polls > (1u << 24)) {
                    // announce, then look once more: a request posted before the announcement is still served
                    if (lane == 0) {
                        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(const_cast<uint32_t *>(&slot->state)), "r"((uint32_t)kMailboxLeaving) : "memory");
                        __threadfence_system();
                    }
                    __syncwarp();
                    leaving = true;
                }
            }
            continue;
        }
        last_seq = seq_a;
        if (leaving) {
            leaving = false;
            if (lane == 0) asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(const_cast<uint32_t *>(&slot->state)), "r"((uint32_t)kMailboxRunning) : "memory");
        }
        if (flags & kMailboxQuit) break;
        const P3 a = {__longlong_as_double((long long)__shfl_sync(0xffffffffu, v, 1)),
                      __longlong_as_double((long long)__shfl_sync(0xffffffffu, v, 2)),
                      __longlong_as_double((long long)__shfl_sync(0xffffffffu, v, 3))};
        const P3 b = {__longlong_as_double((long long)__shfl_sync(0xffffffffu, v, 4)),
                      __longlong_as_double((long long)__shfl_sync(0xffffffffu, v, 5)),
                      __longlong_as_double((long long)__shfl_sync(0xffffffffu, v, 6))};
        const bool cpg = (flags & kMailboxCanPassGate) != 0;
        bool hit = false;
        if (flags & kMailboxSegment) {
            P3 mn, mx;
            ray_box(a, b, mn, mx);
            if (BP == EPP_BP_GRID) {
                if (lane == 0) hit = rw.invalid(wv, a, b, cpg);
            } else {
                for (int k = lane; k < wv.M && !hit; k += 32) hit = ray_candidate_invalid(wv.aabb, wv.obb, k, a, b, mn, mx, cpg);
            }
        } else {
            if (BP == EPP_BP_GRID) {
                if (lane == 0) hit = pw.invalid(wv, a, kModeBool, cpg, 0.0);
            } else {
                for (int k = lane; k < wv.M && !hit; k += 32) hit = point_candidate_invalid(wv.aabb, wv.obb, k, a, kModeBool, cpg, 0.0);
            }
        }
        const bool invalid = __any_sync(0xffffffffu, hit);
        if (lane == 0)
            asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(const_cast<uint32_t *>(slot->ack)), "r"(seq_a), "r"(invalid ? 0u : 1u) : "memory");
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_last));
        polls = 0;
    }
    if (lane == 0) {
        __threadfence_system();
        asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(const_cast<uint32_t *>(slot->exited)), "r"(last_seq), "r"(1u) : "memory");
        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(const_cast<uint32_t *>(&slot->state)), "r"((uint32_t)kMailboxExited) : "memory");
    }
}

// ---- broad-phase statistics (bench.py: measured M_visited and k of SURVEY.md 8(d)'s ops-per-check formula) -----------

// out[0] += broad-phase candidates visited, out[1] += candidates whose stored AABB passes the reference's exact test
// (strict containment for points, closed box/box for segments: the R-tree hits of src/World.cpp:83,143).
template <int BP>
__global__ void __launch_bounds__(kBlock) k_count_candidates(WorldView wv, const double *__restrict__ pa,
                                                             const double *__restrict__ pb, size_t n, int aabb_in_smem,
                                                             unsigned long long *out)
{
    PointWorld<BP> pw;
    RayWorld<BP> rw;
    if (pb) rw.init(wv, aabb_in_smem != 0);
    else pw.init(wv, aabb_in_smem != 0);
    unsigned long long visited = 0, hits = 0;
    for (size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (size_t)gridDim.x * kBlock) {
        const P3 a = {__ldg(pa + 3 * i), __ldg(pa + 3 * i + 1), __ldg(pa + 3 * i + 2)};
        if (pb) {
            const P3 b = {__ldg(pb + 3 * i), __ldg(pb + 3 * i + 1), __ldg(pb + 3 * i + 2)};
            P3 mn, mx;
            ray_box(a, b, mn, mx);
            rw.any(wv, mn, mx, [&](int k) -> bool {
                ++visited;
                if (boxes_touch(wv.aabb + k, mn, mx)) ++hits;
                return false;
            });
        } else {
            pw.any(wv, a, [&](int k) -> bool {
                ++visited;
                if (within_strict(wv.aabb + k, a)) ++hits;
                return false;
            });
        }
    }
    for (int o = 16; o; o >>= 1) {
        visited += __shfl_xor_sync(0xffffffffu, visited, o);
        hits += __shfl_xor_sync(0xffffffffu, hits, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(out, visited);
        atomicAdd(out + 1, hits);
    }
}

// Dynamic shared memory a kernel variant needs for this world.
size_t point_smem_bytes(const WorldView &wv, int bp, bool &aabb_in_smem)
{
    aabb_in_smem = false;
    if (bp == EPP_BP_BITS) return (size_t)(wv.cmask_words ? wv.cmask_off + wv.cmask_words : 3 * wv.nb * wv.Ws) * 4;
    if (bp == EPP_BP_BRUTE && wv.M <= 2048) {
        aabb_in_smem = true;
        return (size_t)wv.M * sizeof(AabbRec);
    }
    return 0;
}

size_t ray_smem_bytes(const WorldView &wv, int bp, bool &aabb_in_smem)
{
    aabb_in_smem = false;
    if (bp == EPP_BP_BITS) return (wv.range_d > 0 ? (size_t)2 * wv.nb * wv.range_d * wv.Ws : (size_t)6 * wv.nb * wv.Ws) * 4;
    if (bp == EPP_BP_BRUTE && wv.M <= 2048) {
        aabb_in_smem = true;
        return (size_t)wv.M * sizeof(AabbRec);
    }
    return 0;
}

// Opt in to more than 48 KB of (static + dynamic) shared memory per block when needed.
template <typename K>
cudaError_t ensure_smem(K kernel, size_t bytes, size_t static_bytes = 0)
{
    if (bytes + static_bytes <= 40 * 1024) return cudaSuccess;
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

// The coarse-cell word masks pay for kernels that copy the table once per SM (the warp-pipelined point kernels); direct
// forms copy the table per block (the 22 KB of masks made K3b 20 % slower) and the ray queries loop over coarse cells
// (K2 10 % slower): those take the view without them.
WorldView without_cmask(const WorldView &wv)
{
    WorldView v = wv;
    v.cmask_words = 0;
    return v;
}

unsigned persistent_grid(const LaunchCfg &cfg, size_t tiles, int blocks_per_sm)
{
    size_t cap = (size_t)cfg.sm_count * (size_t)blocks_per_sm;
    size_t g = tiles < cap ? tiles : cap;
    return (unsigned)(g ? g : 1);
}

} // namespace

std::atomic<uint64_t> g_launches{0};

#ifdef EPP_TRACE
extern "C" int epp_debug_trace_base(unsigned base, void *stream)
{
    k_trace_base<<<1, 1, 0, (cudaStream_t)stream>>>(base);
    return (int)cudaGetLastError();
}

extern "C" int epp_debug_trace_copy(unsigned long long *host, size_t n, int clear)
{
    cudaDeviceSynchronize();
    if (n > (1u << 20)) n = 1u << 20;
    cudaError_t e = cudaMemcpyFromSymbol(host, g_trace, n * 8);
    if (e == cudaSuccess && clear) {
        void *p;
        cudaGetSymbolAddress(&p, g_trace);
        cudaMemset(p, 0, sizeof(unsigned long long) << 20);
    }
    return (int)e;
}
#endif

#define EPP_DISPATCH_BP(bp, ...)                                                 \
    switch (bp) {                                                                \
    case EPP_BP_BRUTE: { constexpr int BP = EPP_BP_BRUTE; __VA_ARGS__; } break;  \
    case EPP_BP_BITS: { constexpr int BP = EPP_BP_BITS; __VA_ARGS__; } break;    \
    default: { constexpr int BP = EPP_BP_GRID; __VA_ARGS__; } break;             \
    }

cudaError_t launch_gather_push_impl(const LaunchCfg &cfg, const GatherView &gv, cudaStream_t stream);
bool gather_is_bulk(const LaunchCfg &cfg, const GatherView &gv);

// TR: the units are the rows of ta's matrices (n = trajectories * rows is only used for the size checks).
template <int BP, int S, bool RS, bool GA, bool TR = false>
cudaError_t launch_states_warp(const LaunchCfg &cfg, const WorldView &wv, const double *d_xyz, size_t n, int mode,
                               int can_pass_gate, double min_distance, bool in_smem, size_t table_bytes, int stages,
                               uint32_t *d_mask, cudaStream_t stream, bool &launched, const GatherView &gv,
                               TrajArgs ta = TrajArgs(), size_t n_traj = 0)
{
    launched = false;
    if (TR) {
        ta.wpt = (uint32_t)((ta.rows + 31) / 32);
        ta.cpt = (uint32_t)((ta.rows + S * 32 - 1) / (S * 32));
    }
    const size_t words = TR ? n_traj * ta.cpt * S : (n + 31) / 32;
    if (words >= 0xfff00000ull) return cudaSuccess;
    const size_t rec_bytes = RS ? ((size_t)wv.M * kRecDoubles * 8 + 127) / 128 * 128 : 0;
    const size_t warp_bytes = warp_area_bytes(BP, S), slot_bytes = (size_t)S * kWordBytes;
    // resident blocks per SM (tuning, default 1) and the shared-memory budget of each
    int bps = cfg.states_warp_blocks_per_sm > 0 ? cfg.states_warp_blocks_per_sm : 1;
    int threads = cfg.states_warp_threads;
    // Batches of a multi-batch call overlap (late dependency wait): that only works if a block of the next launch
    // fits next to a block of this one, so they run as blocks of 512 threads with half the ring each, ONE per SM in
    // the grid and room for the next launch's block beside it: two launches share every SM, one in its ramp or tail
    // while the other streams (1 M states on the track: 6.8 us per batch; 7.4 with both of a launch's blocks on the
    // SM; 8.2 with one 1024-thread block; single launches are ~8 % slower in this shape, so only multi-batch calls use it).
    if (cfg.states_multi > 0 && !GA && !TR && bps == 1 && threads == 1024) {
        bps = cfg.states_multi;
        threads = 1024 / bps;
    }
    bps = bps > 8 ? 8 : bps;
    int cons = (threads > 0 ? threads / 32 : kMaxWarps) - 1;
    cons = cons < 1 ? 1 : (cons > kMaxWarps - 1 ? kMaxWarps - 1 : cons);
    // the kernel is compiled for one 1024-thread block per SM (up to 64 registers per thread): several
    // resident blocks share that budget
    if ((cons + 1) * 32 * bps > 1024) cons = 1024 / (32 * bps) - 1;
    cons = cons < 1 ? 1 : cons;
    const size_t chunks_total = (words + S - 1) / S;
    // multi-batch launches may put FEWER blocks into the grid than fit (states_multi_grid): the blocks of the next
    // launch then run next to this launch's blocks for its whole duration, not only in its tail
    const int gps = (cfg.states_multi > 0 && cfg.states_multi_grid > 0 && cfg.states_multi_grid < bps) ? cfg.states_multi_grid : bps;
    const size_t blocks_cap = (size_t)cfg.sm_count * gps;
    // ownership is handed out in words for state batches (12-word units with the fused exchange: whole chunks and whole
    // 3-word exchange units) and in whole chunks for trajectory matrices
    const size_t gran = GA ? 12 : 1;
    const size_t own_units = TR ? chunks_total : (words + gran - 1) / gran;
    size_t grid = chunks_total < blocks_cap ? chunks_total : blocks_cap;
    if (grid > own_units) grid = own_units;
    const size_t words_per_block = ((own_units + grid - 1) / grid) * (TR ? S : gran);
    const size_t chunks_per_block = (words_per_block + S - 1) / S;
    if ((size_t)cons > chunks_per_block) cons = (int)chunks_per_block;
    const size_t budget = (size_t)(228 * 1024) / bps - 1024 - 4096; // 1 KB reserved per block + static + slack
    // GA: a block parks its verdict words and sends them in groups of whole exchange units; batches whose blocks own
    // more words than the parking area holds (> 4.8 M states per rank) take the plain kernel + the push kernel
    uint32_t ga_group = 0;
    if (GA) {
        ga_group = (uint32_t)(cfg.gather_group_chunks / 3 * 3);
        ga_group = ga_group < 3u ? 3u : ga_group;
        if (words_per_block > kGaStageWords || gather_is_bulk(cfg, gv)) { // every rank must take the same form: decided by the slot size
            if constexpr (GA) {
                const GatherView none = {};
                cudaError_t e2 = launch_states_warp<BP, S, RS, false, false>(cfg, wv, d_xyz, n, mode, can_pass_gate, min_distance,
                                                                            in_smem, table_bytes, stages, d_mask, stream, launched, none);
                if (e2 != cudaSuccess || !launched) return e2;
                return launch_gather_push_impl(cfg, gv, stream);
            }
        }
    }
    const size_t ga_bytes = ga_group ? (size_t)kGaStageWords * 4 + (kGaStageWords / S) * 4 : 0;
    const size_t fixed = table_bytes + rec_bytes + cons * warp_bytes + ga_bytes;
    if (fixed + 2 * slot_bytes > budget) return cudaSuccess;
    size_t slots = (budget - fixed) / slot_bytes;
    slots = slots > (size_t)kMaxSlots ? (size_t)kMaxSlots : slots;
    if (slots > chunks_per_block) slots = chunks_per_block < 2 ? 2 : chunks_per_block;
    if (cfg.states_warp_log2_slots > 0 && ((size_t)1 << cfg.states_warp_log2_slots) < slots) slots = (size_t)1 << cfg.states_warp_log2_slots;
    // A consumer that takes chunk i waits on phase (i / slots) of the slot's `full` barrier; that is only
    // well defined if chunk i - slots has been consumed, which in-order hand-out guarantees iff slots >= consumers.
    if ((size_t)cons > slots) cons = (int)slots;
    // records are folded by one consumer thread each: a batch too small for M consumer threads per block takes
    // the records from global memory instead (the caller retries with RS = false)
    if (RS && cons * 32 < wv.M) return cudaSuccess;
    const size_t smem = fixed + slot_bytes * slots;
    cudaError_t err = cudaFuncSetAttribute(k_states_warp<BP, S, RS, GA, TR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)grid);
    lc.blockDim = dim3((unsigned)(cons + 1) * 32);
    lc.dynamicSmemBytes = smem;
    lc.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = attr;
    lc.numAttrs = cfg.states_pdl ? 1 : 0;
    err = cudaLaunchKernelEx(&lc, k_states_warp<BP, S, RS, GA, TR>, wv, d_xyz, n,
                             mode | (cfg.states_debug << 8) | ((cfg.states_late_wait && cfg.states_pdl && !GA && !TR) ? (1 << 16) : 0),
                             can_pass_gate, min_distance, in_smem ? 1 : 0, (uint32_t)slots, (uint32_t)table_bytes,
                             (uint32_t)rec_bytes, (uint32_t)(own_units / grid), (uint32_t)(own_units % grid), d_mask, gv,
                             ga_group, ta, (uint32_t)gran);
    if (err != cudaSuccess) return err;
    launched = true;
    return cudaGetLastError();
}

cudaError_t launch_gather_push(const LaunchCfg &cfg, const GatherView &gv, cudaStream_t stream)
{
    return launch_gather_push_impl(cfg, gv, stream);
}

bool gather_is_bulk(const LaunchCfg &cfg, const GatherView &gv)
{
    return gv.world > 1 && cfg.gather_bulk_min_words > 0 && gv.slot_words >= (uint64_t)cfg.gather_bulk_min_words;
}

cudaError_t launch_gather_deferred(const LaunchCfg &cfg, const GatherView &gv, int flush, cudaStream_t stream)
{
    size_t blocks = (gv.slot_words / 4 + 255) / 256;
    const size_t cap = (size_t)cfg.sm_count * 2;
    blocks = blocks < 1 ? 1 : (blocks > cap ? cap : blocks);
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)blocks);
    lc.blockDim = dim3(256);
    lc.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = attr;
    lc.numAttrs = 1;
    cudaError_t err = cudaLaunchKernelEx(&lc, k_gather_deferred, gv, flush);
    g_launches++;
    return err != cudaSuccess ? err : cudaGetLastError();
}

cudaError_t launch_gather_push_impl(const LaunchCfg &cfg, const GatherView &gv, cudaStream_t stream)
{
    if (gather_is_bulk(cfg, gv)) {
        size_t nb = (gv.slot_words / 4 + 255) / 256;
        const size_t capb = (size_t)cfg.sm_count * 8;
        nb = nb < 1 ? 1 : (nb > capb ? capb : nb);
        k_gather_bulk<<<(unsigned)nb, 256, 0, stream>>>(gv);
        g_launches++;
        return cudaGetLastError();
    }
    if (gv.deferred) return launch_gather_deferred(cfg, gv, 0, stream);
    size_t blocks = (gv.slot_words / 4 + 255) / 256;
    const size_t cap = (size_t)cfg.sm_count * 4;
    blocks = blocks < 1 ? 1 : (blocks > cap ? cap : blocks);
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)blocks);
    lc.blockDim = dim3(256);
    lc.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = attr;
    lc.numAttrs = 1;
    cudaError_t err = cudaLaunchKernelEx(&lc, k_gather_push, gv);
    g_launches++;
    return err != cudaSuccess ? err : cudaGetLastError();
}

cudaError_t launch_states(const LaunchCfg &cfg, const WorldView &wv, int bp, const double *d_xyz, size_t n, int mode,
                          int can_pass_gate, double min_distance, uint32_t *d_mask, cudaStream_t stream,
                          const GatherView *gv)
{
    // A gathering call is a collective: it announces its (empty) slot and unpacks the peers' even without units.
    if (n == 0) return gv ? launch_gather_push(cfg, *gv, stream) : cudaSuccess;
    const GatherView no_gather = {};
    bool in_smem;
    const size_t smem = point_smem_bytes(wv, bp, in_smem);
    cudaError_t err = cudaSuccess;
    // Warp-pipelined form: any alignment, any n; worth its prologue from a few thousand states up.
    if ((cfg.states_variant == 3 || (cfg.states_variant == 0 && n >= (size_t)cfg.states_warp_min_n))) {
        size_t table_bytes = ((bp == EPP_BP_GRID ? (size_t)wv.occ_bytes : smem) + 127) / 128 * 128;
        const int stages = 0; // ring depth comes from the shared-memory budget
        const int S = cfg.states_per_thread == 8 ? 8 : 4;
        const int rec_max = cfg.states_rec_smem_max > 0 ? (cfg.states_rec_smem_max < 992 ? cfg.states_rec_smem_max : 992) : kRecSmemMax;
        bool rs = wv.M <= rec_max && wv.zrot && cfg.states_rec_smem != 0;
        const int wbp = (bp == EPP_BP_BITS && wv.W == 1 && cfg.states_bits1 != 0) ? kBpBits1 : bp;
        bool launched = false;
#define EPP_LAUNCH_WARP(S_, RS_, GA_)                                                                                 \
    err = launch_states_warp<BP, S_, RS_, GA_>(cfg, wv, d_xyz, n, mode, can_pass_gate, min_distance, in_smem,          \
                                               table_bytes, stages, d_mask, stream, launched, gv ? *gv : no_gather)
        // the fused exchange exists for the default chunk size; S = 8 (a tuning knob) takes the push kernel
#define EPP_LAUNCH_WARP_S()                                  \
    if (S == 8 && rs) EPP_LAUNCH_WARP(8, true, false);       \
    else if (S == 8) EPP_LAUNCH_WARP(8, false, false);       \
    else if (rs && gv) EPP_LAUNCH_WARP(4, true, true);       \
    else if (rs) EPP_LAUNCH_WARP(4, true, false);            \
    else if (gv) EPP_LAUNCH_WARP(4, false, true);            \
    else EPP_LAUNCH_WARP(4, false, false)
        for (int attempt = 0; attempt < 2 && !launched && err == cudaSuccess; ++attempt) {
            if (attempt == 1) {
                if (!rs) break;
                rs = false; // too few consumer threads for the records: second try with global records
            }
            if (wbp == kBpBits1) {
                constexpr int BP = kBpBits1;
                EPP_LAUNCH_WARP_S();
            } else {
                EPP_DISPATCH_BP(bp, { EPP_LAUNCH_WARP_S(); });
            }
        }
#undef EPP_LAUNCH_WARP_S
#undef EPP_LAUNCH_WARP
        if (err != cudaSuccess) return err;
        if (launched) {
            g_launches++;
            if (gv && S == 8) return launch_gather_push(cfg, *gv, stream);
            return cudaSuccess;
        }
    }
    const size_t tiles = (n + kBlock - 1) / kBlock;
    const unsigned grid = persistent_grid(cfg, tiles, cfg.states_blocks_per_sm);
    const WorldView wd = without_cmask(wv);
    const size_t smem_d = point_smem_bytes(wd, bp, in_smem);
    EPP_DISPATCH_BP(bp, {
        err = ensure_smem(k_states<BP>, smem_d);
        if (err == cudaSuccess) {
            k_states<BP><<<grid, kBlock, smem_d, stream>>>(wd, d_xyz, n, mode, can_pass_gate, min_distance, in_smem ? 1 : 0, d_mask);
            err = cudaGetLastError();
        }
    });
    g_launches++;
    if (err == cudaSuccess && gv) err = launch_gather_push(cfg, *gv, stream);
    return err;
}

template <int BP, bool RS>
cudaError_t launch_segments_warp(const LaunchCfg &cfg, const WorldView &wv, const double *d_a, const double *d_b, size_t n,
                                 int can_pass_gate, size_t table_bytes, bool in_smem, uint32_t *d_mask,
                                 cudaStream_t stream, bool &launched)
{
    launched = false;
    const size_t words = (n + 31) / 32;
    if (words >= 0xfff00000ull) return cudaSuccess;
    const size_t rec_bytes = RS ? ((size_t)wv.M * kRayRec * 8 + 127) / 128 * 128 : 0;
    const size_t chunks_total = (words + kSegS - 1) / kSegS;
    const size_t grid = chunks_total < (size_t)cfg.sm_count ? chunks_total : (size_t)cfg.sm_count;
    const size_t chunks_per_block = (chunks_total + grid - 1) / grid;
    // Batches of a multi-batch call overlap (late dependency wait).  Grid worlds: a block of the next launch fits beside
    // a block of this one — blocks of 1024 / segments_multi threads with that share of the shared memory, one per SM in
    // the grid, as k_states_warp does (4 M edges, 10 000 OBBs: 107 -> 91 us per batch).  Bits / brute worlds keep the
    // 1024-thread block: their table and folded records take 65 KB of every block, so half an SM leaves 13 ring slots
    // and 13 consumer warps (1 M segments, 256 OBBs: 42 us per batch against 22.7); with the late wait alone the next
    // launch's block starts on an SM the moment this launch's block leaves it: 21.2 us (profiles/r02_segmulti_ab.log).
    const int bps = (BP == EPP_BP_GRID && cfg.segments_multi > 1) ? (cfg.segments_multi > 4 ? 4 : cfg.segments_multi) : 1;
    int cons = (bps > 1 ? 1024 / bps / 32 : (cfg.segments_warp_threads > 0 ? cfg.segments_warp_threads / 32 : kMaxWarps)) - 1;
    cons = cons < 1 ? 1 : (cons > kMaxWarps - 1 ? kMaxWarps - 1 : cons);
    if ((size_t)cons > chunks_per_block) cons = (int)chunks_per_block;
    const size_t budget = (size_t)(228 * 1024) / bps - 1024 - 4096;
    const size_t fixed = table_bytes + rec_bytes + (size_t)cons * kSegWarpBytes;
    if (fixed + 2 * (size_t)kSegSlotBytes > budget) return cudaSuccess;
    size_t slots = (budget - fixed) / kSegSlotBytes;
    if (slots > chunks_per_block) slots = chunks_per_block < 2 ? 2 : chunks_per_block;
    if (cfg.segments_warp_slots > 0 && (size_t)cfg.segments_warp_slots < slots) slots = cfg.segments_warp_slots;
    slots = slots >= (size_t)kSegMaxSlots ? (size_t)kSegMaxSlots : slots;
    if (slots < 2) slots = 2;
    if ((size_t)cons > slots) cons = (int)slots; // in-order hand-out needs slots >= consumers (see k_states_warp)
    const size_t smem = fixed + slots * kSegSlotBytes;
    cudaError_t err = cudaFuncSetAttribute(k_segments_warp<BP, RS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)grid);
    lc.blockDim = dim3((unsigned)(cons + 1) * 32);
    lc.dynamicSmemBytes = smem;
    lc.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = attr;
    lc.numAttrs = cfg.states_pdl ? 1 : 0;
    err = cudaLaunchKernelEx(&lc, k_segments_warp<BP, RS>, wv, d_a, d_b, n, can_pass_gate, in_smem ? 1 : 0, (uint32_t)slots,
                             (uint32_t)table_bytes, (uint32_t)rec_bytes, (uint32_t)(words / grid), (uint32_t)(words % grid),
                             (cfg.segments_late_wait && cfg.states_pdl) ? 1 : 0, d_mask);
    if (err != cudaSuccess) return err;
    launched = true;
    return cudaGetLastError();
}

cudaError_t launch_segments(const LaunchCfg &cfg, const WorldView &wv_in, int bp, const double *d_a, const double *d_b,
                            size_t n, int can_pass_gate, uint32_t *d_mask, cudaStream_t stream, const GatherView *gv)
{
    const WorldView wv = without_cmask(wv_in);
    if (n == 0) return gv ? launch_gather_push(cfg, *gv, stream) : cudaSuccess;
    bool in_smem;
    const size_t smem = ray_smem_bytes(wv, bp, in_smem);
    cudaError_t err = cudaSuccess;
    const size_t tiles = (n + kBlock - 1) / kBlock;
    if (cfg.segments_variant == 2 || (cfg.segments_variant == 0 && n >= (size_t)cfg.segments_warp_min_n)) {
        // warp-pipelined form: end points through the cp.async.bulk ring
        const bool rs = wv.zrot && wv.M <= kRayRecMax && bp != EPP_BP_GRID;
        bool in_smem_w = in_smem && !rs;
        size_t table_bytes = bp == EPP_BP_GRID ? (size_t)wv.occ_bytes : ((rs && bp == EPP_BP_BRUTE) ? 0 : smem);
        table_bytes = (table_bytes + 127) / 128 * 128;
        bool launched = false;
        EPP_DISPATCH_BP(bp, {
            if (rs) err = launch_segments_warp<BP, true>(cfg, wv, d_a, d_b, n, can_pass_gate, table_bytes, in_smem_w, d_mask, stream, launched);
            else err = launch_segments_warp<BP, false>(cfg, wv, d_a, d_b, n, can_pass_gate, table_bytes, in_smem_w, d_mask, stream, launched);
        });
        if (err != cudaSuccess) return err;
        if (launched) {
            g_launches++;
            if (gv) err = launch_gather_push(cfg, *gv, stream);
            return err;
        }
    }
    if (cfg.segments_variant != 1 && wv.zrot && wv.M <= kRayRecMax) {
        // records in shared memory (the broad-phase table of the brute form is the records themselves)
        const size_t table_bytes = bp == EPP_BP_BRUTE ? 0 : (smem + 127) / 128 * 128;
        const size_t total = table_bytes + (size_t)wv.M * kRayRec * 8;
        // persistent blocks: as many per SM as the shared memory holds at once (a second wave would start late)
        int bps = (int)((size_t)(227 * 1024) / (total + 1024));
        bps = bps < 1 ? 1 : (bps > cfg.segments_blocks_per_sm ? cfg.segments_blocks_per_sm : bps);
        // threads per block (tuning knob segments_block): the records are per block, so bigger blocks put more warps
        // behind the same shared memory (3 x 256 threads fit an SM with 71 KB each, 2 x 512 or 1 x 1024 hold 32 warps)
        // 0 = auto: the biggest block that still gives every SM a tile
        int tb = cfg.segments_block;
        if (tb <= 0) tb = n >= (size_t)cfg.sm_count * 1024 ? 1024 : (n >= (size_t)cfg.sm_count * 512 ? 512 : 256);
        tb = tb >= 1024 ? 1024 : (tb >= 512 ? 512 : 256);
#define EPP_LAUNCH_SEG_RS(TB_)                                                                                        \
    {                                                                                                                 \
        const int bps_tb = bps < 1024 / TB_ ? bps : 1024 / TB_; /* 64 registers per thread: 1024 threads per SM */    \
        const unsigned grid_tb = persistent_grid(cfg, (n + TB_ - 1) / TB_, bps_tb);                                   \
        err = ensure_smem(k_segments_rs<BP, TB_>, total);                                                             \
        if (err == cudaSuccess) {                                                                                     \
            k_segments_rs<BP, TB_><<<grid_tb, TB_, total, stream>>>(wv, d_a, d_b, n, can_pass_gate,                    \
                                                                   (uint32_t)table_bytes, d_mask);                     \
            err = cudaGetLastError();                                                                                 \
        }                                                                                                             \
    }
        EPP_DISPATCH_BP(bp, {
            if (tb == 1024) EPP_LAUNCH_SEG_RS(1024)
            else if (tb == 512) EPP_LAUNCH_SEG_RS(512)
            else EPP_LAUNCH_SEG_RS(256)
        });
#undef EPP_LAUNCH_SEG_RS
        g_launches++;
        if (err == cudaSuccess && gv) err = launch_gather_push(cfg, *gv, stream);
        return err;
    }
    const unsigned grid = persistent_grid(cfg, tiles, cfg.segments_blocks_per_sm);
    EPP_DISPATCH_BP(bp, {
        err = ensure_smem(k_segments<BP>, smem);
        if (err == cudaSuccess) {
            k_segments<BP><<<grid, kBlock, smem, stream>>>(wv, d_a, d_b, n, can_pass_gate, in_smem ? 1 : 0, d_mask);
            err = cudaGetLastError();
        }
    });
    g_launches++;
    if (err == cudaSuccess && gv) err = launch_gather_push(cfg, *gv, stream);
    return err;
}

cudaError_t launch_trajectories(const LaunchCfg &cfg, const WorldView &wv, int bp, const double *d_mats, size_t n_traj,
                                size_t rows, size_t ld, size_t traj_stride, double min_distance, uint32_t *d_traj_valid,
                                uint32_t *d_sample_mask, int64_t *d_first_invalid, long long *d_first_scratch,
                                cudaStream_t stream)
{
    if (n_traj == 0) return cudaSuccess;
    k_fill_i64<<<(unsigned)((n_traj + 255) / 256), 256, 0, stream>>>((int64_t *)d_first_scratch, n_traj, LLONG_MAX);
    g_launches++;
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    bool done = false;
    if (rows > 0 && rows < 0xfffffff0ull && n_traj * rows >= (size_t)cfg.states_warp_min_n && cfg.traj_variant != 1) {
        // warp-pipelined form (k_states_warp in its trajectory mode): columns 0, 3, 6 arrive by cp.async.bulk
        bool in_smem;
        const size_t smem = point_smem_bytes(wv, bp, in_smem);
        size_t table_bytes = ((bp == EPP_BP_GRID ? (size_t)wv.occ_bytes : smem) + 127) / 128 * 128;
        const int rec_max = cfg.states_rec_smem_max > 0 ? (cfg.states_rec_smem_max < 992 ? cfg.states_rec_smem_max : 992) : kRecSmemMax;
        bool rs = wv.M <= rec_max && wv.zrot && cfg.states_rec_smem != 0;
        const int wbp = (bp == EPP_BP_BITS && wv.W == 1 && cfg.states_bits1 != 0) ? kBpBits1 : bp;
        TrajArgs ta = {};
        ta.mats = d_mats;
        ta.rows = rows;
        ta.ld = ld;
        ta.traj_stride = traj_stride;
        ta.aligned = ((reinterpret_cast<uintptr_t>(d_mats) & 15u) == 0 && (ld & 1) == 0 && (traj_stride & 1) == 0) ? 1 : 0;
        ta.sample_mask = d_sample_mask;
        ta.first_row = d_first_scratch;
        const GatherView no_gather = {};
#define EPP_LAUNCH_TRAJ(RS_)                                                                                                \
    err = launch_states_warp<BP, 4, RS_, false, true>(cfg, wv, nullptr, n_traj * rows, kModeMinDist, 0, min_distance, in_smem, \
                                                      table_bytes, 0, nullptr, stream, done, no_gather, ta, n_traj)
        for (int attempt = 0; attempt < 2 && !done && err == cudaSuccess; ++attempt) {
            if (attempt == 1) {
                if (!rs) break;
                rs = false;
            }
            if (wbp == kBpBits1) {
                constexpr int BP = kBpBits1;
                if (rs) EPP_LAUNCH_TRAJ(true); else EPP_LAUNCH_TRAJ(false);
            } else {
                EPP_DISPATCH_BP(bp, { if (rs) EPP_LAUNCH_TRAJ(true); else EPP_LAUNCH_TRAJ(false); });
            }
        }
#undef EPP_LAUNCH_TRAJ
        if (err != cudaSuccess) return err;
        if (done) g_launches++;
    }
    if (rows > 0 && !done) {
        bool in_smem;
        const WorldView wd = without_cmask(wv);
        const size_t smem = point_smem_bytes(wd, bp, in_smem);
        const size_t row_tiles = (rows + kBlock - 1) / kBlock;
        if (row_tiles > 65535 || n_traj > 0x7fffffffull) return cudaErrorInvalidValue;
        dim3 grid((unsigned)n_traj, (unsigned)row_tiles);
        EPP_DISPATCH_BP(bp, {
            err = ensure_smem(k_traj_samples<BP>, smem);
            if (err == cudaSuccess) {
                k_traj_samples<BP><<<grid, kBlock, smem, stream>>>(wd, d_mats, rows, ld, traj_stride, min_distance,
                                                                   in_smem ? 1 : 0, d_sample_mask, d_first_scratch);
                err = cudaGetLastError();
            }
        });
        g_launches++;
        if (err != cudaSuccess) return err;
    }
    k_traj_finalize<<<(unsigned)((n_traj + 255) / 256), 256, 0, stream>>>(d_first_scratch, d_first_invalid, n_traj,
                                                                         d_traj_valid);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_poly_trajectories(const LaunchCfg &cfg, const WorldView &wv_in, int bp, const double *d_coeffs,
                                     const double *d_durations, const double *d_dt, size_t n_traj, int n_segments,
                                     int n_coeffs, double min_distance, size_t max_samples, uint32_t *d_traj_valid,
                                     uint32_t *d_sample_mask, uint32_t *d_n_samples, int64_t *d_first_invalid,
                                     long long *d_first_scratch, cudaStream_t stream)
{
    const WorldView wv = without_cmask(wv_in);
    if (n_traj == 0) return cudaSuccess;
    k_fill_i64<<<(unsigned)((n_traj + 255) / 256), 256, 0, stream>>>((int64_t *)d_first_scratch, n_traj, LLONG_MAX);
    g_launches++;
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    if (d_sample_mask) {
        const size_t words = n_traj * (max_samples / 32);
        k_fill_u32<<<(unsigned)((words + 255) / 256), 256, 0, stream>>>(d_sample_mask, words, 0u);
        g_launches++;
    }
    bool in_smem;
    const size_t smem = point_smem_bytes(wv, bp, in_smem);
    // Trajectories per block G (= lanes of its producer warp that work), threads' registers and ring depth.  The sweep is
    // latency-bound: what matters first is that every SM holds as many blocks as fit, then that blocks split evenly
    // over the SMs, and a block's 7 consumer warps keep up with its producer only up to G ~ 18.  Measured on 10 000 /
    // 20 000 / 40 000 / 2 000 trajectories of ~1000 rows (profiles/r02_fin_ab_poly.log, r02_wide_ab_poly.log):
    //  - one wave of 4 x SMs blocks of 64-register threads (no spills; 5 x SMs of 48-register threads otherwise) with a
    //    ring of 6 rounds while that needs G <= 18: 10 000 trajectories 176 us (625 blocks of 16, the round-2 rule: 5
    //    blocks on some SMs, 4 on others) -> 169 us (715 of 14) -> 158 us (556 of 18, 64 registers, ring 6);
    //  - else blocks of 16 in several waves: 40 000 trajectories 841 us (1250 blocks of 32) -> 579 us; 20 000: 304 us
    //    (one wave of 715 blocks of 28: 505 us; 64-register threads: 333 us);
    //  - batches that cannot fill the SMs with G >= 12: G halves from 32 while there are fewer than 4 blocks per SM
    //    (2000 trajectories: G = 4, 88 us; G = 14 would leave one block per SM: 94 us).
    int G = cfg.poly_traj_per_warp;
    const size_t sms = (size_t)(cfg.sm_count > 0 ? cfg.sm_count : 148);
    // Block shapes (threads x resident blocks per SM, registers per thread): 256 x 4, 63 registers ("wide": batches that
    // run as one wave); 224 x 5, 56 registers (1 producer + 6 consumer warps; everything else by default: no spills,
    // where 256 x 5 at 48 registers reloads spilled loop constants in every work item — 40 000 trajectories 578 -> 515 us,
    // 20 000: 306 -> 290, 2 000: 90 -> 86; 10 000 as one wave: 168 against the wide shape's 160;
    // profiles/r02_slim_ab.log); on request (knob poly_threads) 192 x 6 at 56 registers (40 000: 509 us, 20 000: 296),
    // 256 x 5 at 48 and 512 x 2.
    const bool big = cfg.poly_threads >= 512;
    const bool legacy = cfg.poly_threads == 256; // 256 x 5, 48 registers
    int slim = (cfg.poly_threads == 224 || cfg.poly_threads == 192) ? cfg.poly_threads : 0;
    bool wide = !big && !slim && cfg.poly_blocks_per_sm == 4; // 64 registers per thread, 4 resident blocks
    const bool wide_auto = !big && !slim && !legacy && cfg.poly_blocks_per_sm == 0;
    bool one_wave = false;
    if (G <= 0 && cfg.poly_balance && n_traj >= 12 * 4 * sms) {
        if (wide_auto) wide = n_traj <= 18 * 4 * sms;
        const size_t resident = (size_t)(big ? 2 : (slim == 192 ? 6 : (wide ? 4 : 5))) * sms;
        one_wave = n_traj <= 18 * resident;
        G = one_wave ? (int)((n_traj + resident - 1) / resident) : 16;
        G = (G + 1) & ~1; // pairs of trajectories are the consumers' work items
    } else if (G <= 0) {
        G = 32;
        while (G > 4 && (n_traj + G - 1) / (size_t)G < sms * 4) G >>= 1;
    }
    G = G > 32 ? 32 : (G < 1 ? 1 : G);
    if (!big && !wide && !legacy && !slim) slim = 224;
    int ring = cfg.poly_ring > 0 ? cfg.poly_ring : (wide && one_wave ? 6 : 4);
    ring = ring < 2 ? 2 : (ring > kPolyRingMax ? kPolyRingMax : ring);
    const unsigned grid = (unsigned)((n_traj + G - 1) / G);
    const size_t table_bytes = (smem + 127) / 128 * 128;
    const size_t total = table_bytes + (size_t)ring * ((size_t)G * kPolyPitch * 10 + (size_t)G * 4 + 32);
    // 16-byte coefficient loads need a 16 B-aligned batch (a segment's 3 x 10 coefficients are 240 B)
    const bool n10 = n_coeffs == 10 && (reinterpret_cast<uintptr_t>(d_coeffs) & 15u) == 0;
    EPP_DISPATCH_BP(bp, {
        auto kern = big    ? (n10 ? k_poly_traj<BP, true, 512> : k_poly_traj<BP, false, 512>)
                    : wide ? (n10 ? k_poly_traj<BP, true, 256, 4> : k_poly_traj<BP, false, 256, 4>)
                           : (n10 ? k_poly_traj<BP, true, 256> : k_poly_traj<BP, false, 256>);
        if (slim == 224) kern = n10 ? k_poly_traj<BP, true, 224, 5> : k_poly_traj<BP, false, 224, 5>;
        if (slim == 192) kern = n10 ? k_poly_traj<BP, true, 192, 6> : k_poly_traj<BP, false, 192, 6>;
        err = ensure_smem(kern, total);
        if (err == cudaSuccess) {
            kern<<<grid, big ? 512 : (slim ? slim : 256), total, stream>>>(wv, d_coeffs, d_durations, d_dt, n_traj, n_segments, n_coeffs,
                                                        min_distance, max_samples, in_smem ? 1 : 0, G, (uint32_t)table_bytes,
                                                        ring, cfg.poly_wait ? 1 : 0, d_sample_mask, d_n_samples, d_first_scratch);
            err = cudaGetLastError();
        }
    });
    g_launches++;
    if (err != cudaSuccess) return err;
    k_traj_finalize<<<(unsigned)((n_traj + 255) / 256), 256, 0, stream>>>(d_first_scratch, d_first_invalid, n_traj,
                                                                         d_traj_valid);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_mailbox(const WorldView &wv, int bp, MailboxSlot *slot, uint32_t last_seq, uint64_t idle_ns,
                           cudaStream_t stream)
{
    cudaError_t err = cudaSuccess;
    if (bp == EPP_BP_GRID) k_mailbox<EPP_BP_GRID><<<1, 32, 0, stream>>>(wv, slot, last_seq, (unsigned long long)idle_ns);
    else k_mailbox<EPP_BP_BRUTE><<<1, 32, 0, stream>>>(wv, slot, last_seq, (unsigned long long)idle_ns); // all OBBs, 32 at a time
    err = cudaGetLastError();
    g_launches++;
    return err;
}

cudaError_t launch_count_candidates(const LaunchCfg &cfg, const WorldView &wv, int bp, const double *d_a, const double *d_b,
                                    size_t n, uint64_t *d_out, cudaStream_t stream)
{
    cudaError_t err = cudaMemsetAsync(d_out, 0, 2 * sizeof(uint64_t), stream);
    if (err != cudaSuccess || n == 0) return err;
    bool in_smem;
    const size_t smem = d_b ? ray_smem_bytes(wv, bp, in_smem) : point_smem_bytes(wv, bp, in_smem);
    const unsigned grid = persistent_grid(cfg, (n + kBlock - 1) / kBlock, 4);
    EPP_DISPATCH_BP(bp, {
        err = ensure_smem(k_count_candidates<BP>, smem);
        if (err == cudaSuccess) {
            k_count_candidates<BP><<<grid, kBlock, smem, stream>>>(wv, d_a, d_b, n, in_smem ? 1 : 0, (unsigned long long *)d_out);
            err = cudaGetLastError();
        }
    });
    g_launches++;
    return err;
}

cudaError_t launch_gen_states(uint64_t seed, uint64_t first, size_t n, const double lo[3], const double hi[3],
                              double *d_xyz, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    Box3 box;
    for (int d = 0; d < 3; ++d) {
        box.lo[d] = lo[d];
        box.hi[d] = hi[d];
    }
    const size_t elems = 3 * n;
    k_gen_states<<<(unsigned)((elems + 255) / 256), 256, 0, stream>>>(seed, first, n, box, d_xyz);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_gen_segments(uint64_t seed, uint64_t first, size_t n, const double lo[3], const double hi[3],
                                double max_len, double *d_a, double *d_b, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    Box3 box;
    for (int d = 0; d < 3; ++d) {
        box.lo[d] = lo[d];
        box.hi[d] = hi[d];
    }
    k_gen_segments<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(seed, first, n, box, max_len, d_a, d_b);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_mask_summary(const LaunchCfg &cfg, const uint32_t *d_mask, size_t n_words, uint64_t first_word,
                                uint64_t *d_out, cudaStream_t stream)
{
    cudaError_t err = cudaMemsetAsync(d_out, 0, 2 * sizeof(uint64_t), stream);
    if (err != cudaSuccess || n_words == 0) return err;
    size_t blocks = (n_words + 255) / 256;
    const size_t cap = (size_t)cfg.sm_count * 8;
    if (blocks > cap) blocks = cap;
    k_mask_summary<<<(unsigned)blocks, 256, 0, stream>>>(d_mask, n_words, first_word, (unsigned long long *)d_out);
    g_launches++;
    return cudaGetLastError();
}

} // namespace epp
