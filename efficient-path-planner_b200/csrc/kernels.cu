// kernels.cu — sm_100a kernels of the validity-check hot path and their launchers.
//
// K1  k_states        World::checkPointValidity (both overloads) over an n x 3 AoS batch
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
constexpr int kWarpsPerBlock = kBlock / 32;

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
template <>
struct PointWorld<EPP_BP_BITS> {
    const uint32_t *inside;
    __device__ void init(const WorldView &wv, bool)
    {
        uint32_t *dst = reinterpret_cast<uint32_t *>(g_smem);
        block_copy_words(dst, wv.bits_inside, 3 * wv.nb * wv.W);
        __syncthreads();
        inside = dst;
    }
    __device__ static uint32_t table(const WorldView &wv, bool, const void *&src)
    {
        src = wv.bits_inside;
        return (uint32_t)(3 * wv.nb * wv.W * 4);
    }
    __device__ void bind(const WorldView &, bool) { inside = reinterpret_cast<const uint32_t *>(g_smem); }
    template <class F>
    __device__ __forceinline__ bool any(const WorldView &wv, const P3 &p, F test) const
    {
        const int nb = wv.nb, W = wv.W;
        const uint32_t *tx = inside + dev_key_bin(p.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb) * W;
        const uint32_t *ty = inside + (nb + dev_key_bin(p.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb)) * W;
        const uint32_t *tz = inside + (2 * nb + dev_key_bin(p.z, wv.kb.off[2], wv.kb.h0[2], wv.kb.shift[2], nb)) * W;
        for (int w = 0; w < W; ++w) {
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
        const int nb = wv.nb, W = wv.W;
        const uint32_t *tx = inside + dev_key_bin(p.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb) * W;
        const uint32_t *ty = inside + (nb + dev_key_bin(p.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb)) * W;
        const uint32_t *tz = inside + (2 * nb + dev_key_bin(p.z, wv.kb.off[2], wv.kb.h0[2], wv.kb.shift[2], nb)) * W;
        if (W == 1) return (tx[0] & ty[0] & tz[0]) != 0u;
        uint32_t m = 0u;
        for (int w = 0; w < W; ++w) m |= tx[w] & ty[w] & tz[w];
        return m != 0u;
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

// Uniform grid: candidates = CSR list of the cell holding p.
template <>
struct PointWorld<EPP_BP_GRID> {
    __device__ void init(const WorldView &, bool) {}
    __device__ static uint32_t table(const WorldView &, bool, const void *&src)
    {
        src = nullptr;
        return 0u;
    }
    __device__ void bind(const WorldView &, bool) {}
    __device__ __forceinline__ static void cell_range(const WorldView &wv, const P3 &p, uint32_t &s, uint32_t &e)
    {
        const int bx = dev_bin(p.x, wv.gbins.lo[0], wv.gbins.inv[0], wv.gbins.n[0]);
        const int by = dev_bin(p.y, wv.gbins.lo[1], wv.gbins.inv[1], wv.gbins.n[1]);
        const int bz = dev_bin(p.z, wv.gbins.lo[2], wv.gbins.inv[2], wv.gbins.n[2]);
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
    __device__ __forceinline__ bool has_candidates(const WorldView &wv, const P3 &p) const
    {
        uint32_t s, e;
        cell_range(wv, p, s, e);
        return e != s;
    }
    __device__ __forceinline__ bool invalid(const WorldView &wv, const P3 &p, int mode, bool cpg, double md) const
    {
        return any(wv, p, [&](int k) { return point_candidate_invalid(wv.aabb, wv.obb, k, p, mode, cpg, md); });
    }
};

// ---- segment queries ---------------------------------------------------------------------

template <int BP>
struct RayWorld;

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

// candidates = AND over axes of max_ge[bin(rayMin)] & min_le[bin(rayMax)].
template <>
struct RayWorld<EPP_BP_BITS> {
    const uint32_t *min_le;
    const uint32_t *max_ge;
    __device__ void init(const WorldView &wv, bool)
    {
        const int words = 3 * wv.nb * wv.W;
        uint32_t *dst = reinterpret_cast<uint32_t *>(g_smem);
        block_copy_words(dst, wv.bits_min_le, words);
        block_copy_words(dst + words, wv.bits_max_ge, words);
        __syncthreads();
        min_le = dst;
        max_ge = dst + words;
    }
    __device__ __forceinline__ bool invalid(const WorldView &wv, const P3 &a, const P3 &b, bool cpg) const
    {
        P3 mn, mx;
        ray_box(a, b, mn, mx);
        const int nb = wv.nb, W = wv.W;
        const uint32_t *gx = max_ge + dev_key_bin(mn.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb) * W;
        const uint32_t *gy = max_ge + (nb + dev_key_bin(mn.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb)) * W;
        const uint32_t *gz = max_ge + (2 * nb + dev_key_bin(mn.z, wv.kb.off[2], wv.kb.h0[2], wv.kb.shift[2], nb)) * W;
        const uint32_t *lx = min_le + dev_key_bin(mx.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb) * W;
        const uint32_t *ly = min_le + (nb + dev_key_bin(mx.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb)) * W;
        const uint32_t *lz = min_le + (2 * nb + dev_key_bin(mx.z, wv.kb.off[2], wv.kb.h0[2], wv.kb.shift[2], nb)) * W;
        for (int w = 0; w < W; ++w) {
            uint32_t m = gx[w] & gy[w] & gz[w] & lx[w] & ly[w] & lz[w];
            while (m) {
                const int k = (w << 5) + __ffs(m) - 1;
                m &= m - 1;
                if (ray_candidate_invalid(wv.aabb, wv.obb, k, a, b, mn, mx, cpg)) return true;
            }
        }
        return false;
    }
};

// Cells overlapped by the ray box; a candidate listed in several of them is evaluated only
// in the first cell of (ray cells) x (candidate cells).
template <>
struct RayWorld<EPP_BP_GRID> {
    __device__ void init(const WorldView &, bool) {}
    __device__ __forceinline__ bool invalid(const WorldView &wv, const P3 &a, const P3 &b, bool cpg) const
    {
        P3 mn, mx;
        ray_box(a, b, mn, mx);
        const AxisBins &g = wv.gbins;
        const int x0 = dev_bin(mn.x, g.lo[0], g.inv[0], g.n[0]), x1 = dev_bin(mx.x, g.lo[0], g.inv[0], g.n[0]);
        const int y0 = dev_bin(mn.y, g.lo[1], g.inv[1], g.n[1]), y1 = dev_bin(mx.y, g.lo[1], g.inv[1], g.n[1]);
        const int z0 = dev_bin(mn.z, g.lo[2], g.inv[2], g.n[2]), z1 = dev_bin(mx.z, g.lo[2], g.inv[2], g.n[2]);
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
                        if (ray_candidate_invalid(wv.aabb, wv.obb, k, a, b, mn, mx, cpg)) return true;
                    }
                }
            }
        return false;
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

constexpr int kMaxStages = 8;

// ---- K1 (streaming form): TMA bulk-copy ring + block-wide candidate compaction -----------------------
//
// Persistent blocks; tile = S * kBlock consecutive states = one contiguous span of the AoS batch that
// thread 0 fetches with cp.async.bulk into a ring of `stages` shared-memory slots.
//
// The direct kernel above lets every lane walk its own candidate list, so a warp pays the full exact test
// whenever ANY lane has a candidate, at ~50 % lane utilisation and with one dependent chain per warp.
// Here a tile is handled in two dense phases:
//   phase 1  every thread locates the candidates of its S states (independent chains -> ILP) and appends
//            (state, OBB) pairs to a block-wide queue; one shared-memory atomic per warp reserves the space;
//   phase 2  the block walks the queue with all lanes busy, running the reference's exact per-candidate
//            test (strict stored-AABB test, type rule, OBB test) and OR-ing hits into an invalid bitmap.
// The verdict of a state is an OR over its candidates, so neither the order of the queue nor the loss of the
// early exit changes a single bit.  Pairs that do not fit the queue are tested in place by their lane.
constexpr uint32_t kPairSentinel = 0xffffffffu;
constexpr int kPairUnitShift = 22; // pair = unit << 22 | obb index (M < 2^22)

template <int BP>
struct PointCand;
template <>
struct PointCand<EPP_BP_BRUTE> {};
template <>
struct PointCand<EPP_BP_BITS> {
    int ox, oy, oz; // row offsets (in words) of the three bin rows
    uint32_t m;     // W == 1: the candidate word itself
};
template <>
struct PointCand<EPP_BP_GRID> {
    uint32_t s, e;
};

template <int BP>
struct PointEnum;

template <>
struct PointEnum<EPP_BP_BRUTE> {
    __device__ __forceinline__ static int locate(const PointWorld<EPP_BP_BRUTE> &w, const WorldView &wv, const P3 &p,
                                                 PointCand<EPP_BP_BRUTE> &)
    {
        int c = 0;
        for (int k = 0; k < wv.M; ++k) c += within_strict(w.aabb + k, p) ? 1 : 0;
        return c;
    }
    template <class F>
    __device__ __forceinline__ static void each(const PointWorld<EPP_BP_BRUTE> &w, const WorldView &wv, const P3 &p,
                                                const PointCand<EPP_BP_BRUTE> &, F f)
    {
        for (int k = 0; k < wv.M; ++k)
            if (within_strict(w.aabb + k, p)) f(k);
    }
};

template <>
struct PointEnum<EPP_BP_BITS> {
    __device__ __forceinline__ static int locate(const PointWorld<EPP_BP_BITS> &w, const WorldView &wv, const P3 &p,
                                                 PointCand<EPP_BP_BITS> &c)
    {
        const int nb = wv.nb, W = wv.W;
        const int bx = dev_key_bin(p.x, wv.kb.off[0], wv.kb.h0[0], wv.kb.shift[0], nb);
        const int by = nb + dev_key_bin(p.y, wv.kb.off[1], wv.kb.h0[1], wv.kb.shift[1], nb);
        const int bz = 2 * nb + dev_key_bin(p.z, wv.kb.off[2], wv.kb.h0[2], wv.kb.shift[2], nb);
        if (W == 1) { // worlds of up to 32 OBBs (the race track): one word per bin, no loops
            c.m = w.inside[bx] & w.inside[by] & w.inside[bz];
            return __popc(c.m);
        }
        c.ox = bx * W;
        c.oy = by * W;
        c.oz = bz * W;
        int n = 0;
        for (int i = 0; i < W; ++i) n += __popc(w.inside[c.ox + i] & w.inside[c.oy + i] & w.inside[c.oz + i]);
        return n;
    }
    template <class F>
    __device__ __forceinline__ static void each(const PointWorld<EPP_BP_BITS> &w, const WorldView &wv, const P3 &,
                                                const PointCand<EPP_BP_BITS> &c, F f)
    {
        if (wv.W == 1) {
            uint32_t m = c.m;
            while (m) {
                f(__ffs(m) - 1);
                m &= m - 1;
            }
            return;
        }
        for (int i = 0; i < wv.W; ++i) {
            uint32_t m = w.inside[c.ox + i] & w.inside[c.oy + i] & w.inside[c.oz + i];
            while (m) {
                f((i << 5) + __ffs(m) - 1);
                m &= m - 1;
            }
        }
    }
};

template <>
struct PointEnum<EPP_BP_GRID> {
    __device__ __forceinline__ static int locate(const PointWorld<EPP_BP_GRID> &, const WorldView &wv, const P3 &p,
                                                 PointCand<EPP_BP_GRID> &c)
    {
        const int bx = dev_bin(p.x, wv.gbins.lo[0], wv.gbins.inv[0], wv.gbins.n[0]);
        const int by = dev_bin(p.y, wv.gbins.lo[1], wv.gbins.inv[1], wv.gbins.n[1]);
        const int bz = dev_bin(p.z, wv.gbins.lo[2], wv.gbins.inv[2], wv.gbins.n[2]);
        const size_t cell = ((size_t)bz * wv.gbins.n[1] + by) * wv.gbins.n[0] + bx;
        c.s = __ldg(wv.cell_start + cell);
        c.e = __ldg(wv.cell_start + cell + 1);
        return (int)(c.e - c.s);
    }
    template <class F>
    __device__ __forceinline__ static void each(const PointWorld<EPP_BP_GRID> &, const WorldView &wv, const P3 &,
                                                const PointCand<EPP_BP_GRID> &c, F f)
    {
        for (uint32_t j = c.s; j < c.e; ++j) f((int)__ldg(wv.cell_items + j));
    }
};

// Shared-memory bookkeeping of the compaction, double buffered by tile parity.
template <int UNITS>
struct PairState {
    int count[2];
    uint32_t inv[2][UNITS / 32];
};

// One tile: `units` (<= S*kBlock) states at `pts` (AoS doubles in shared memory).  Ends with the block's
// mask words written; contains two __syncthreads.  buf alternates 0/1 per call.
template <int BP, int S>
__device__ __forceinline__ void states_tile(const PointWorld<BP> &world, const WorldView &wv, const double *pts,
                                            int units, int mode, bool cpg, double md, uint32_t *queue, int qcap,
                                            PairState<S * kBlock> &ps, int buf, uint32_t *__restrict__ mask_out)
{
    // ---- phase 1
    P3 p[S];
    PointCand<BP> cand[S];
    int cnt[S];
    int total = 0;
#pragma unroll
    for (int u = 0; u < S; ++u) {
        const int unit = u * kBlock + threadIdx.x;
        cnt[u] = 0;
        if (unit < units) {
            p[u].x = pts[3 * unit];
            p[u].y = pts[3 * unit + 1];
            p[u].z = pts[3 * unit + 2];
            cnt[u] = PointEnum<BP>::locate(world, wv, p[u], cand[u]);
        }
        total += cnt[u];
    }
    if (total != 0) {
        // one shared-memory atomic per lane that has candidates (a minority of lanes in sparse worlds)
        int base = atomicAdd(&ps.count[buf], total);
        if (base + total <= qcap) {
#pragma unroll
            for (int u = 0; u < S; ++u) {
                if (cnt[u] == 0) continue;
                const uint32_t tag = (uint32_t)(u * kBlock + threadIdx.x) << kPairUnitShift;
                PointEnum<BP>::each(world, wv, p[u], cand[u], [&](int k) { queue[base++] = tag | (uint32_t)k; });
            }
        } else { // queue full: blank my reserved slots and test in place
            for (int j = base; j < qcap && j < base + total; ++j) queue[j] = kPairSentinel;
#pragma unroll
            for (int u = 0; u < S; ++u) {
                if (cnt[u] == 0) continue;
                bool hit = false;
                PointEnum<BP>::each(world, wv, p[u], cand[u], [&](int k) {
                    hit = hit || point_candidate_invalid(wv.aabb, wv.obb, k, p[u], mode, cpg, md);
                });
                const int unit = u * kBlock + threadIdx.x;
                if (hit) atomicOr(&ps.inv[buf][unit >> 5], 1u << (unit & 31));
            }
        }
    }
    __syncthreads();
    // ---- phase 2
    const int pairs = ps.count[buf] < qcap ? ps.count[buf] : qcap;
    for (int q = threadIdx.x; q < pairs; q += kBlock) {
        const uint32_t pr = queue[q];
        if (pr == kPairSentinel) continue;
        const int unit = (int)(pr >> kPairUnitShift);
        const int k = (int)(pr & ((1u << kPairUnitShift) - 1u));
        const P3 q3 = {pts[3 * unit], pts[3 * unit + 1], pts[3 * unit + 2]};
        if (point_candidate_invalid(wv.aabb, wv.obb, k, q3, mode, cpg, md))
            atomicOr(&ps.inv[buf][unit >> 5], 1u << (unit & 31));
    }
    __syncthreads();
    // ---- verdicts: bit = valid; bits of units beyond `units` stay 0
    if (threadIdx.x < (units + 31) / 32) {
        const int w = threadIdx.x;
        const int live = units - 32 * w;
        const uint32_t active = live >= 32 ? 0xffffffffu : ((1u << live) - 1u);
        mask_out[w] = ~ps.inv[buf][w] & active;
        ps.inv[buf][w] = 0u;
    }
    if (threadIdx.x == 0) ps.count[buf] = 0;
}

template <int BP, int S>
__global__ void __launch_bounds__(kBlock) k_states_stream(WorldView wv, const double *__restrict__ xyz, size_t n, int mode,
                                                          int can_pass_gate, double min_distance, int aabb_in_smem,
                                                          int stages, uint32_t table_bytes, int qcap,
                                                          uint32_t *__restrict__ mask)
{
    constexpr int kTile = S * kBlock;
    constexpr uint32_t kTileBytes = kTile * 3 * sizeof(double);
    __shared__ __align__(8) uint64_t s_full[kMaxStages];
    __shared__ __align__(8) uint64_t s_table;
    __shared__ PairState<kTile> s_ps;
    unsigned char *ring = g_smem + table_bytes; // table_bytes is a multiple of 128
    uint32_t *queue = reinterpret_cast<uint32_t *>(ring + (size_t)stages * kTileBytes);
    const size_t full_tiles = n / kTile;
    const size_t first = blockIdx.x;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) mbar_init(&s_full[s], 1);
        mbar_init(&s_table, 1);
        mbar_init_fence();
        // broad-phase table first (needed by every tile), then the first `stages` tiles
        const void *tsrc;
        const uint32_t tbytes = (PointWorld<BP>::table(wv, aabb_in_smem != 0, tsrc) + 15u) & ~15u; // blob parts are padded
        if (tbytes) {
            mbar_expect_tx(&s_table, tbytes);
            bulk_g2s(g_smem, tsrc, tbytes, &s_table);
        }
        for (int s = 0; s < stages; ++s) {
            const size_t tile = first + (size_t)s * gridDim.x;
            if (tile < full_tiles) {
                mbar_expect_tx(&s_full[s], kTileBytes);
                bulk_g2s(ring + (size_t)s * kTileBytes, xyz + tile * (size_t)(3 * kTile), kTileBytes, &s_full[s]);
            }
        }
        s_ps.count[0] = s_ps.count[1] = 0;
    }
    if (threadIdx.x < kTile / 32) s_ps.inv[0][threadIdx.x] = s_ps.inv[1][threadIdx.x] = 0u;
    PointWorld<BP> world;
    world.bind(wv, aabb_in_smem != 0);
    __syncthreads(); // publishes the barriers and the zeroed bookkeeping
    {
        const void *tsrc;
        if (PointWorld<BP>::table(wv, aabb_in_smem != 0, tsrc)) mbar_wait(&s_table, 0);
    }

    const bool cpg = can_pass_gate != 0;
    int s = 0, buf = 0;
    uint32_t parity = 0;
    for (size_t tile = first; tile < full_tiles; tile += gridDim.x) {
        mbar_wait(&s_full[s], parity);
        const double *pts = reinterpret_cast<const double *>(ring + (size_t)s * kTileBytes);
        states_tile<BP, S>(world, wv, pts, kTile, mode, cpg, min_distance, queue, qcap, s_ps, buf,
                           mask + tile * (kTile / 32));
        // every read of the slot happened before the second barrier inside states_tile
        if (threadIdx.x == 0) {
            const size_t nxt = tile + (size_t)stages * gridDim.x;
            if (nxt < full_tiles) {
                mbar_expect_tx(&s_full[s], kTileBytes);
                bulk_g2s(ring + (size_t)s * kTileBytes, xyz + nxt * (size_t)(3 * kTile), kTileBytes, &s_full[s]);
            }
        }
        buf ^= 1;
        if (++s == stages) {
            s = 0;
            parity ^= 1u;
        }
    }

    // ragged tail (n % kTile states): plain loads into slot 0, same tile routine
    const size_t tail0 = full_tiles * kTile;
    if (tail0 < n && blockIdx.x == (unsigned)(full_tiles % gridDim.x)) {
        const int rem = (int)(n - tail0);
        double *slot = reinterpret_cast<double *>(ring);
        __syncthreads(); // phase 2 of another warp may still read slot 0 only before the barriers above; be explicit
        for (int e = threadIdx.x; e < 3 * rem; e += kBlock) slot[e] = __ldg(xyz + 3 * tail0 + e);
        __syncthreads();
        states_tile<BP, S>(world, wv, slot, rem, mode, cpg, min_distance, queue, qcap, s_ps, buf,
                           mask + tail0 / 32);
    }
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
constexpr int kRecDoubles = 17;
constexpr int kRecSmemMax = 64;
constexpr int kWordBytes = 32 * 3 * 8; // one mask word of states = 768 B of coordinates

__device__ __forceinline__ bool rec_hit(const double *__restrict__ r, const P3 &p)
{
    const bool in = (p.x > r[0]) & (p.x < r[3]) & (p.y > r[1]) & (p.y < r[4]) & (p.z > r[2]) & (p.z < r[5]);
    const double d0 = dsub(p.x, r[6]), d1 = dsub(p.y, r[7]), d2 = dsub(p.z, r[8]);
    const double lx = dadd(dmul(r[9], d0), dmul(r[11], d1));  // R(0,0) d0 + R(1,0) d1   (src/OBB.cpp:66)
    const double ly = dadd(dmul(r[10], d0), dmul(r[12], d1)); // R(0,1) d0 + R(1,1) d1
    return in & (fabs(lx) <= r[13]) & (fabs(ly) <= r[14]) & (fabs(d2) <= r[15]);
}

template <int BP, int S, bool RS>
__device__ __forceinline__ void warp_states_step(const PointWorld<BP> &world, const WorldView &wv,
                                                 const double *__restrict__ rec, const double *__restrict__ pts,
                                                 int units, int mode, bool cpg, double md, unsigned char *queue,
                                                 uint32_t *inv, uint32_t *__restrict__ mask_out, int lane)
{
    // ---- phase 1: which of my S states have a broad-phase candidate at all
    const uint32_t lt = (1u << lane) - 1u;
    int queued = 0;
    const int dbg = mode >> 8; // profiling only (epp_set_tuning states_debug): 1 skips phase 2, 2 also phase 1
    mode &= 0xff;
#pragma unroll
    for (int u = 0; u < S; ++u) {
        const int unit = u * 32 + lane;
        bool has = false;
        if (unit < units) {
            const P3 p = {pts[3 * unit], pts[3 * unit + 1], pts[3 * unit + 2]};
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
            const P3 p = {pts[3 * unit], pts[3 * unit + 1], pts[3 * unit + 2]};
            const bool hit = world.any(wv, p, [&](int k) -> bool {
                if (RS) return rec_hit(rec + k * kRecDoubles, p);
                return point_candidate_invalid(wv.aabb, wv.obb, k, p, mode, cpg, md);
            });
            if (hit) atomicOr(&inv[unit >> 5], 1u << (unit & 31));
        }
        __syncwarp();
    }
    // ---- verdicts: bit = valid; bits of units beyond `units` stay 0
    if (lane < S && lane * 32 < units) {
        const int live = units - 32 * lane;
        const uint32_t active = live >= 32 ? 0xffffffffu : ((1u << live) - 1u);
        mask_out[lane] = ~inv[lane] & active;
        inv[lane] = 0u;
    }
    __syncwarp();
}

// Warp g (of gridDim.x * kWarpsPerBlock) owns words [g*wpw + min(g, rem), +wpw + (g < rem)); the host passes
// wpw = words / warps and rem = words % warps so no thread divides.
template <int BP, int S, bool RS>
__global__ void __launch_bounds__(kBlock) k_states_warp(WorldView wv, const double *__restrict__ xyz, size_t n, int mode,
                                                        int can_pass_gate, double min_distance, int aabb_in_smem,
                                                        int stages, uint32_t table_bytes, uint32_t rec_bytes,
                                                        uint32_t wpw, uint32_t rem, uint32_t *__restrict__ mask)
{
    __shared__ __align__(8) uint64_t s_full[kWarpsPerBlock][kMaxStages];
    __shared__ __align__(8) uint64_t s_table;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool cpg = can_pass_gate != 0;
    constexpr uint32_t kSlotBytes = S * kWordBytes;

    // shared memory: [broad-phase table][records][per warp: ring | byte queue (S*32) | invalid words (32 B)]
    double *rec = reinterpret_cast<double *>(g_smem + table_bytes);
    const uint32_t ring_bytes = (uint32_t)stages * kSlotBytes;
    const uint32_t warp_bytes = ring_bytes + S * 32u + 32u;
    unsigned char *mine = g_smem + table_bytes + rec_bytes + (uint32_t)warp * warp_bytes;
    unsigned char *queue = mine + ring_bytes;
    uint32_t *inv = reinterpret_cast<uint32_t *>(queue + S * 32);

    const uint32_t g = blockIdx.x * kWarpsPerBlock + warp;
    const uint32_t w0 = g * wpw + (g < rem ? g : rem);
    const uint32_t w1 = w0 + wpw + (g < rem ? 1u : 0u);
    const uint32_t words_total = (uint32_t)((n + 31) / 32);
    const bool aligned = (reinterpret_cast<uintptr_t>(xyz) & 15u) == 0;
    // states in the chunk that starts at word cw (< w1)
    auto chunk_units = [&](uint32_t cw) -> int {
        const uint32_t sw = (w1 - cw) < (uint32_t)S ? (w1 - cw) : (uint32_t)S;
        int units = (int)(sw * 32u);
        if (cw + sw == words_total) units = (int)(n - (size_t)cw * 32u); // the batch's ragged last word
        return units;
    };
    // cp.async.bulk moves multiples of 16 B from 16 B-aligned addresses: an odd number of states (only
    // possible in the last chunk of the batch) or an 8 B-aligned batch is loaded by the lanes instead.
    auto issue = [&](uint32_t cw, int slot) {
        const uint32_t bytes = (uint32_t)chunk_units(cw) * 24u;
        mbar_expect_tx(&s_full[warp][slot], bytes);
        bulk_g2s(mine + (uint32_t)slot * kSlotBytes, xyz + (size_t)cw * 96, bytes, &s_full[warp][slot]);
    };
    auto by_tma = [&](uint32_t cw) -> bool { return aligned && (chunk_units(cw) & 1) == 0; };

    if (lane == 0) {
        for (int s = 0; s < stages; ++s) mbar_init(&s_full[warp][s], 1);
        if (warp == 0) mbar_init(&s_table, 1);
        mbar_init_fence();
        if (warp == 0) {
            const void *tsrc;
            const uint32_t tbytes = (PointWorld<BP>::table(wv, aabb_in_smem != 0, tsrc) + 15u) & ~15u;
            if (tbytes) {
                mbar_expect_tx(&s_table, tbytes);
                bulk_g2s(g_smem, tsrc, tbytes, &s_table);
            }
        }
        uint32_t cw = w0;
        for (int s = 0; s < stages && cw < w1; ++s, cw += S)
            if (by_tma(cw)) issue(cw, s);
    }
    if (lane < 8) inv[lane] = 0u;
    if (RS) {
        for (int k = threadIdx.x; k < wv.M; k += kBlock) {
            const AabbRec *a = wv.aabb + k;
            const ObbRec *o = wv.obb + k;
            double *r = rec + k * kRecDoubles;
            for (int d = 0; d < 3; ++d) {
                r[d] = a->mn[d];
                r[3 + d] = a->mx[d];
                r[6 + d] = o->center[d];
            }
            r[9] = o->rot[0];
            r[10] = o->rot[1];
            r[11] = o->rot[3];
            r[12] = o->rot[4];
            const bool filling = (o->flags & kFlagFilling) != 0;
            const bool skip = (mode & 0xff) == kModeBool ? (filling && cpg) : filling; // src/World.cpp:92-95, :116-119
            const double infl = (mode & 0xff) == kModeBool ? o->inflate : min_distance; // :97, :121
            for (int d = 0; d < 3; ++d)
                r[13 + d] = skip ? -1.0 : (filling ? o->half[d] : dadd(o->half[d], infl)); // src/OBB.cpp:67-72
            r[16] = 0.0;
        }
    }
    PointWorld<BP> world;
    world.bind(wv, aabb_in_smem != 0);
    __syncthreads(); // publishes s_table's init, the records and the zeroed bitmaps
    {
        const void *tsrc;
        if (PointWorld<BP>::table(wv, aabb_in_smem != 0, tsrc)) mbar_wait(&s_table, 0);
    }

    int slot = 0;
    uint32_t parity = 0;
    for (uint32_t cw = w0; cw < w1; cw += S) {
        const int units = chunk_units(cw);
        double *pts = reinterpret_cast<double *>(mine + (uint32_t)slot * kSlotBytes);
        if (by_tma(cw)) {
            mbar_wait(&s_full[warp][slot], parity);
        } else {
            const double *src = xyz + (size_t)cw * 96;
            for (int e = lane; e < 3 * units; e += 32) pts[e] = __ldg(src + e);
            __syncwarp();
        }
        warp_states_step<BP, S, RS>(world, wv, rec, pts, units, mode, cpg, min_distance, queue, inv, mask + cw, lane);
        // the step ended with __syncwarp: every lane's reads of the slot are done
        const uint32_t nxt = cw + (uint32_t)stages * S;
        if (lane == 0 && nxt < w1 && by_tma(nxt)) issue(nxt, slot);
        if (++slot == stages) {
            slot = 0;
            parity ^= 1u;
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

constexpr int kPolyWarps = 4;
constexpr int kPolyPitch = 33; // odd pitch: conflict-free column writes and row reads

// Polynomial::evaluate(t, 0) (polynomial.h:136-150): Horner from the top coefficient.
__device__ __forceinline__ double horner(const double *__restrict__ c, int n, double t)
{
    double r = __ldg(c + n - 1);
    for (int j = n - 2; j >= 0; --j) r = dadd(dmul(r, t), __ldg(c + j));
    return r;
}

// A warp owns 32 trajectories.  Phase 1: lane l advances the reference's sequential sampler
// (trajectory.cpp:121-140: running sums, not k*dt) of ITS trajectory by up to 32 samples and
// parks (segment, time_in_segment) in shared memory.  Phase 2: for each of the 32 trajectories
// in turn, the 32 lanes evaluate its 32 parked samples (Horner x 3 + world check) and ballot
// them into one mask word.
template <int BP>
__global__ void __launch_bounds__(kPolyWarps * 32) k_poly_traj(WorldView wv, const double *__restrict__ coeffs,
                                                               const double *__restrict__ durations,
                                                               const double *__restrict__ dts, size_t n_traj,
                                                               int n_segments, int n_coeffs, double min_distance,
                                                               size_t max_samples, int aabb_in_smem,
                                                               uint32_t *__restrict__ sample_mask,
                                                               uint32_t *__restrict__ n_samples_out,
                                                               long long *__restrict__ first_row)
{
    __shared__ double s_t[kPolyWarps][32 * kPolyPitch];
    __shared__ unsigned short s_seg[kPolyWarps][32 * kPolyPitch];
    PointWorld<BP> world;
    world.init(wv, aabb_in_smem != 0);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t traj0 = ((size_t)blockIdx.x * kPolyWarps + warp) * 32;
    if (traj0 >= n_traj) return;
    const size_t my = traj0 + lane;
    const bool have = my < n_traj;
    const size_t words_per_traj = max_samples / 32;

    // Sampler state of my trajectory: evaluateRange(0, maxTime, dt), trajectory.cpp:95-119.
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

    for (size_t chunk = 0; chunk < words_per_traj; ++chunk) {
        // ---- phase 1: up to 32 samples of my trajectory
        int cnt = 0;
        while (!done && cnt < 32) {
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
            s_t[warp][lane * kPolyPitch + cnt] = tin; // :132
            s_seg[warp][lane * kPolyPitch + cnt] = (unsigned short)seg;
            ++cnt;
            tin = dadd(tin, dt); // :138-139
            acc = dadd(acc, dt);
        }
        emitted_total += (uint32_t)cnt;
        __syncwarp();
        if (__ballot_sync(0xffffffffu, cnt > 0) == 0) break;

        // ---- phase 2: evaluate trajectory j's parked samples, one per lane
        for (int j = 0; j < 32; ++j) {
            const int cj = __shfl_sync(0xffffffffu, cnt, j);
            if (cj == 0) continue;
            const size_t tj = traj0 + j;
            bool valid = false;
            if (lane < cj) {
                const double t = s_t[warp][j * kPolyPitch + lane];
                const int sg = s_seg[warp][j * kPolyPitch + lane];
                const double *c = coeffs + ((tj * n_segments + sg) * 3) * (size_t)n_coeffs; // segment.cpp:51-58
                const P3 p = {horner(c, n_coeffs, t), horner(c + n_coeffs, n_coeffs, t),
                              horner(c + 2 * n_coeffs, n_coeffs, t)};
                valid = !world.invalid(wv, p, kModeMinDist, false, min_distance);
            }
            const uint32_t word = __ballot_sync(0xffffffffu, valid);
            if (lane == 0) {
                if (sample_mask) sample_mask[tj * words_per_traj + chunk] = word;
                const uint32_t active = cj >= 32 ? 0xffffffffu : ((1u << cj) - 1u);
                const uint32_t bad = active & ~word;
                if (bad) atomicMin(first_row + tj, (long long)(chunk * 32 + (size_t)(__ffs(bad) - 1)));
            }
        }
        __syncwarp();
    }
    if (have && n_samples_out) n_samples_out[my] = emitted_total;
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

// Dynamic shared memory a kernel variant needs for this world.
size_t point_smem_bytes(const WorldView &wv, int bp, bool &aabb_in_smem)
{
    aabb_in_smem = false;
    if (bp == EPP_BP_BITS) return (size_t)3 * wv.nb * wv.W * 4;
    if (bp == EPP_BP_BRUTE && wv.M <= 2048) {
        aabb_in_smem = true;
        return (size_t)wv.M * sizeof(AabbRec);
    }
    return 0;
}

size_t ray_smem_bytes(const WorldView &wv, int bp, bool &aabb_in_smem)
{
    aabb_in_smem = false;
    if (bp == EPP_BP_BITS) return (size_t)6 * wv.nb * wv.W * 4;
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

unsigned persistent_grid(const LaunchCfg &cfg, size_t tiles, int blocks_per_sm)
{
    size_t cap = (size_t)cfg.sm_count * (size_t)blocks_per_sm;
    size_t g = tiles < cap ? tiles : cap;
    return (unsigned)(g ? g : 1);
}

} // namespace

std::atomic<uint64_t> g_launches{0};

#define EPP_DISPATCH_BP(bp, ...)                                                 \
    switch (bp) {                                                                \
    case EPP_BP_BRUTE: { constexpr int BP = EPP_BP_BRUTE; __VA_ARGS__; } break;  \
    case EPP_BP_BITS: { constexpr int BP = EPP_BP_BITS; __VA_ARGS__; } break;    \
    default: { constexpr int BP = EPP_BP_GRID; __VA_ARGS__; } break;             \
    }

template <int BP, int S>
cudaError_t launch_states_stream(const LaunchCfg &cfg, const WorldView &wv, const double *d_xyz, size_t n, int mode,
                                 int can_pass_gate, double min_distance, bool in_smem, size_t table_bytes, int stages,
                                 uint32_t *d_mask, cudaStream_t stream, bool &launched)
{
    constexpr int kTile = S * kBlock;
    const int qcap = cfg.states_queue_pairs > 0 ? cfg.states_queue_pairs : 2 * kTile;
    const size_t smem = table_bytes + (size_t)stages * kTile * 24 + (size_t)qcap * 4;
    launched = false;
    if (smem > 200 * 1024) return cudaSuccess;
    cudaError_t err = cudaFuncSetAttribute(k_states_stream<BP, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    const size_t tiles = (n + kTile - 1) / kTile;
    const unsigned grid = persistent_grid(cfg, tiles, cfg.states_stream_blocks_per_sm);
    k_states_stream<BP, S><<<grid, kBlock, smem, stream>>>(wv, d_xyz, n, mode, can_pass_gate, min_distance, in_smem ? 1 : 0,
                                                           stages, (uint32_t)table_bytes, qcap, d_mask);
    launched = true;
    return cudaGetLastError();
}

template <int BP, int S, bool RS>
cudaError_t launch_states_warp(const LaunchCfg &cfg, const WorldView &wv, const double *d_xyz, size_t n, int mode,
                               int can_pass_gate, double min_distance, bool in_smem, size_t table_bytes, int stages,
                               uint32_t *d_mask, cudaStream_t stream, bool &launched)
{
    const size_t rec_bytes = RS ? ((size_t)wv.M * kRecDoubles * 8 + 127) / 128 * 128 : 0;
    const size_t warp_bytes = (size_t)stages * S * kWordBytes + S * 32 + 32;
    const size_t smem = table_bytes + rec_bytes + kWarpsPerBlock * warp_bytes;
    launched = false;
    const size_t words = (n + 31) / 32;
    if (smem > 200 * 1024 || words >= 0xfff00000ull) return cudaSuccess;
    cudaError_t err = cudaFuncSetAttribute(k_states_warp<BP, S, RS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    // resident blocks per SM: the tuning value, limited by shared memory (228 KB per SM, 1 KB reserved per block)
    int bps = cfg.states_warp_blocks_per_sm > 0 ? cfg.states_warp_blocks_per_sm : 3;
    const int fit = (int)((228 * 1024) / (smem + 1024 + 768));
    if (bps > fit) bps = fit;
    if (bps < 1) bps = 1;
    const size_t warps_needed = (words + S - 1) / S;
    const size_t blocks_needed = (warps_needed + kWarpsPerBlock - 1) / kWarpsPerBlock;
    const unsigned grid = persistent_grid(cfg, blocks_needed, bps);
    const uint32_t warps = grid * kWarpsPerBlock;
    k_states_warp<BP, S, RS><<<grid, kBlock, smem, stream>>>(wv, d_xyz, n, mode | (cfg.states_debug << 8), can_pass_gate, min_distance,
                                                             in_smem ? 1 : 0, stages, (uint32_t)table_bytes,
                                                             (uint32_t)rec_bytes, (uint32_t)(words / warps),
                                                             (uint32_t)(words % warps), d_mask);
    launched = true;
    return cudaGetLastError();
}

cudaError_t launch_states(const LaunchCfg &cfg, const WorldView &wv, int bp, const double *d_xyz, size_t n, int mode,
                          int can_pass_gate, double min_distance, uint32_t *d_mask, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    bool in_smem;
    const size_t smem = point_smem_bytes(wv, bp, in_smem);
    cudaError_t err = cudaSuccess;
    // Warp-pipelined form: any alignment, any n; worth its prologue from a few thousand states up.
    if ((cfg.states_variant == 3 || (cfg.states_variant == 0 && n >= (size_t)cfg.states_warp_min_n)) &&
        wv.M < (1 << kPairUnitShift)) {
        const size_t table_bytes = (smem + 127) / 128 * 128;
        const int stages = cfg.states_stages < 1 ? 1 : (cfg.states_stages > kMaxStages ? kMaxStages : cfg.states_stages);
        const int S = cfg.states_per_thread == 4 ? 4 : 8;
        const bool rs = wv.M <= kRecSmemMax && wv.zrot && cfg.states_rec_smem != 0;
        const int wbp = (bp == EPP_BP_BITS && wv.W == 1 && cfg.states_bits1 != 0) ? kBpBits1 : bp;
        bool launched = false;
#define EPP_LAUNCH_WARP(S_, RS_)                                                                                       \
    err = launch_states_warp<BP, S_, RS_>(cfg, wv, d_xyz, n, mode, can_pass_gate, min_distance, in_smem, table_bytes, \
                                          stages, d_mask, stream, launched)
#define EPP_LAUNCH_WARP_S()                      \
    if (S == 4 && rs) EPP_LAUNCH_WARP(4, true);  \
    else if (S == 4) EPP_LAUNCH_WARP(4, false);  \
    else if (rs) EPP_LAUNCH_WARP(8, true);       \
    else EPP_LAUNCH_WARP(8, false)
        if (wbp == kBpBits1) {
            constexpr int BP = kBpBits1;
            EPP_LAUNCH_WARP_S();
        } else {
            EPP_DISPATCH_BP(bp, { EPP_LAUNCH_WARP_S(); });
        }
#undef EPP_LAUNCH_WARP_S
#undef EPP_LAUNCH_WARP
        if (err != cudaSuccess) return err;
        if (launched) {
            g_launches++;
            return cudaSuccess;
        }
    }
    // Streaming form: batches worth a ring, 16 B aligned (cp.async.bulk), obb index fits the pair encoding.
    const bool stream_ok = (reinterpret_cast<uintptr_t>(d_xyz) & 15u) == 0 && wv.M < (1 << kPairUnitShift);
    const bool want_stream = cfg.states_variant == 2 || (cfg.states_variant == 0 && n >= 8192);
    if (stream_ok && want_stream) {
        const size_t table_bytes = (smem + 127) / 128 * 128;
        const int stages = cfg.states_stages < 2 ? 2 : (cfg.states_stages > kMaxStages ? kMaxStages : cfg.states_stages);
        const int S = cfg.states_per_thread == 4 ? 4 : 2;
        bool launched = false;
        EPP_DISPATCH_BP(bp, {
            if (S == 4)
                err = launch_states_stream<BP, 4>(cfg, wv, d_xyz, n, mode, can_pass_gate, min_distance, in_smem, table_bytes,
                                                  stages, d_mask, stream, launched);
            else
                err = launch_states_stream<BP, 2>(cfg, wv, d_xyz, n, mode, can_pass_gate, min_distance, in_smem, table_bytes,
                                                  stages, d_mask, stream, launched);
        });
        if (err != cudaSuccess) return err;
        if (launched) {
            g_launches++;
            return cudaSuccess;
        }
    }
    const size_t tiles = (n + kBlock - 1) / kBlock;
    const unsigned grid = persistent_grid(cfg, tiles, cfg.states_blocks_per_sm);
    EPP_DISPATCH_BP(bp, {
        err = ensure_smem(k_states<BP>, smem);
        if (err == cudaSuccess) {
            k_states<BP><<<grid, kBlock, smem, stream>>>(wv, d_xyz, n, mode, can_pass_gate, min_distance, in_smem ? 1 : 0, d_mask);
            err = cudaGetLastError();
        }
    });
    g_launches++;
    return err;
}

cudaError_t launch_segments(const LaunchCfg &cfg, const WorldView &wv, int bp, const double *d_a, const double *d_b,
                            size_t n, int can_pass_gate, uint32_t *d_mask, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    bool in_smem;
    const size_t smem = ray_smem_bytes(wv, bp, in_smem);
    const size_t tiles = (n + kBlock - 1) / kBlock;
    const unsigned grid = persistent_grid(cfg, tiles, cfg.segments_blocks_per_sm);
    cudaError_t err = cudaSuccess;
    EPP_DISPATCH_BP(bp, {
        err = ensure_smem(k_segments<BP>, smem);
        if (err == cudaSuccess) {
            k_segments<BP><<<grid, kBlock, smem, stream>>>(wv, d_a, d_b, n, can_pass_gate, in_smem ? 1 : 0, d_mask);
            err = cudaGetLastError();
        }
    });
    g_launches++;
    return err;
}

cudaError_t launch_trajectories(const LaunchCfg &, const WorldView &wv, int bp, const double *d_mats, size_t n_traj,
                                size_t rows, size_t ld, size_t traj_stride, double min_distance, uint32_t *d_traj_valid,
                                uint32_t *d_sample_mask, int64_t *d_first_invalid, long long *d_first_scratch,
                                cudaStream_t stream)
{
    if (n_traj == 0) return cudaSuccess;
    k_fill_i64<<<(unsigned)((n_traj + 255) / 256), 256, 0, stream>>>((int64_t *)d_first_scratch, n_traj, LLONG_MAX);
    g_launches++;
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    if (rows > 0) {
        bool in_smem;
        const size_t smem = point_smem_bytes(wv, bp, in_smem);
        const size_t row_tiles = (rows + kBlock - 1) / kBlock;
        if (row_tiles > 65535 || n_traj > 0x7fffffffull) return cudaErrorInvalidValue;
        dim3 grid((unsigned)n_traj, (unsigned)row_tiles);
        EPP_DISPATCH_BP(bp, {
            err = ensure_smem(k_traj_samples<BP>, smem);
            if (err == cudaSuccess) {
                k_traj_samples<BP><<<grid, kBlock, smem, stream>>>(wv, d_mats, rows, ld, traj_stride, min_distance,
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

cudaError_t launch_poly_trajectories(const LaunchCfg &, const WorldView &wv, int bp, const double *d_coeffs,
                                     const double *d_durations, const double *d_dt, size_t n_traj, int n_segments,
                                     int n_coeffs, double min_distance, size_t max_samples, uint32_t *d_traj_valid,
                                     uint32_t *d_sample_mask, uint32_t *d_n_samples, int64_t *d_first_invalid,
                                     long long *d_first_scratch, cudaStream_t stream)
{
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
    const size_t per_block = (size_t)kPolyWarps * 32;
    const unsigned grid = (unsigned)((n_traj + per_block - 1) / per_block);
    EPP_DISPATCH_BP(bp, {
        err = ensure_smem(k_poly_traj<BP>, smem, sizeof(double) * kPolyWarps * 32 * kPolyPitch + 2 * kPolyWarps * 32 * kPolyPitch);
        if (err == cudaSuccess) {
            k_poly_traj<BP><<<grid, kPolyWarps * 32, smem, stream>>>(wv, d_coeffs, d_durations, d_dt, n_traj, n_segments,
                                                                     n_coeffs, min_distance, max_samples, in_smem ? 1 : 0,
                                                                     d_sample_mask, d_n_samples, d_first_scratch);
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
