// floor.cu — what is the fastest a 1M-state (24 MiB) pass can possibly be on this box?
// Times, from a CUDA graph of 16 launches over a 16-batch ring (> L2), per launch:
//   empty      : an empty kernel (launch floor inside a graph)
//   ldg        : plain 16 B loads, xor-reduced, 1 word written per warp (read floor, no smem)
//   tma        : per-warp cp.async.bulk of 3 KB chunks into smem, 1 LDS per lane per chunk, 1 word per warp-chunk
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o floor floor.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void k_empty() {}

__global__ void __launch_bounds__(256) k_ldg(const uint4 *__restrict__ p, size_t n16, uint32_t *out)
{
    uint32_t acc = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = __ldg(p + i);
        acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    for (int o = 16; o; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) out[((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5] = acc;
}

__device__ __forceinline__ uint32_t sa(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int CHUNK>
__global__ void __launch_bounds__(256) k_tma(const double *__restrict__ xyz, uint32_t chunks_per_warp, int stages, uint32_t *out)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[8][8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char *mine = smem + (size_t)warp * stages * CHUNK;
    const uint32_t g = blockIdx.x * 8 + warp;
    const size_t c0 = (size_t)g * chunks_per_warp;
    if (lane == 0) {
        for (int s = 0; s < stages; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sa(&bars[warp][s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int s = 0; s < stages && s < (int)chunks_per_warp; ++s) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sa(&bars[warp][s])), "r"(CHUNK) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sa(mine + s * CHUNK)),
                         "l"((const char *)xyz + (c0 + s) * CHUNK), "r"(CHUNK), "r"(sa(&bars[warp][s])) : "memory");
        }
    }
    __syncwarp();
    int slot = 0;
    uint32_t parity = 0, acc = 0;
    for (uint32_t c = 0; c < chunks_per_warp; ++c) {
        uint32_t ok = 0;
        while (!ok) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(sa(&bars[warp][slot])), "r"(parity) : "memory");
        }
        const uint32_t *w = reinterpret_cast<const uint32_t *>(mine + slot * CHUNK);
        for (int i = lane; i < CHUNK / 4; i += 32 * 8) acc ^= w[i];
        __syncwarp();
        if (lane == 0 && c + stages < chunks_per_warp) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sa(&bars[warp][slot])), "r"(CHUNK) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sa(mine + slot * CHUNK)),
                         "l"((const char *)xyz + (c0 + c + stages) * CHUNK), "r"(CHUNK), "r"(sa(&bars[warp][slot])) : "memory");
        }
        if (++slot == stages) { slot = 0; parity ^= 1; }
    }
    for (int o = 16; o; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) out[g] = acc;
}

template <class F>
float time_graph(cudaStream_t st, int ring, int reps, F launch)
{
    cudaGraph_t g;
    cudaGraphExec_t ge;
    cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal);
    for (int r = 0; r < ring; ++r) launch(r);
    cudaStreamEndCapture(st, &g);
    cudaGraphInstantiate(&ge, g, 0);
    cudaGraphLaunch(ge, st);
    cudaStreamSynchronize(st);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a, st);
    for (int i = 0; i < reps; ++i) cudaGraphLaunch(ge, st);
    cudaEventRecord(b, st);
    cudaStreamSynchronize(st);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("error: %s\n", cudaGetErrorString(e));
    return ms * 1e3f / (reps * ring);
}

int main(int argc, char **argv)
{
    const size_t n = argc > 1 ? (size_t)atol(argv[1]) : (1u << 20);
    const size_t bytes = n * 24;
    const int ring = 16;
    std::vector<double *> bufs(ring);
    for (int r = 0; r < ring; ++r) {
        CK(cudaMalloc(&bufs[r], bytes));
        CK(cudaMemset(bufs[r], r + 1, bytes));
    }
    uint32_t *out;
    CK(cudaMalloc(&out, 1 << 22));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    const double ideal = bytes / 6539.9e9 * 1e6;
    printf("n %zu states, %.1f MiB per launch, ideal at 6539.9 GB/s = %.2f us\n", n, bytes / 1048576.0, ideal);
    printf("empty                       %7.2f us\n", time_graph(st, ring, 200, [&](int) { k_empty<<<1, 32, 0, st>>>(); }));
    printf("empty 592x256               %7.2f us\n", time_graph(st, ring, 200, [&](int) { k_empty<<<592, 256, 0, st>>>(); }));
    for (int bps : {2, 4, 8}) {
        float us = time_graph(st, ring, 100, [&](int r) { k_ldg<<<148 * bps, 256, 0, st>>>((const uint4 *)bufs[r], bytes / 16, out); });
        printf("ldg.128 persistent %d blk/SM %7.2f us  frac %.3f\n", bps, us, ideal / us);
    }
    {
        const unsigned grid = (unsigned)((bytes / 16 + 255) / 256);
        float us = time_graph(st, ring, 100, [&](int r) { k_ldg<<<grid, 256, 0, st>>>((const uint4 *)bufs[r], bytes / 16, out); });
        printf("ldg.128 one elem/thread     %7.2f us  frac %.3f\n", us, ideal / us);
    }
    for (int bps : {2, 3, 4}) {
        for (int stages : {1, 2, 4}) {
            constexpr int CH = 3072;
            const uint32_t chunks = (uint32_t)(bytes / CH);
            const uint32_t warps = 148 * bps * 8;
            const uint32_t cpw = chunks / warps; // drops the remainder: floor measurement only
            const size_t smem = (size_t)8 * stages * CH;
            if (smem * bps > 200 * 1024) continue;
            cudaFuncSetAttribute(k_tma<CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            float us = time_graph(st, ring, 100, [&](int r) { k_tma<CH><<<148 * bps, 256, smem, st>>>(bufs[r], cpw, stages, out); });
            printf("tma 3KB/warp-chunk bps %d stages %d cpw %u  %7.2f us  frac %.3f\n", bps, stages, cpw, us, ideal / us * (double)cpw * warps / chunks);
        }
    }
    return 0;
}
