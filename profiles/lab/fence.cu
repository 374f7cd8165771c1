// fence.cu — cycles of the barrier-init prologue pieces (one warp's lane 0, 1024-thread block, 148 blocks)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ uint32_t sa(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void k(long long *out, int variant)
{
    __shared__ __align__(8) uint64_t bars[64];
    __shared__ __align__(128) unsigned char buf[4096];
    long long t0 = clock64();
    if (threadIdx.x < 64) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sa(&bars[threadIdx.x])));
    long long t1 = clock64();
    if (threadIdx.x < 64) {
        if (variant & 1) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    long long t2 = clock64();
    if (threadIdx.x < 64) {
        if (variant & 2) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    long long t3 = clock64();
    __syncthreads();
    long long t4 = clock64();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sa(&bars[0])), "r"(3072) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sa(buf)),
                     "l"(out + 1024), "r"(3072), "r"(sa(&bars[0])) : "memory");
    }
    long long t5 = clock64();
    uint32_t ok = 0;
    while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(sa(&bars[0])), "r"(0) : "memory");
    long long t6 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        out[0] = t1 - t0; out[1] = t2 - t1; out[2] = t3 - t2; out[3] = t4 - t3; out[4] = t5 - t4; out[5] = t6 - t5;
    }
}
int main()
{
    long long *d, h[6];
    cudaMalloc(&d, 1 << 20);
    cudaMemset(d, 0, 1 << 20);
    for (int v = 0; v < 4; ++v)
        for (int rep = 0; rep < 3; ++rep) {
            k<<<148, 1024>>>(d, v);
            cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
            printf("variant %d (mbarrier_init fence %d, proxy fence %d): init %lld  fence1 %lld  fence2 %lld  syncthreads %lld  issue %lld  tma 3KB L2 round trip %lld cycles  [%s]\n",
                   v, v & 1, (v >> 1) & 1, h[0], h[1], h[2], h[3], h[4], h[5], cudaGetErrorString(cudaGetLastError()));
        }
    return 0;
}
