// Microbenchmark: how fast can warps read TMEM (tcgen05.ld) WHILE the tensor core runs tf32 MMAs,
// and how fast do the MMAs run with A from shared memory vs A from TMEM, N = 128 vs 256.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../c3dgs_b200/csrc -o tmem_mma_bench tmem_mma_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"

using namespace c3vq::ptx;

constexpr int DP = 48, KSTEPS = DP / 8;

__device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ float min4(uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    return fminf(fminf(fminf(__uint_as_float(a), __uint_as_float(b)), __uint_as_float(c)), __uint_as_float(d));
}

// LDMODE 0: x16 double buffered (the kernel's drain), 1: x32 + wait, 2: x16 + wait (no overlap)
template <int N, bool A_TMEM, int LDMODE>
__global__ void __launch_bounds__(640, 1) k(int n_jobs, int do_mma, int drain_iters, unsigned long long *out, float *sink)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *sA = smem;                         // 128 x 48 tf32, canonical K-major (LBO 2048)
    uint8_t *sB = smem + 128 * DP * 4;          // N x 48
    __shared__ uint32_t tptr;
    __shared__ uint64_t bar;
    __shared__ volatile int done;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (128 + N) * DP; i += blockDim.x) reinterpret_cast<float *>(smem)[i] = 0.f;
    if (threadIdx.x == 0) {
        mbar_init(smem_u32(&bar), 1);
        fence_mbar_init();
        done = 0;
    }
    if (warp == 2) {
        tmem_alloc(smem_u32(&tptr), 512);
        tmem_relinquish();
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tptr;
    constexpr int NSLOT = A_TMEM ? (512 - 64) / N : 512 / N;
    const uint32_t a_tmem = tb + NSLOT * N;     // 48 columns of A when A_TMEM

    if (warp == 1 && lane == 0) {
        if (do_mma) {
            constexpr uint32_t IDESC = idesc_tf32(128, N);
            const unsigned long long t0 = clock64();
            for (int j = 0; j < n_jobs; j++) {
                const uint32_t d = tb + (j % NSLOT) * N;
#pragma unroll
                for (int ks = 0; ks < KSTEPS; ks++) {
                    const uint64_t bd = smem_desc(smem_u32(sB) + ks * 2 * (N * 16), N * 16, 128);
                    if (A_TMEM) {
                        umma_tf32_ts(d, a_tmem + ks * 8, bd, IDESC, ks > 0);
                    } else {
                        const uint64_t ad = smem_desc(smem_u32(sA) + ks * 2 * 2048, 2048, 128);
                        umma_tf32(d, ad, bd, IDESC, ks > 0);
                    }
                }
            }
            umma_commit(smem_u32(&bar));
            mbar_wait(smem_u32(&bar), 0);
            const unsigned long long t1 = clock64();
            out[blockIdx.x * 32 + 0] = t1 - t0;
        }
        done = 1;
    } else if (warp >= 4) {
        const uint32_t base = tb + ((uint32_t)((warp & 3) * 32) << 16);
        float acc = 0.f;
        unsigned long long iters = 0;
        const unsigned long long t0 = clock64();
        for (int it = 0;; it++) {
            if (do_mma ? (done != 0) : (it >= drain_iters)) break;
            const uint32_t col = (it * 32 + (warp >> 2) * 64) & 255;
            if (LDMODE == 0) {
                uint32_t va[16], vb[16];
                tmem_ld16_async(base + col, va);
                tmem_wait_ld16(va);
                tmem_ld16_async(base + col + 16, vb);
                acc += fminf(fminf(min4(va[0], va[1], va[2], va[3]), min4(va[4], va[5], va[6], va[7])),
                             fminf(min4(va[8], va[9], va[10], va[11]), min4(va[12], va[13], va[14], va[15])));
                tmem_wait_ld16(vb);
                acc += fminf(fminf(min4(vb[0], vb[1], vb[2], vb[3]), min4(vb[4], vb[5], vb[6], vb[7])),
                             fminf(min4(vb[8], vb[9], vb[10], vb[11]), min4(vb[12], vb[13], vb[14], vb[15])));
            } else if (LDMODE == 1) {
                float v[32];
                tmem_ld32(base + col, v);
                float m = v[0];
#pragma unroll
                for (int i = 1; i < 32; i += 2) m = fminf(fminf(m, v[i]), v[(i + 1) & 31]);
                acc += m;
            } else {
                uint32_t va[16];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    tmem_ld16_async(base + col + 16 * h, va);
                    tmem_wait_ld16(va);
                    acc += fminf(fminf(min4(va[0], va[1], va[2], va[3]), min4(va[4], va[5], va[6], va[7])),
                                 fminf(min4(va[8], va[9], va[10], va[11]), min4(va[12], va[13], va[14], va[15])));
                }
            }
            iters++;
        }
        const unsigned long long t1 = clock64();
        if (lane == 0) {
            out[blockIdx.x * 32 + warp] = t1 - t0;
            out[148 * 32 + blockIdx.x * 32 + warp] = iters;
        }
        sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tb, 512);
    }
}

template <int N, bool A_TMEM, int LDMODE>
void run(int drain_warps, int do_mma, const char *name)
{
    unsigned long long *out;
    float *sink;
    cudaMalloc(&out, 2 * 148 * 32 * 8);
    cudaMemset(out, 0, 2 * 148 * 32 * 8);
    cudaMalloc(&sink, 148 * 1024 * 4);
    const int threads = (4 + drain_warps) * 32;
    const size_t smem = (128 + N) * DP * 4;
    cudaFuncSetAttribute(k<N, A_TMEM, LDMODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int n_jobs = 20000;                  // x 6 MMAs
    for (int rep = 0; rep < 2; rep++) k<N, A_TMEM, LDMODE><<<148, threads, smem>>>(n_jobs, do_mma, 20000, out, sink);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[2 * 148 * 32];
    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    double mma_cyc = do_mma ? (double)h[0] / (n_jobs * 6.0) : 0.0;
    double bytes = 0, cyc = 1;
    for (int w = 4; w < 4 + drain_warps; w++) {
        bytes += (double)h[148 * 32 + w] * 4096.0;
        cyc = (double)h[w];
    }
    printf("%-44s N=%3d A=%s drain_warps=%2d mma=%d : %7.1f cyc/MMA (floor %3d)   drain %8.1f B/cyc/SM (%s)\n", name, N,
           A_TMEM ? "tmem" : "smem", drain_warps, do_mma, mma_cyc, N / 2, drain_warps ? bytes / cyc : 0.0,
           cudaGetErrorString(e));
    cudaFree(out);
    cudaFree(sink);
}

int main()
{
    // MMA alone
    run<128, false, 0>(0, 1, "mma only");
    run<256, false, 0>(0, 1, "mma only");
    run<128, true, 0>(0, 1, "mma only");
    run<256, true, 0>(0, 1, "mma only");   // NSLOT = 1
    // drain alone
    for (int w : {4, 8, 16}) run<128, false, 0>(w, 0, "drain only, x16 double buffered");
    for (int w : {8, 16}) run<128, false, 1>(w, 0, "drain only, x32 + wait");
    for (int w : {8, 16}) run<128, false, 2>(w, 0, "drain only, x16 + wait");
    // both
    for (int w : {4, 8, 16}) run<128, false, 0>(w, 1, "both, x16 double buffered");
    for (int w : {8, 16}) run<256, false, 0>(w, 1, "both, x16 double buffered");
    for (int w : {8, 16}) run<128, true, 0>(w, 1, "both, x16 double buffered");
    for (int w : {8, 16}) run<128, false, 1>(w, 1, "both, x32 + wait");
    for (int w : {8, 16}) run<128, true, 1>(w, 1, "both, x32 + wait");
    return 0;
}
