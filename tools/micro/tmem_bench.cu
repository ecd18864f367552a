// Microbenchmark: TMEM -> register read throughput (tcgen05.ld) per SM, for different shapes / warp counts.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_bench tmem_bench.cu && ./tmem_bench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int NREG>
__device__ __forceinline__ void ld(uint32_t taddr, uint32_t (&r)[128]);

#define R4(i) "%" #i
__device__ __forceinline__ void ld32(uint32_t taddr, uint32_t *r)
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void ld16(uint32_t taddr, uint32_t *r)
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// mode 0: x32 + wait each; mode 1: two x32 in flight then wait; mode 2: x16 + wait each; mode 3: x32 + wait + min32 work
template <int MODE>
__global__ void k(int iters, unsigned long long *out, float *sink)
{
    __shared__ uint32_t tptr;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tptr)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t base = tptr + ((uint32_t)((warp & 3) * 32) << 16);
    float acc = 0.f;
    const unsigned long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        const uint32_t col = (it * 32 + (warp >> 2) * 64) & 255;
        if (MODE == 0 || MODE == 3) {
            uint32_t r[32];
            ld32(base + col, r);
            wait_ld();
            if (MODE == 3) {
                float m = __uint_as_float(r[0]);
#pragma unroll
                for (int i = 1; i < 32; i += 2) m = fminf(fminf(m, __uint_as_float(r[i])), __uint_as_float(r[(i + 1) & 31]));
                acc += m;
            } else {
                acc += __uint_as_float(r[it & 31]);
            }
        } else if (MODE == 1) {
            uint32_t r[32], q[32];
            ld32(base + col, r);
            ld32(base + ((col + 32) & 255), q);
            wait_ld();
            acc += __uint_as_float(r[it & 31]) + __uint_as_float(q[it & 31]);
        } else {
            uint32_t r[16];
            ld16(base + col, r);
            wait_ld();
            acc += __uint_as_float(r[it & 15]);
        }
    }
    const unsigned long long t1 = clock64();
    if (lane == 0) out[blockIdx.x * 32 + warp] = t1 - t0;
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tptr), "r"(512));
}

template <int MODE>
void run(int warps, const char *name, int bytes_per_iter)
{
    unsigned long long *out;
    float *sink;
    cudaMalloc(&out, 148 * 32 * 8);
    cudaMalloc(&sink, 148 * 1024 * 4);
    const int iters = 4096;
    k<MODE><<<148, warps * 32, 0>>>(iters, out, sink);
    k<MODE><<<148, warps * 32, 0>>>(iters, out, sink);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[32];
    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    double cyc = (double)h[0] / iters;
    printf("%-34s warps=%2d  %7.1f cyc/iter/warp  -> %7.1f B/cyc/SM  (%s)\n", name, warps, cyc,
           (double)bytes_per_iter * warps / cyc, cudaGetErrorString(e));
    cudaFree(out);
    cudaFree(sink);
}

int main()
{
    for (int w : {1, 4, 8, 16}) run<0>(w, "32x32b.x32 + wait", 4096);
    for (int w : {4, 8, 16}) run<1>(w, "2 x (32x32b.x32) + wait", 8192);
    for (int w : {4, 8}) run<2>(w, "32x32b.x16 + wait", 2048);
    for (int w : {4, 8, 16}) run<3>(w, "32x32b.x32 + wait + min32", 4096);
    return 0;
}
