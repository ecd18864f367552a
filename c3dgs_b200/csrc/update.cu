// update.cu — sensitivity-weighted centroid sums and the EMA epilogue.
//
// Replaces the two torch_scatter.scatter calls and the materialised
// `x * importance[:, None]` temporary of VectorQuantize.update
// (compression/vq.py:31, :33), and ema_inplace + divide + trace normalise
// (compression/vq.py:32, :34, :45-46, :76-77).
//
// Accumulator layout in HBM/L2: acc[k][stride] fp32, stride = acc_stride(d)
// (16-byte aligned rows): columns 0..d-1 = sum w*x, column d = sum w.  At K=4096,
// d=45 the table is 768 KB and stays L2 resident; rows of x are streamed exactly once
// (algorithmic bytes per row: 4*d + 4 (w) + 8 (idx)).  Each row contributes one
// vectorised red.global.add.v4.f32 per 16-byte accumulator chunk when the row can be
// loaded as float4 (d % 4 == 0), else one scalar red per element issued by consecutive
// lanes onto consecutive addresses.
#include "common.cuh"

namespace c3vq {

__device__ __forceinline__ void red_add_v4(float *addr, float a, float b, float c, float d)
{
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

// Generic path: LPR lanes cooperate on one row (LPR = 8, 16 or 32), 32/LPR rows per warp pass.
// Element c < d carries w*x_c (product rounded to fp32 first, like the reference's temporary),
// element c == wcol carries w.
template <int LPR>
__global__ void __launch_bounds__(256)
k_accumulate(const float *__restrict__ x, const float *__restrict__ w, const int64_t *__restrict__ rows,
             const int64_t *__restrict__ idx, int64_t n, int d, int64_t k, float *__restrict__ acc, int stride,
             int wcol)
{
    constexpr int RPW = 32 / LPR;  // rows per warp pass
    const int lane = threadIdx.x & 31;
    const int sub = lane / LPR;    // which row of the pass
    const int c0 = lane % LPR;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int ncol = (wcol >= 0) ? d + 1 : d;
    for (int64_t i0 = warp * RPW; i0 < n; i0 += nwarps * RPW) {
        const int64_t i = i0 + sub;
        if (i >= n) continue;
        const int64_t j = idx[i];
        if ((uint64_t)j >= (uint64_t)k) continue;  // out-of-range code: ignore the row
        const int64_t src = rows ? rows[i] : i;
        const float wi = w ? w[src] : 1.f;
        const float *xp = x + src * d;
        float *ap = acc + j * stride;
        for (int c = c0; c < ncol; c += LPR) {
            const float v = (c < d) ? __fmul_rn(xp[c], wi) : wi;
            atomicAdd(ap + ((c < d) ? c : wcol), v);
        }
    }
}

// float4 path (d % 4 == 0, 16-byte aligned x): one lane per 16-byte chunk, vector reductions.
__global__ void __launch_bounds__(256)
k_accumulate_v4(const float *__restrict__ x, const float *__restrict__ w, const int64_t *__restrict__ rows,
                const int64_t *__restrict__ idx, int64_t n, int d, int64_t k, float *__restrict__ acc, int stride,
                int wcol)
{
    const int d4 = d >> 2;
    const int cpr = d4 + (wcol >= 0 ? 1 : 0);  // chunks per row (last one = weight chunk)
    const int64_t total = n * cpr;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = e / cpr;
        const int c = (int)(e - i * cpr);
        const int64_t j = idx[i];
        if ((uint64_t)j >= (uint64_t)k) continue;
        const int64_t src = rows ? rows[i] : i;
        const float wi = w ? w[src] : 1.f;
        float *ap = acc + j * stride;
        if (c < d4) {
            const float4 v = *reinterpret_cast<const float4 *>(x + src * d + 4 * c);
            red_add_v4(ap + 4 * c, __fmul_rn(v.x, wi), __fmul_rn(v.y, wi), __fmul_rn(v.z, wi), __fmul_rn(v.w, wi));
        } else {
            atomicAdd(ap + wcol, wi);
        }
    }
}

int launch_accumulate(const float *x, const float *w, const int64_t *rows, const int64_t *idx, int64_t n, int d,
                      int64_t k, float *acc, int stride, int wcol, cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    const int sms = sm_count();
    const bool v4 = (d % 4 == 0) && (stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) &&
                    ((reinterpret_cast<uintptr_t>(acc) & 15) == 0);
    if (v4) {
        const int64_t total = n * (d / 4 + (wcol >= 0 ? 1 : 0));
        int64_t blocks = (total + 255) / 256;
        if (blocks > (int64_t)sms * 32) blocks = (int64_t)sms * 32;
        k_accumulate_v4<<<(unsigned)blocks, 256, 0, s>>>(x, w, rows, idx, n, d, k, acc, stride, wcol);
    } else {
        const int ncol = d + (wcol >= 0 ? 1 : 0);
        const int lpr = ncol <= 8 ? 8 : ncol <= 16 ? 16 : 32;
        const int64_t warps_needed = (n + (32 / lpr) - 1) / (32 / lpr);
        int64_t blocks = (warps_needed + 7) / 8;
        if (blocks > (int64_t)sms * 16) blocks = (int64_t)sms * 16;
        if (lpr == 8)
            k_accumulate<8><<<(unsigned)blocks, 256, 0, s>>>(x, w, rows, idx, n, d, k, acc, stride, wcol);
        else if (lpr == 16)
            k_accumulate<16><<<(unsigned)blocks, 256, 0, s>>>(x, w, rows, idx, n, d, k, acc, stride, wcol);
        else
            k_accumulate<32><<<(unsigned)blocks, 256, 0, s>>>(x, w, rows, idx, n, d, k, acc, stride, wcol);
    }
    C3VQ_LAUNCH_CHECK("k_accumulate");
    return C3VQ_OK;
}

// One warp per codeword, lanes over columns (coalesced); K is a few thousand, so this launch is latency only.
// Arithmetic mirrors torch: mul_(decay) rounds, add_(new, alpha) is fma(alpha, new, self).
__global__ void __launch_bounds__(256)
k_ema(float *__restrict__ cb, float *__restrict__ ei, const float *__restrict__ acc, int k, int d, int stride,
      float decay, float alpha, float eps, int scale_normalize)
{
    const int lane = threadIdx.x & 31;
    const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (j >= k) return;
    const float *a = acc + (size_t)j * stride;
    const float sw = a[d];
    if (lane == 0) ei[j] = fmaf(alpha, sw, __fmul_rn(ei[j], decay));   // vq.py:32
    const float denom = __fadd_rn(sw, eps);                             // vq.py:34
    float *row = cb + (size_t)j * d;
    float v[2];
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int c = lane + 32 * h;
        v[h] = 0.f;
        if (c < d) v[h] = fmaf(alpha, __fdiv_rn(a[c], denom), __fmul_rn(row[c], decay));
    }
    if (scale_normalize && d >= 6) {                                    // vq.py:76-77
        const float r0 = __shfl_sync(0xffffffffu, v[0], 0);
        const float r3 = __shfl_sync(0xffffffffu, v[0], 3);
        const float r5 = __shfl_sync(0xffffffffu, v[0], 5);
        const float tr = __fadd_rn(__fadd_rn(r0, r3), r5);
        v[0] = __fdiv_rn(v[0], tr);
        v[1] = __fdiv_rn(v[1], tr);
    }
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int c = lane + 32 * h;
        if (c < d) row[c] = v[h];
    }
}

int launch_ema(float *codebook, float *entry_importance, const float *acc, int k, int d, double decay, double eps,
               int scale_normalize, cudaStream_t s)
{
    if (k == 0) return C3VQ_OK;
    k_ema<<<(k + 7) / 8, 256, 0, s>>>(codebook, entry_importance, acc, k, d, acc_stride(d), (float)decay,
                                          (float)(1.0 - decay), (float)eps, scale_normalize);
    C3VQ_LAUNCH_CHECK("k_ema");
    return C3VQ_OK;
}

}  // namespace c3vq
