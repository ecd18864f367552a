// assign_exact.cu — SIMT in-order fp32 nearest-codeword search.
//
// Replaces the reference kernel distance_reduce / distance
// (submodules/weighted_distance/weighted_distance.cu:9-44) with the same arithmetic —
// acc = fma(x_c - cb_c, x_c - cb_c, acc) for c = 0..d-1, ascending codeword scan,
// strict `<` — so results are bit-identical by construction.  It is the parity
// anchor for the tensor-core path and its fallback for rows whose candidate list
// overflows; it is NOT the fast path.
//
// Layout: each thread owns RPT feature rows in registers (zero padded to DP floats;
// fma(0,0,acc) == acc keeps the chain bit-identical), a block of 128 threads streams
// the codebook through a 64-codeword shared-memory tile read with broadcast LDS.128.
// Algorithmic work: 2*d fp32 ops per (row, codeword); CUDA-core bound.
#include "common.cuh"

namespace c3vq {

constexpr int EX_THREADS = 128;
constexpr int EX_TILE = 64;  // codewords per shared-memory tile

template <int DP, int RPT>
__global__ void __launch_bounds__(EX_THREADS)
k_assign_exact(const float *__restrict__ x, const int64_t *__restrict__ rows, int64_t n, int d,
               const float *__restrict__ cb, int k, float *__restrict__ out_dist, int64_t *__restrict__ out_idx)
{
    __shared__ __align__(16) float cs[EX_TILE * DP];
    const int tid = threadIdx.x;
    const int64_t base = (int64_t)blockIdx.x * (EX_THREADS * RPT);

    float xr[RPT][DP];
#pragma unroll
    for (int r = 0; r < RPT; r++) {
        const int64_t i = base + tid + r * EX_THREADS;
        const bool valid = i < n;
        const int64_t src = valid ? (rows ? rows[i] : i) : 0;
        const float *xp = x + src * d;
#pragma unroll
        for (int c = 0; c < DP; c++) xr[r][c] = (valid && c < d) ? xp[c] : 0.f;
    }
    for (int e = tid; e < EX_TILE * DP; e += EX_THREADS) cs[e] = 0.f;  // padding columns stay zero

    float best[RPT];
    int best_j[RPT];
#pragma unroll
    for (int r = 0; r < RPT; r++) { best[r] = FLT_MAX; best_j[r] = 0; }

    for (int j0 = 0; j0 < k; j0 += EX_TILE) {
        const int cnt = min(EX_TILE, k - j0);
        __syncthreads();
        const float *src = cb + (size_t)j0 * d;
        for (int e = tid; e < cnt * d; e += EX_THREADS) {
            const int jj = e / d;
            cs[jj * DP + (e - jj * d)] = src[e];
        }
        __syncthreads();
        for (int jj = 0; jj < cnt; jj++) {
            float acc[RPT];
#pragma unroll
            for (int r = 0; r < RPT; r++) acc[r] = 0.f;
            const float4 *cp = reinterpret_cast<const float4 *>(cs + jj * DP);
#pragma unroll
            for (int c4 = 0; c4 < DP / 4; c4++) {
                const float4 cv = cp[c4];
#pragma unroll
                for (int r = 0; r < RPT; r++) {
                    float df;
                    df = xr[r][4 * c4 + 0] - cv.x; acc[r] = fmaf(df, df, acc[r]);
                    df = xr[r][4 * c4 + 1] - cv.y; acc[r] = fmaf(df, df, acc[r]);
                    df = xr[r][4 * c4 + 2] - cv.z; acc[r] = fmaf(df, df, acc[r]);
                    df = xr[r][4 * c4 + 3] - cv.w; acc[r] = fmaf(df, df, acc[r]);
                }
            }
#pragma unroll
            for (int r = 0; r < RPT; r++) {
                if (acc[r] < best[r]) { best[r] = acc[r]; best_j[r] = j0 + jj; }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < RPT; r++) {
        const int64_t i = base + tid + r * EX_THREADS;
        if (i < n) {
            if (out_dist) out_dist[i] = best[r];
            if (out_idx) out_idx[i] = best_j[r];
        }
    }
}

template <int DP, int RPT>
static int launch_t(const float *x, const int64_t *rows, int64_t n, int d, const float *cb, int k, float *out_dist,
                    int64_t *out_idx, cudaStream_t s)
{
    const int64_t per_block = EX_THREADS * RPT;
    const int64_t blocks = (n + per_block - 1) / per_block;
    k_assign_exact<DP, RPT><<<(unsigned)blocks, EX_THREADS, 0, s>>>(x, rows, n, d, cb, k, out_dist, out_idx);
    C3VQ_LAUNCH_CHECK("k_assign_exact");
    return C3VQ_OK;
}

int launch_assign_exact(const float *x, const int64_t *rows, int64_t n, int d, const float *cb, int k,
                        float *out_dist, int64_t *out_idx, cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    C3VQ_CHECK_ARG(n / (EX_THREADS) < 0x7fffffffLL, "n=%lld too large", (long long)n);
    switch (exact_dp(d)) {
        case 8: return launch_t<8, 4>(x, rows, n, d, cb, k, out_dist, out_idx, s);
        case 16: return launch_t<16, 4>(x, rows, n, d, cb, k, out_dist, out_idx, s);
        case 32: return launch_t<32, 2>(x, rows, n, d, cb, k, out_dist, out_idx, s);
        case 48: return launch_t<48, 2>(x, rows, n, d, cb, k, out_dist, out_idx, s);
        default: return launch_t<64, 1>(x, rows, n, d, cb, k, out_dist, out_idx, s);
    }
}

}  // namespace c3vq
