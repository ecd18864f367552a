// assign_tc.cu — tensor-core nearest-codeword search (+ fused weighted centroid sums) for sm_100a.
//
// Replaces distance_reduce (submodules/weighted_distance/weighted_distance.cu:20-44) and, when `acc`
// is given, the two scatter sums of VectorQuantize.update (compression/vq.py:31, :33), reading every
// feature row from HBM exactly once per iteration.
//
// Method (results are bit-identical to the reference's fp32 argmin):
//   1. SCREEN  s_ij = |c'_j|^2 - 2 x'_i . c'_j  with tcgen05.mma into TMEM (kind::f16 on operands scaled by a power
//      of two, or kind::tf32), where x' = x - mu, c' = c - mu (mu = codebook mean; a distance-preserving shift that
//      shrinks the cancellation error).  Operands are rounded by us, |c'|^2 rides along as three extra K columns
//      (hi/mid/lo pieces against a constant in the A operand), so the epilogue is a pure min: one FMNMX3 per three scores.
//   2. BAND    every GROUP of 3 consecutive codes (16 at the one-K-step widths) whose minimum score is within band_i of
//      the running row minimum is recorded, where band_i = 2*E_i + slack bounds the worst-case screening error
//      (input rounding 2^-11 per operand, tensor-core accumulation, norm split) plus the reference chain's own
//      rounding; the fp32 argmin of the reference is therefore always inside a recorded group.  The
//      running minimum starts from the score of the row's winner in the previous call on the same workspace,
//      re-evaluated against the current codebook (an upper bound of the minimum whatever the workspace holds).
//   3. RE-RANK all codes of the surviving groups with the reference arithmetic (in-order fma chain, first index wins
//      ties).  Rows whose group list overflows, or whose magnitudes break the error model, go to an exact scan
//      (k_assign_fallback_scan / _finish).
//
// Kernel organisation (persistent, one CTA per SM, 896 threads, warp specialised, setmaxnreg per role):
//   warps 0-15  drain: three tcgen05.ld of 16 columns in flight, running min + group list per row; two threads per
//               row (one per column half of every accumulator slot; one per slot at the narrow widths); at the end of a
//               pass the two exchange minima and push the surviving groups into a work queue
//   warps 16-23 feature producers AND re-rankers: stage the prefetched next 256 raw rows in shared memory
//               (double buffered), write them centred + rounded into TMEM as the A operand together with the row's
//               band and seed; then, while the tensor core and the drain warps work on that pass, re-rank the previous
//               pass' queue from the staged rows (a lane per code, shared-memory 64-bit atomicMin per row), write
//               idx/dist, scale the rows by their weights and bulk-reduce them into the centroid sums; then prefetch.
//   warp 24     bulk-copy (TMA engine, cp.async.bulk) producer: codebook image stages -> smem ring
//   warps 25,26 MMA issuers, one per row tile: per stage KSTEPS tcgen05.mma (M=128, N=BN), commits to mbarriers;
//               warp 26 also allocates the TMEM (512 columns: four 128 x BN fp32 accumulator slots, two per row
//               tile, then the A operand of both row tiles)
// A pass is 256 rows x all K codes.  Each row tile owns two accumulator slots, so the MMA warp fills one while
// the tile's 256 drain threads read the other: MMA, drain and re-rank all overlap.
#include <cuda_fp16.h>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace c3vq {

using namespace ptx;

#ifdef C3VQ_TC_PROF
#define PROF_DECL unsigned long long prof_t = clock64(); unsigned long long prof_a[8] = {0, 0, 0, 0, 0, 0, 0, 0}
#define PROF(i) do { const unsigned long long _n = clock64(); prof_a[i] += _n - prof_t; prof_t = _n; } while (0)
#define PROF_FLUSH(base, cond) do { if (cond) { for (int _i = 0; _i < 8; _i++) atomicAdd(&p.hdr->prof[(base) + _i], prof_a[_i]); } } while (0)
#else
#define PROF_DECL
#define PROF(i)
#define PROF_FLUSH(base, cond)
#endif

constexpr int TC_BM = 128;            // rows per UMMA tile
constexpr int TC_PR = 2 * TC_BM;      // rows per CTA pass ("pair" of tiles)
// codes per stage / per MMA: four accumulator slots plus the A operand of both row tiles share the 512 TMEM columns
// (64 codes per MMA with three accumulator slots per row tile was an experiment of the TF32 era: 4.19 ms per launch
// against 3.87 ms, the extra barrier round trips cost more than the deeper ring gained)
__host__ __device__ constexpr int tc_bn(int dp) { return dp <= 64 ? 96 : 64; }
__host__ __device__ constexpr int tc_spt(int dp) { return 2; }   // accumulator slots per row tile
// fp16 screening: the A operand takes half the TMEM columns, which would leave room for 112 codes per MMA (4 x 112 +
// 2 x DP/2 <= 512 up to DP = 64): 37 steps per pass over 4096 codewords instead of 43.  Built, bit-exact (full GPU
// suite green) and measured in round 2 (5.8 M rows, fused sums, mean of iterations 1-7): 3.61 / 3.78 / 3.19 ms at
// D = 45 / 48 / 7 against 3.43 / 3.70 / 3.15 ms with 96 -- the 56 columns a drain thread then holds per step spill
// (72 registers at 896 threads), which costs more than the saved barrier waits.  Kept behind C3VQ_TC_F16_BN112.
#ifdef C3VQ_TC_F16_BN112
__host__ __device__ constexpr int tc_bn_f16(int dp) { return dp <= 64 ? 112 : 96; }
#else
__host__ __device__ constexpr int tc_bn_f16(int dp) { return 96; }
#endif
// Candidate granularity: the drain records GROUPS of GS consecutive codes (one mask bit each) and the re-rank
// evaluates every code of a recorded group exactly.  Larger groups mean fewer minima / mask instructions per score in
// the drain (which bounds the kernel) and proportionally more exact distances in the re-rank (which has slack).
// Groups of 3 (the default for the wide rows): a group minimum is ONE three-input FMNMX3, so the 48 scores of a half job
// cost 16 + 8 minimum instructions instead of 24 + 6 with groups of 4 -- the drain is bound by the ALU pipe these run on
// -- and a candidate group costs three exact distances instead of four.
#ifndef C3VQ_TC_GS
#define C3VQ_TC_GS 3
#endif
#ifndef C3VQ_TC_GS_LOW
#define C3VQ_TC_GS_LOW 16
#endif
// (groups of 3 need a half job of 48 codes: BN = 96)
__host__ __device__ constexpr int tc_gs(int dp, int bn)
{
    return bn != 96 ? 4 : (dp <= 16 ? C3VQ_TC_GS_LOW : C3VQ_TC_GS);
}
#ifndef C3VQ_TC_LCAP
#define C3VQ_TC_LCAP 6
#endif
__host__ __device__ constexpr int tc_lcap(int dp) { return dp <= 48 ? C3VQ_TC_LCAP : 4; }   // chunk-list capacity per drain thread
constexpr int TC_QCAP = 1024;         // re-rank queue capacity per pass (groups of 4 codes)
constexpr int TC_MAXSTAGES = 6;
constexpr int TC_DRAIN_WARP0 = 0;    // warp order = issue priority on B200 (highest warp id first): misc warps last
constexpr int TC_NDRAIN = 16;          // drain warps: (tile t, lane quarter q, column half h)
constexpr int TC_DT = TC_NDRAIN * 32; // drain threads
constexpr int TC_PROD_WARP0 = TC_DRAIN_WARP0 + TC_NDRAIN;
constexpr int TC_MISC_WARP0 = TC_PROD_WARP0 + 8;        // TMA producer, MMA issuer, TMEM allocator, spare
constexpr int TC_THREADS = (TC_MISC_WARP0 + 4) * 32;   // 896
// register budgets per role (setmaxnreg): the CTA owns 896 * 72 = 64512 registers = 4*32*40 + 16*32*64 + 8*32*104.
// The unrolled drain loop keeps its addresses in uniform registers and fits 64; the producers (12 float4 of a codebook
// row plus the packed A operand) spilled at 88 (all 36 instantiations together: 3052 bytes of spill stores, 412 now)
constexpr int TC_REGS_MISC = 40, TC_REGS_DRAIN = 64, TC_REGS_PROD = 104;
constexpr uint32_t TC_SBO = 128;
constexpr float TC_BIG = 8.507059173023462e37f;  // 2^126: score of dead / padding codes
constexpr float TC_HUGE2 = 1.2676506002282294e30f;  // 2^100: squared-norm limit of the error model
constexpr int TC_SMEM_LIMIT = 227 * 1024;
// fp16 screening (kind::f16, twice the tensor rate of kind::tf32 at the same 2^-11 unit roundoff): operands are
// scaled by gamma = 2^g so that gamma * max |c'_j| < 2^7; the norm pieces ride against a constant 256 in A so that
// three pieces of 65504 mark a dead / padding code far above every real score of a row inside the range limit
// Fused sums, large inputs: the CTAs add into TC_ACC_PARTS partial tables that are folded pairwise afterwards.  A
// table row that owns 10^5-10^6 rows would otherwise be the end of as long a chain of fp32 additions at the L2
// (2.5e-5 of centroid error against float64 at 5.8 M rows; 3e-6 with 32 partial tables, profiles/r2f_sums_probe.txt).
constexpr int TC_ACC_PARTS = 32;
constexpr int64_t TC_ACC_PARTS_MIN_ROWS = (1 << 18) + 1;   // above the reference's colour mini-batch (2^18 rows): there
                                                           // the chains are short and the fold's ~12 us would show
constexpr float TC_F16_ACONST = 256.f;
constexpr float TC_F16_DEAD = 3.f * 65504.f * 256.f;     // 5.03e7
constexpr float TC_F16_XLIM = 1073741824.f;             // 2^30: limit of gamma^2 |x'|^2 (every element below 2^15)

// Workspace header (device memory, zeroed per call).
struct TcHeader {
    unsigned long long rows;        // stats: rows processed by the tensor-core kernel
    unsigned long long cands;       // stats: codes re-ranked exactly (4 per surviving group)
    unsigned long long fb_rows;     // stats: rows sent to the exact fallback
    unsigned int fb_count;          // fallback list length
    unsigned int force_fallback;    // set by prep when a finite code breaks the magnitude limit
    int cmax_bits;                  // max |c'_j| over the regular codes, float bits (non-negative)
    int live;                       // number of regular codes
    float mu_sum[72];               // column sums over the regular codes
    int thr_exp;                    // regular code <=> finite and exponent of |c_j|^2 <= thr_exp (k_tc_prep_mean)
    int xs_max_bits;                // float bits of max |x_i|^2 over a sample of the rows (k_tc_prep_hist)
    int excl_gap_bits;              // 0x7f800000 - float bits of min |c'_j| over the finite outlier codes (0: none)
    unsigned int exp_hist[256];     // exponent histogram of |c_j|^2 over the finite codes (k_tc_prep_hist)
    unsigned long long prof[32];    // developer build (C3VQ_TC_PROF): cycle counters per role/phase
    int g_exp;                      // fp16 screening: operands are scaled by 2^g_exp (k_tc_prep_mean)
};

// padded K extent: tf32 steps of 8 with at least two norm columns, fp16 steps of 16 with three
__host__ __device__ inline int tc_dp(int d, bool f16 = false) { return f16 ? ((d + 3 + 15) & ~15) : ((d + 2 + 7) & ~7); }
__host__ __device__ inline int tc_nc(int d, bool f16 = false) { int r = tc_dp(d, f16) - d; return r < 3 ? r : 3; }  // norm columns

struct TcParams {
    const float *x;
    const float *w;
    const int64_t *rows;
    int64_t n;
    int d, k, nsteps, nstages;
    int64_t npairs;
    const float *bimg;      // [k_pad/bn][dp*esize/16][bn][16 bytes]: tf32 (or scaled fp16) image of -2c' and the norm pieces
    // raw codebook for the exact distances, float4-column major: float4 c4 of code j at cbp4[c4 * kq + j] (columns
    // zero padded to dpc, a multiple of 4; kq = k rounded up to 8).  The lanes that re-rank the consecutive codes of a
    // group then read consecutive 16-byte pieces: one L1 wavefront per group and load instead of one per code (with
    // row-major rows the re-rank's divergent loads were half of all L1 wavefronts of the kernel, r2r ncu capture)
    const float *cbp;
    int dpc, kq;
    const float *mu;        // [d] centre
    TcHeader *hdr;
    float *out_dist;
    int64_t *out_idx;
    float *acc;             // fused sums: table (or acc_parts partial tables of acc_part_elems floats each)
    int acc_stride;
    int acc_parts;          // CTA b adds into partial table b % acc_parts
    size_t acc_part_elems;
    unsigned int *fb_list;
    unsigned long long *fb_best;   // per fallback row: running (distance bits, code) minimum
    float *dbg_scores, *dbg_norm2, *dbg_band;
    float gamma2;           // relative slack for the reference chain's own rounding
    // Seeds (full-batch loops call again and again with the same rows): the winner of row i in the previous call on this
    // workspace.  Its distance under the CURRENT codebook is an upper bound of the row's minimum, so the drain starts
    // with a tight threshold instead of learning it chunk by chunk (far fewer trips through the list path).  Any value
    // is safe: the bound is evaluated against the current codebook, stale or foreign entries only make it loose.
    int *prev;              // [n] or null (rows indirection, debug build)
    const uint8_t *reg;     // [k] 1 = regular code (has a real score in the screening)
};

// ------------------------------------------------------------------------------------------------
// Outlier codewords.  The error band of a row grows with Cmax = max |c'_j|, so a single far-away codeword would put
// every code of every row inside the band.  They are a fact of the reference's own pipeline: the shape codebook is
// divided by its trace after every step (vq.py:76-77), and a codeword nobody uses keeps its random initial value, whose
// trace can be anything, 0 included.  A code is "regular" when its squared norm lies at most one binade above the
// larger of (the 75th percentile over the codebook, the largest |x_i|^2 among 2 k rows (K / 2) spread over the input) and at most 8
// binades above that percentile; live codewords are weighted means of rows and so always regular, a freshly drawn
// uniform-box codebook has its norms within one binade of its own 75th percentile.  The other codes are kept out of
// the screening (score 2^126, like a non-finite code) and out of mu / Cmax.  That is exact, not
// a heuristic, for every row that passes the test the producer warps make after the re-rank: with Cx = min |c'_j|
// over the outliers, every outlier is at least Cx - |x'| away from the row, and the row keeps the winner among the
// regular codes only if 0.98 (Cx - 1.001 |x'|)^2 exceeds that winner's exact distance — a margin no fp32 rounding of
// the reference chain, of the centring or of the norms can bridge.  Rows that fail it (the ones that really live out
// there) take the exact scan.  The percentile only decides how many rows that is.
constexpr int TC_OUTLIER_BINADES = 8;
constexpr int TC_PREP_CODES = 32;       // codes per block of the two reduction kernels (4 per warp: short latency chains)

// prep 0: exponent histogram of |c_j|^2 over the finite codes (warp per code)
__global__ void __launch_bounds__(256) k_tc_prep_hist(const float *__restrict__ cb, int k, int d,
                                                      const float *__restrict__ x, const int64_t *__restrict__ rows,
                                                      int64_t n, TcHeader *hdr)
{
    __shared__ unsigned int sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    {   // data scale: the largest |x_i|^2 among gridDim.x * 16 rows spread evenly over the input (warp per row)
        const int64_t S = (int64_t)gridDim.x * 16;
        int bits = 0;
#pragma unroll
        for (int t = 0; t < 2; t++) {
            const int64_t si = (int64_t)blockIdx.x * 16 + wib * 2 + t;
            if (si < n) {
                const int64_t i = S < n ? (int64_t)(((unsigned long long)si * (unsigned long long)n) / (unsigned long long)S) : si;
                const float *row = x + (size_t)(rows ? rows[i] : i) * d;
                const float v0 = lane < d ? row[lane] : 0.f;
                const float v1 = lane + 32 < d ? row[lane + 32] : 0.f;
                float r = fmaf(v0, v0, v1 * v1);
                for (int o = 16; o; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
                if (isfinite(r)) bits = max(bits, __float_as_int(r));
            }
        }
        if (lane == 0 && bits) atomicMax(&hdr->xs_max_bits, bits);
    }
    const int j0 = blockIdx.x * TC_PREP_CODES;
#pragma unroll
    for (int j = j0 + wib; j < min(k, j0 + TC_PREP_CODES); j += 8) {
        const float *row = cb + (size_t)j * d;
        const float v0 = lane < d ? row[lane] : 0.f;
        const float v1 = lane + 32 < d ? row[lane + 32] : 0.f;
        float r = fmaf(v0, v0, v1 * v1);
        for (int o = 16; o; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
        if (lane == 0 && isfinite(r)) atomicAdd(&sh[(__float_as_uint(r) >> 23) & 0xff], 1u);
    }
    __syncthreads();
    if (sh[threadIdx.x]) atomicAdd(&hdr->exp_hist[threadIdx.x], sh[threadIdx.x]);
}

// prep 1: column sums over the regular codes (for the centre mu)
__global__ void __launch_bounds__(256) k_tc_prep_mean(const float *__restrict__ cb, int k, int d, TcHeader *hdr)
{
    // warp per code (lanes over columns, coalesced row reads), TC_PREP_CODES codes per block
    __shared__ float ssum[72];
    __shared__ int slive;
    __shared__ int sthr;
    __shared__ unsigned int shist[256];
    if (threadIdx.x < 72) ssum[threadIdx.x] = 0.f;
    shist[threadIdx.x] = hdr->exp_hist[threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0) {
        slive = 0;
        unsigned int total = 0, run = 0;
        for (int e = 0; e < 256; e++) total += shist[e];
        int e75 = 255;
        for (int e = 0; e < 256; e++) {
            run += shist[e];
            if (4ull * run >= 3ull * total) { e75 = e; break; }
        }
        const int ex = (hdr->xs_max_bits >> 23) & 0xff;
        sthr = min(min(max(ex, e75) + 1, e75 + TC_OUTLIER_BINADES), 254);
#ifdef C3VQ_TC_NO_OUTLIERS
        sthr = 254;                  // developer build: every finite code is regular (the behaviour before r1q)
#endif
        if (blockIdx.x == 0) {
            hdr->thr_exp = sthr;
            // fp16 screening: gamma = 2^g with gamma * |c'_j| < 2^7 for every regular code (|c_j|^2 < 2^(E+1) and
            // |mu| <= max |c_j| give |c'_j| < 2^((E+3)/2)).  The scale follows the regular codes only (sthr is a
            // percentile of the codebook, at most 8 binades above it): a single far-away ROW must not shrink gamma,
            // or every ordinary operand would sink into the fp16 subnormals.  Rows more than 2^8 above that scale
            // exceed the range limit and take the exact scan.
            hdr->g_exp = 7 - ((sthr - 127 + 4) >> 1);
        }
    }
    __syncthreads();
    const int thr = sthr;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    float s0 = 0.f, s1 = 0.f;
    int live = 0;
    const int j0 = blockIdx.x * TC_PREP_CODES;
#pragma unroll
    for (int j = j0 + wib; j < min(k, j0 + TC_PREP_CODES); j += 8) {
        const float *row = cb + (size_t)j * d;
        const float v0 = lane < d ? row[lane] : 0.f;
        const float v1 = lane + 32 < d ? row[lane + 32] : 0.f;
        float r = fmaf(v0, v0, v1 * v1);
        for (int o = 16; o; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
        const bool fin = __all_sync(0xffffffffu, isfinite(v0) && isfinite(v1)) && isfinite(r) &&
                         (int)((__float_as_uint(r) >> 23) & 0xff) <= thr;
        if (fin) { s0 += v0; s1 += v1; live++; }
    }
    if (lane < d) atomicAdd(&ssum[lane], s0);
    if (lane + 32 < d) atomicAdd(&ssum[lane + 32], s1);
    if (lane == 0) atomicAdd(&slive, live);
    __syncthreads();
    if (threadIdx.x < d) atomicAdd(&hdr->mu_sum[threadIdx.x], ssum[threadIdx.x]);
    if (threadIdx.x == 0) atomicAdd(&hdr->live, slive);
}

// prep 2: one thread per (padded) code: centre, norms, TF32 image, padded raw copy, Cmax
__global__ void k_tc_prep_image(const float *__restrict__ cb, int k, int k_pad, int d, int dp, int nc, int dpc,
                                int bn, int f16, TcHeader *hdr, float *__restrict__ mu_out, float *__restrict__ bimg,
                                float *__restrict__ cbp, int kq, uint8_t *__restrict__ reg)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k_pad) return;
    const int live = hdr->live;
    const float inv = live > 0 ? 1.f / (float)live : 0.f;
    float *img = bimg + (size_t)(j / bn) * (bn * dp) + (size_t)(j % bn) * 4;   // + chunk*BN*4 + (kk&3)
    // fp16 image: 16-byte chunks of 8 elements, + chunk*BN*8 + (kk&7)
    __half *img16 = reinterpret_cast<__half *>(bimg) + (size_t)(j / bn) * (bn * dp) + (size_t)(j % bn) * 8;
    const int g = hdr->g_exp;
    bool fin = j < k;
    double n2 = 0.0;
    float raw2 = 0.f;
    if (j < k) {
        const float *row = cb + (size_t)j * d;
#pragma unroll 8
        for (int c = 0; c < d; c++) {                        // unrolled: the row loads of a group issue together
            const float v = row[c];
            fin = fin && isfinite(v);
            const float m = hdr->mu_sum[c] * inv;
            if (j == 0) mu_out[c] = m;
            const float cc = v - m;
            n2 += (double)cc * (double)cc;
            raw2 = fmaf(v, v, raw2);
            cbp[((size_t)(c >> 2) * kq + j) * 4 + (c & 3)] = v;
        }
        for (int c = d; c < dpc; c++) cbp[((size_t)(c >> 2) * kq + j) * 4 + (c & 3)] = 0.f;
    }
    const float nf = (float)n2;
    if (fin && !(isfinite(raw2) && (int)((__float_as_uint(raw2) >> 23) & 0xff) <= hdr->thr_exp)) {
        // finite outlier: out of the screening; the producers check every row against the nearest of them
        atomicMax(&hdr->excl_gap_bits, 0x7f800000 - __float_as_int(fminf(sqrtf(nf), __int_as_float(0x7f800000))));
        fin = false;
    }
    if (fin && !(nf < TC_HUGE2)) {      // finite code outside the error model: everything goes exact
        atomicOr(&hdr->force_fallback, 1u);
    }
    {   // Cmax: one atomic per warp instead of one per code (k_pad is a multiple of BN, which need not be a multiple of
        // 32: the last warp may be partial, its lanes past k_pad have returned above)
        const int cb_bits = __reduce_max_sync(__activemask(), fin ? __float_as_int(sqrtf(nf)) : 0);
        if ((threadIdx.x & 31) == 0 && cb_bits) atomicMax(&hdr->cmax_bits, cb_bits);
    }
    if (j < k) reg[j] = (fin && nf < TC_HUGE2) ? 1 : 0;     // the code takes part in the screening with its real score
    if (f16) {
        // scaled fp16 operands: -2 gamma c' and gamma^2 |c'|^2 / 256 in three pieces (exact splits: every
        // remainder fits the next piece unless it falls below the fp16 subnormal step, 2^-25 absolute)
        const float vq = ldexpf(nf, 2 * g - 8);
        const __half h1 = __float2half_rn(vq);
        const float r1 = vq - __half2float(h1);
        const __half h2 = __float2half_rn(r1);
        const __half h3 = __float2half_rn(r1 - __half2float(h2));
        for (int c = 0; c < dp; c++) {
            __half v = __float2half_rn(0.f);
            if (fin && nf < TC_HUGE2) {
                if (c < d) {
                    const float cc = cb[(size_t)j * d + c] - hdr->mu_sum[c] * inv;
                    v = __float2half_rn(ldexpf(-2.f * cc, g));
                } else if (c < d + 3) {
                    v = (c == d) ? h1 : (c == d + 1) ? h2 : h3;
                }
            } else if (c < d + 3 && c >= d) {
                v = __float2half_rn(65504.f);   // dead or padding code: score 3 * 65504 * 256, never inside any band
            }
            img16[(size_t)(c >> 3) * (bn * 8) + (c & 7)] = v;
        }
        return;
    }
    for (int c = 0; c < dp; c++) {
        float v = 0.f;
        if (fin && nf < TC_HUGE2) {
            if (c < d) {
                const float cc = cb[(size_t)j * d + c] - hdr->mu_sum[c] * inv;
                v = -2.f * __uint_as_float(to_tf32(cc));
            } else if (c < d + nc) {
                const float h = __uint_as_float(to_tf32(nf));
                const float r1 = nf - h;
                const float mpart = __uint_as_float(to_tf32(r1));
                const float l = __uint_as_float(to_tf32(r1 - mpart));
                v = (c == d) ? h : (c == d + 1) ? mpart : l;
            }
        } else if (c == d) {
            v = TC_BIG;                  // dead or padding code: score 2^126, never inside any band
        }
        img[(size_t)(c >> 2) * (bn * 4) + (c & 3)] = v;
    }
}

// ------------------------------------------------------------------------------------------------
// Shared-memory plan of k_assign_tc (bytes; every region 16-byte aligned, sB/sA 128-byte aligned)
template <int DP, bool F16>
struct TcSmem {
    static constexpr int KSTEPS = F16 ? DP / 16 : DP / 8;                  // MMAs per stage (K = 16 fp16 or 8 tf32 each)
    static constexpr int BN = F16 ? tc_bn_f16(DP) : tc_bn(DP);                       // codes per stage / per MMA
    static constexpr int ESIZE = F16 ? 2 : 4;                              // bytes per operand element
    static constexpr int ACOLS = F16 ? DP / 2 : DP;                        // TMEM columns of one row tile's A operand
    static constexpr int GS = tc_gs(DP, BN);                               // codes per candidate group (3, 4, 8 or 16)
    static constexpr int NG = BN / (2 * GS);                               // groups per drain thread and (half) job
    static_assert(GS == 3 || GS == 4 || GS == 8 || GS == 16, "group size");
    // Drain warps per (tile, lane quarter, SLOT) instead of (tile, lane quarter, column half): a warp takes every other
    // step and reads all 96 columns of it in two batches, one wait and one hand-back per 96 columns instead of per 48.
    // Measured (same box, 5.8 M rows, mean of iterations 1-9): D = 7 2.03 ms against 2.17 ms; D = 45 2.56 against 2.57,
    // D = 48 2.79 against 2.79 -- the narrow widths, where the drain is all there is, take it; the wide rows stay on
    // the column-half mapping (C3VQ_TC_SLOTWARPS_ALL switches them too).
#ifdef C3VQ_TC_SLOTWARPS_ALL
    static constexpr bool SLOTWARPS = BN == 96 && tc_spt(DP) == 2;
#else
    static constexpr bool SLOTWARPS = BN == 96 && tc_spt(DP) == 2 && DP <= 16;
#endif
    static constexpr int LPI = GS == 3 ? 3 : 4;                            // re-rank lanes per queue item
    static constexpr int CPL = GS == 3 ? 1 : GS / 4;                       // codes per re-rank lane
    // Tried for the groups of 16 and dropped (round 2): group r = the 16 columns r, r + 3, r + 6, ... so that every first
    // minimum takes one column of each of the three TMEM loads and ptxas cannot sink a load below the reduction of the
    // one before it (it does with contiguous groups of 8 and 16).  5.8 M x 7: 3.35 ms against 3.41 ms for groups of 4 on
    // the same box, where contiguous groups of 16 with their serialised loads gain 6 % (3.05 vs 3.23 ms): the 48 live
    // columns spill, and the drain is bound by issue slots, not by the TMEM latency.
    static_assert((BN / 2) % GS == 0 && (BN != 112 || GS == 4), "group size must divide the half job");
    static constexpr int LCAP = tc_lcap(DP);
    static constexpr uint32_t LBO_B = BN * 16;
    static constexpr uint32_t BSTAGE = BN * DP * ESIZE;
    // staged row stride in floats: rows are 16-byte aligned (bulk reduce source) and XS / 4 is odd, which keeps the
    // lane <-> row accesses at 4-way bank conflicts (lane <-> column accesses are conflict free)
    static constexpr int XS = DP + 4;
    static constexpr uint32_t XS_BYTES = 2u * TC_PR * XS * 4;              // raw rows, two passes in flight
    static constexpr uint32_t W_BYTES = 2u * TC_PR * 4;                    // row weights, two passes in flight
    static constexpr uint32_t BAND_BYTES = 2u * TC_PR * 4;                 // per-row error band, two passes in flight
    static constexpr uint32_t SEED_BYTES = 2u * TC_PR * 4;                 // per-row initial running minimum, two passes
    // TMEM columns: NSLOT accumulator slots of BN (two per row tile), then the A operand of both row tiles
    static constexpr int SPT = tc_spt(DP);
    static constexpr int NSLOT = 2 * SPT;
    static constexpr uint32_t ACOL = NSLOT * BN;
    static_assert(ACOL + 2 * ACOLS <= 512, "TMEM budget");
    static constexpr uint32_t LIST_BYTES = 2u * TC_DT * LCAP * 4;          // score words + payload words
    static constexpr uint32_t QUEUE_BYTES = 2u * TC_QCAP * 4;              // two passes in flight
    static constexpr uint32_t BEST_BYTES = TC_PR * 8;
    static constexpr uint32_t FLAG_BYTES = 4u * TC_PR;                     // per-row, per-half status, two passes in flight
    static constexpr uint32_t XCHG_BYTES = 4u * TC_PR * 4;                 // running minima exchanged between column halves
    static constexpr uint32_t MISC_BYTES = 80 * 4 /*mu*/ + 64 /*counters, tmem ptr*/ + 40 * 8 /*barriers*/;
    static constexpr uint32_t FIXED = XS_BYTES + W_BYTES + BAND_BYTES + SEED_BYTES + LIST_BYTES + QUEUE_BYTES + BEST_BYTES + FLAG_BYTES +
                                      XCHG_BYTES + MISC_BYTES + 128 /*alignment*/;
    static int stages()
    {
        int s = (TC_SMEM_LIMIT - (int)FIXED) / (int)BSTAGE;
        return s > TC_MAXSTAGES ? TC_MAXSTAGES : s;
    }
    static size_t bytes(int s) { return (size_t)FIXED + (size_t)s * BSTAGE; }
};

// minimum of 4 consecutive scores: FMNMX3 + FMNMX
__device__ __forceinline__ float min4(uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    return fminf(fminf(fminf(__uint_as_float(a), __uint_as_float(b)), __uint_as_float(c)), __uint_as_float(d));
}
__device__ __forceinline__ float min3f(float a, float b, float c) { return fminf(fminf(a, b), c); }
// minimum of 8 consecutive scores: three FMNMX3 + one FMNMX, depth 3
__device__ __forceinline__ float min8(const uint32_t *a)
{
    const float r0 = min3f(__uint_as_float(a[0]), __uint_as_float(a[1]), __uint_as_float(a[2]));
    const float r1 = min3f(__uint_as_float(a[3]), __uint_as_float(a[4]), __uint_as_float(a[5]));
    return fminf(min3f(r0, r1, __uint_as_float(a[6])), __uint_as_float(a[7]));
}
// minimum of 16 consecutive scores: seven FMNMX3 + one FMNMX, depth 3
__device__ __forceinline__ float min16(const uint32_t *a)
{
    const float r0 = min3f(__uint_as_float(a[0]), __uint_as_float(a[1]), __uint_as_float(a[2]));
    const float r1 = min3f(__uint_as_float(a[3]), __uint_as_float(a[4]), __uint_as_float(a[5]));
    const float r2 = min3f(__uint_as_float(a[6]), __uint_as_float(a[7]), __uint_as_float(a[8]));
    const float r3 = min3f(__uint_as_float(a[9]), __uint_as_float(a[10]), __uint_as_float(a[11]));
    const float r4 = min3f(__uint_as_float(a[12]), __uint_as_float(a[13]), __uint_as_float(a[14]));
    return fminf(min3f(r0, r1, r2), min3f(r3, r4, __uint_as_float(a[15])));
}

// row status handed from the drain warps to the re-rank warps
constexpr uint8_t ROW_OK = 0, ROW_FALLBACK = 1, ROW_SKIP = 2;

// DCONST > 0: feature width known at compile time (6, 7, 45, 48); DCONST == 0: runtime width p.d <= DP - 2 (- 3).
// F16: screening with kind::f16 on scaled fp16 operands instead of kind::tf32 (same unit roundoff, twice the rate).
template <int DP, int DCONST, bool DBG, bool F16>
__global__ void __launch_bounds__(TC_THREADS, 1) k_assign_tc(const TcParams p)
{
    using L = TcSmem<DP, F16>;
    constexpr int KSTEPS = L::KSTEPS;
    constexpr int BN = L::BN;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    // aligned by pointer arithmetic on the array itself: a round trip through an integer would lose the address space
    // and turn every access below into a generic LD.E / ST.E (it did until round 2: 407 generic against 77 shared)
    uint8_t *smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);

    const int D = DCONST > 0 ? DCONST : p.d;
    const int S = p.nstages;
    uint8_t *sB = smem;                                        // S * BSTAGE
    float *sX = reinterpret_cast<float *>(sB + (size_t)S * L::BSTAGE);     // [2][PR][XS] raw rows of two passes
    float *sW = sX + 2 * TC_PR * L::XS;                                     // [2][PR] row weights
    float *sBand = sW + 2 * TC_PR;                                          // [2][PR]
    float *sSeed = sBand + 2 * TC_PR;                                       // [2][PR]
    uint32_t *sListS = reinterpret_cast<uint32_t *>(sSeed + 2 * TC_PR);     // [LCAP][DT] chunk-minimum score bits
    uint32_t *sListP = sListS + TC_DT * L::LCAP;                            // [LCAP][DT] (chunk << 12) | group mask
    uint32_t *sQueue = sListP + TC_DT * L::LCAP;                            // [2][QCAP] (row << 16) | group
    unsigned long long *sBest = reinterpret_cast<unsigned long long *>(sQueue + 2 * TC_QCAP);   // [PR]
    uint8_t *sFlag = reinterpret_cast<uint8_t *>(sBest + TC_PR);            // [2 passes][2 halves][PR]
    float *sXchg = reinterpret_cast<float *>(sFlag + 4 * TC_PR);            // [2 passes][2 halves][PR]
    float *sMu = sXchg + 4 * TC_PR;                                         // [80]
    uint32_t *sQCount = reinterpret_cast<uint32_t *>(sMu + 80);             // [2]
    uint32_t *sTmem = sQCount + 2;
    uint64_t *sBar = reinterpret_cast<uint64_t *>(sQCount + 16);
    // barrier indices
    const uint32_t bar0 = smem_u32(sBar);
    auto B_FULL = [&](int s) { return bar0 + 8u * s; };
    auto B_EMPTY = [&](int s) { return bar0 + 8u * (TC_MAXSTAGES + s); };
    const uint32_t A_FULL = bar0 + 8u * (2 * TC_MAXSTAGES), A_EMPTY = A_FULL + 8u;
    auto ACC_FULL = [&](uint32_t sl) { return A_EMPTY + 8u * (1 + sl); };
    auto ACC_EMPTY = [&](uint32_t sl) { return A_EMPTY + 8u * (1 + L::NSLOT + sl); };
    auto Q_FULL = [&](int b) { return A_EMPTY + 8u * (1 + 2 * L::NSLOT + b); };
    auto Q_EMPTY = [&](int b) { return A_EMPTY + 8u * (3 + 2 * L::NSLOT + b); };

    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const int lane = tid & 31;

    if (warp == TC_MISC_WARP0 && lane == 0) {
        for (int s = 0; s < TC_MAXSTAGES; s++) {
            mbar_init(B_FULL(s), 1);
            mbar_init(B_EMPTY(s), 2);                  // one commit per MMA warp (row tile)
        }
        mbar_init(A_FULL, 8);                      // one arrival per producer warp
        mbar_init(A_EMPTY, 2);
        for (int sl = 0; sl < L::NSLOT; sl++) {
            mbar_init(ACC_FULL(sl), 1);
            mbar_init(ACC_EMPTY(sl), L::SLOTWARPS ? 4 : 8);   // one arrival per drain warp of the tile (of the slot)
        }
        for (int b = 0; b < 2; b++) {
            mbar_init(Q_FULL(b), TC_NDRAIN);
            mbar_init(Q_EMPTY(b), 8);
            sQCount[b] = 0;
        }
        fence_mbar_init();
    }
    if (warp == TC_MISC_WARP0 + 2) {
        tmem_alloc(smem_u32(sTmem), 512);
        tmem_relinquish();
        // all 512 columns belong to this CTA, so the allocation starts at lane 0, column 0; the MMA warp relies on it
        if (*(volatile uint32_t *)sTmem != 0u) __trap();
    }
    if (tid < 80) sMu[tid] = tid < D ? p.mu[tid] : 0.f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    const int nsteps = p.nsteps;

    // setmaxnreg sits inside each warpgroup's own branch (ptxas applies a dec to all code reachable from it) and is
    // executed by all four warps of the warpgroup at ONE site (.aligned)
    if (warp >= TC_MISC_WARP0) {
      setmaxnreg_dec<TC_REGS_MISC>();
      if (warp == TC_MISC_WARP0) {
        // ===================== codebook stage producer (TMA engine, 1-D bulk copies) =====================
        if (lane == 0) {
            uint32_t stage = 0, phase = 1;              // ring position; parity of the "empty" phase to wait for
            const uint32_t sb0 = smem_u32(sB);
            PROF_DECL;
            for (int64_t pair = blockIdx.x; pair < p.npairs; pair += gridDim.x) {
                const char *src = reinterpret_cast<const char *>(p.bimg);
                for (int s = 0; s < nsteps; s++, src += L::BSTAGE) {
                    PROF(0);
                    mbar_wait_backoff(B_EMPTY(stage), phase);
                    PROF(1);   // prof[1]: TMA wait B_EMPTY
                    mbar_arrive_expect_tx(B_FULL(stage), L::BSTAGE);
                    bulk_g2s(sb0 + stage * L::BSTAGE, src, L::BSTAGE, B_FULL(stage));
                    if (++stage == (uint32_t)S) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
            PROF_FLUSH(0, true);
        }
      } else if (warp == TC_MISC_WARP0 + 1 || warp == TC_MISC_WARP0 + 2) {
        // ===================== MMA issuers: one warp per row tile =====================
        // Each warp runs its loop convergently (so addresses and descriptors live in uniform registers and the
        // waits are warp-wide); one elected lane issues the tcgen05 instructions.  Two warps because the issue
        // loop is a serial chain of ~50 instructions per job: one warp for both tiles cannot keep the tensor
        // pipe's queue filled.
        {
            // Everything the tcgen05 instructions consume is a compile-time constant or warp-uniform arithmetic
            // on kernel parameters, so it stays in uniform registers: the TMEM base is 0 by construction (the CTA
            // owns all 512 columns; checked after the allocation), and the B descriptor of (stage, ks) is the
            // stage's low word plus ks * DLO_KS under a constant high word.
            const int t = warp - (TC_MISC_WARP0 + 1);                                 // row tile of this warp
            constexpr uint32_t IDESC = F16 ? idesc_f16(TC_BM, BN) : idesc_tf32(TC_BM, BN);
            constexpr uint32_t DHI = ((TC_SBO >> 4) & 0x3FFFu) | (1u << 14);          // SBO, descriptor version 1
            constexpr uint32_t DLO_KS = (2 * L::LBO_B) >> 4, DLO_STAGE = L::BSTAGE >> 4;
            const uint32_t dlo0 = ((smem_u32(sB) & 0x3FFFFu) >> 4) | (((L::LBO_B >> 4) & 0x3FFFu) << 16);
            const uint32_t a_col = L::ACOL + t * L::ACOLS;
            uint32_t pi = 0, stage = 0, phase = 0, dlo = dlo0;
            // every pass starts on the tile's first slot (the drain warps unroll their loop over the slots); with an odd
            // number of steps the slots are used unequally often, so each keeps its own ACC_EMPTY phase bit (bit sk)
            uint32_t sk = 0, sphm = (1u << L::SPT) - 1u;
            PROF_DECL;
            for (int64_t pair = blockIdx.x; pair < p.npairs; pair += gridDim.x, pi++) {
                PROF(0);
                mbar_wait(A_FULL, pi & 1);
                PROF(1);   // prof[8+1]: MMA wait A_FULL
                tc_fence_after();
                if (!L::SLOTWARPS) sk = 0;
                for (int s = 0; s < nsteps; s++) {
                    PROF(0);
                    mbar_wait(B_FULL(stage), phase);
                    PROF(2);   // prof[8+2]: MMA wait B_FULL
                    // the tile owns L::SPT consecutive slots and cycles through them step by step
                    const uint32_t sl = L::SPT * t + sk;
                    mbar_wait(ACC_EMPTY(sl), (sphm >> sk) & 1u);
                    sphm ^= 1u << sk;
                    PROF(3);   // prof[8+3]: MMA wait ACC_EMPTY
                    tc_fence_after();
                    if (elect_one()) {
#pragma unroll
                        for (int ks = 0; ks < KSTEPS; ks++) {
                            const uint64_t bd = ((uint64_t)DHI << 32) | (uint64_t)(dlo + ks * DLO_KS);
                            if (F16) umma_f16_ts(sl * BN, a_col + ks * 8, bd, IDESC, ks > 0 ? 1u : 0u);
                            else umma_tf32_ts(sl * BN, a_col + ks * 8, bd, IDESC, ks > 0 ? 1u : 0u);
                        }
                        umma_commit(ACC_FULL(sl));
                        umma_commit(B_EMPTY(stage));      // second arrival comes from the other tile's warp
                    }
                    __syncwarp();
                    PROF(4);   // prof[8+4]: MMA issue of one job (KSTEPS instructions + commits)
                    dlo += DLO_STAGE;
                    if (++sk == (uint32_t)L::SPT) sk = 0;
                    if (++stage == (uint32_t)S) {
                        stage = 0;
                        phase ^= 1;
                        dlo = dlo0;
                    }
                }
                if (elect_one()) umma_commit(A_EMPTY);
                __syncwarp();
            }
            PROF(0);       // prof[8+0]: MMA issue + everything else
            PROF_FLUSH(8, lane == 0 && warp == TC_MISC_WARP0 + 1);
        }
      }
      // the spare warp: nothing to do until teardown
    } else if (warp >= TC_PROD_WARP0) {
        // ===================== feature producers + re-rankers (256 threads) =====================
        setmaxnreg_inc<TC_REGS_PROD>();
        const int ptid = tid - TC_PROD_WARP0 * 32;
        const int NC = DP - D < 3 ? DP - D : 3;
        // Mapping: producer warp pw owns rows pw*32 .. pw*32+31 of the pass, lane <-> column (kk = lane + 32*h).
        // Global loads are coalesced per row and every smem offset below is a per-thread base plus an immediate.
        constexpr int NH = DP > 32 ? 2 : 1;                       // column groups of 32 (D <= 64)
        const int pw = warp - TC_PROD_WARP0;
        float xr[32 * NH];
        float wr = 0.f;
        int pj = -1;                                             // previous winner of this lane's row (seed), -1 = none
        auto prefetch = [&](int64_t pair) {
            const int64_t r0 = pair * TC_PR + pw * 32;
            pj = (p.prev != nullptr && r0 + lane < p.n) ? p.prev[r0 + lane] : -1;
            if (p.rows == nullptr && r0 + 32 <= p.n) {
                // common case: 32 contiguous full rows -> one base pointer, immediate offsets
                const float *xp = p.x + r0 * D + lane;
#pragma unroll
                for (int r = 0; r < 32; r++) {
#pragma unroll
                    for (int h = 0; h < NH; h++) xr[r * NH + h] = (lane + 32 * h < D) ? xp[r * D + 32 * h] : 0.f;
                }
                wr = p.w ? p.w[r0 + lane] : 1.f;
            } else {
#pragma unroll
                for (int r = 0; r < 32; r++) {
                    const int64_t gi = r0 + r;
                    const bool valid = gi < p.n;
                    const int64_t src = valid ? (p.rows ? p.rows[gi] : gi) : 0;
                    const float *xp = p.x + src * D + lane;
#pragma unroll
                    for (int h = 0; h < NH; h++) xr[r * NH + h] = (valid && lane + 32 * h < D) ? xp[32 * h] : 0.f;
                }
                const int64_t gl = r0 + lane;
                wr = 0.f;
                if (gl < p.n) {
                    const int64_t src = p.rows ? p.rows[gl] : gl;
                    wr = p.w ? p.w[src] : 1.f;
                }
            }
        };
        // warp pw owns rows pw*32..+31 of the pass in both mappings: lane <-> column for the global loads and the
        // staging stores, lane <-> row (= TMEM lane, this warp's quarter pw & 3 of tile pw >> 2) for the A operand
        const uint32_t a_tmem = *(volatile uint32_t *)sTmem + ((uint32_t)((pw & 3) * 32) << 16) + L::ACOL + (pw >> 2) * L::ACOLS;
        // fp16 screening works in units scaled by gam = 2^g (scores and bands by gam^2): exact powers of two
        const float gam = F16 ? ldexpf(1.f, p.hdr->g_exp) : 1.f;
        const float ginv = F16 ? ldexpf(1.f, -p.hdr->g_exp) : 1.f;
        const float cmax = __int_as_float(p.hdr->cmax_bits) * gam;
        const float cexcl = __int_as_float(0x7f800000 - p.hdr->excl_gap_bits);   // nearest outlier code (inf: none)
        const bool has_outliers = p.hdr->excl_gap_bits != 0;
        // constants of the band formula below (C1, C2): 2 C2 Cmax^2 and 1 / (2 C1 Cmax)
        const float outl_c2 = 2.f * (3.814697265625e-6f + 4.76837158203125e-7f) * cmax * cmax * 0.999f;
        const float outl_inv = 1.001f / (2.f * (1.002f * 0.001953125f + 7.62939453125e-6f) * cmax);
        const bool force_fb = p.hdr->force_fallback != 0;
        const bool need_dist = p.out_dist != nullptr;
        unsigned long long st_rows = 0, st_cands = 0, st_fb = 0;

        PROF_DECL;
        // ---- re-rank of one finished pass: queue -> exact distances -> per-row minimum -> outputs ----
        auto rerank = [&](int64_t pair, uint32_t rp) {
            const int qb = rp & 1;
            const float *xq = sX + qb * (TC_PR * L::XS);                    // the staged raw rows of that pass
            mbar_wait_backoff(Q_FULL(qb), (rp >> 1) & 1);
            PROF(6);   // prof[16+6]: re-rank wait Q_FULL
#if defined(C3VQ_TC_EXP_NORERANK) || defined(C3VQ_TC_EXP_NOQUEUE)   // experiment: no exact re-rank (results are garbage)
            const uint32_t nq = 0;
#else
            const uint32_t nq = min(sQCount[qb], (uint32_t)TC_QCAP);
#endif
            const uint32_t *queue = sQueue + qb * TC_QCAP;
            constexpr int GS = L::GS;                                       // lanes per queue item = codes per group
            // Four lanes per queue item, each with GS / 4 codes of the group (lane i: codes i, i + 4, ...: the four lanes
            // read consecutive 16-byte pieces of the column-major copy).  With groups of 16 a lane runs four independent
            // distance chains; one lane per code would make a round of the loop four times as many serial L2 round trips.
            constexpr int CPL = L::CPL, LPI = L::LPI;                       // groups of 3: three lanes per item, ten items per warp
            constexpr int IPW = 32 / LPI;                                   // items per warp and round
            static_assert(CPL == 1 || DP <= 16, "several codes per lane only for the narrow widths (registers)");
            const int li = lane / LPI, lj = lane - li * LPI;                // item slot of this lane, lane within the item
            const bool lane_on = li < IPW;                                  // 32 = 10 * 3 + 2: two lanes sit out
            // (an L1 prefetch of the next round's pieces was measured and removed: once the loads were one wavefront per
            // item, its ~25 instructions per round cost 2.5 % of the kernel and hid nothing)
            sBest[ptid] = ((unsigned long long)__float_as_uint(FLT_MAX)) << 32;      // (FLT_MAX, code 0)
            named_bar_sync(1, TC_PR);
            const uint32_t gmask = ((1u << LPI) - 1u) << (li * LPI);
            for (uint32_t u = pw * IPW + li; lane_on && u < nq; u += 8 * IPW) {
                const uint32_t item = queue[u];
                const int lr = item >> 16;
                const int code0 = (int)(item & 0xFFFFu) * GS + lj;
                const float *xg = xq + lr * L::XS;
                // codes past the end of the codebook (last group) read the last code instead and are dropped below:
                // unconditional loads, no select per component
                const float4 *cb4 = reinterpret_cast<const float4 *>(p.cbp);
                const size_t kq = (size_t)p.kq;
                unsigned long long key = ~0ull;
                float a[CPL];
#pragma unroll
                for (int j = 0; j < CPL; j++) a[j] = 0.f;
                if (DCONST > 0) {
                    constexpr int NF4 = DCONST > 0 ? (DCONST + 3) / 4 : 1;
                    float4 cv[CPL][NF4];
#pragma unroll
                    for (int j = 0; j < CPL; j++)
#pragma unroll
                        for (int c4 = 0; c4 < NF4; c4++)
                            cv[j][c4] = __ldg(cb4 + c4 * kq + min(code0 + LPI * j, p.k - 1));
#pragma unroll
                    for (int c4 = 0; c4 < NF4; c4++) {
                        const float4 xv = reinterpret_cast<const float4 *>(xg)[c4];     // 16-byte aligned rows
#pragma unroll
                        for (int j = 0; j < CPL; j++) {
                            float df;
                            if (4 * c4 + 0 < D) { df = xv.x - cv[j][c4].x; a[j] = fmaf(df, df, a[j]); }
                            if (4 * c4 + 1 < D) { df = xv.y - cv[j][c4].y; a[j] = fmaf(df, df, a[j]); }
                            if (4 * c4 + 2 < D) { df = xv.z - cv[j][c4].z; a[j] = fmaf(df, df, a[j]); }
                            if (4 * c4 + 3 < D) { df = xv.w - cv[j][c4].w; a[j] = fmaf(df, df, a[j]); }
                        }
                    }
                } else {
                    for (int c4 = 0; c4 < p.dpc / 4; c4++) {
#pragma unroll
                        for (int j = 0; j < CPL; j++) {
                            const float4 cv = __ldg(cb4 + c4 * kq + min(code0 + LPI * j, p.k - 1));
                            float df;
                            if (4 * c4 + 0 < D) { df = xg[4 * c4 + 0] - cv.x; a[j] = fmaf(df, df, a[j]); }
                            if (4 * c4 + 1 < D) { df = xg[4 * c4 + 1] - cv.y; a[j] = fmaf(df, df, a[j]); }
                            if (4 * c4 + 2 < D) { df = xg[4 * c4 + 2] - cv.z; a[j] = fmaf(df, df, a[j]); }
                            if (4 * c4 + 3 < D) { df = xg[4 * c4 + 3] - cv.w; a[j] = fmaf(df, df, a[j]); }
                        }
                    }
                }
                // distances are sums of squares from +0: non-negative or NaN; as unsigned bit patterns
                // smaller distance <=> smaller key, NaN/inf sort above FLT_MAX, ties -> lower code
#pragma unroll
                for (int j = 0; j < CPL; j++) {
                    const unsigned long long kj = ((unsigned long long)__float_as_uint(a[j]) << 32) | (unsigned int)(code0 + LPI * j);
                    if (code0 + LPI * j < p.k && kj < key) key = kj;
                }
                if (LPI == 4) {
                    unsigned long long o = __shfl_xor_sync(gmask, key, 1);
                    key = o < key ? o : key;
                    o = __shfl_xor_sync(gmask, key, 2);
                    key = o < key ? o : key;
                } else {                                      // three lanes: each takes the keys of the other two
                    const int b0 = lane - lj;
                    const unsigned long long o1 = __shfl_sync(gmask, key, b0 + (lj + 1) % 3);
                    const unsigned long long o2 = __shfl_sync(gmask, key, b0 + (lj + 2) % 3);
                    key = o1 < key ? o1 : key;
                    key = o2 < key ? o2 : key;
                }
                if (lj == 0) atomicMin(&sBest[lr], key);
            }
            named_bar_sync(1, TC_PR);
            st_cands += (unsigned long long)GS * nq;      // counted once per thread, divided out below
            // ---- outputs (lane <-> row): idx / dist, fallback list, then the weighted row into the centroid sums ----
#ifndef C3VQ_TC_EXP_NORERANK
            {
                const int lr = pw * 32 + lane;
                const uint8_t st0 = sFlag[(qb * 2 + 0) * TC_PR + lr], st1 = sFlag[(qb * 2 + 1) * TC_PR + lr];
                uint8_t status = st0 > st1 ? st0 : st1;
                const int64_t gi = pair * TC_PR + lr;
                const unsigned long long best = sBest[lr];
                const int bj = (int)(best & 0xFFFFFFFFu);
                if (has_outliers && status == ROW_OK) {
                    // the nearest outlier code is at least cexcl - |x'| away; the row keeps its regular winner only
                    // if that bound clears the winner's exact distance with room for every rounding involved
                    // |x'| from the row's band (sBand, written when the row was staged): band >= 1.01 * 2 * (C1 |x'| Cmax +
                    // C2 Cmax^2), so the quotient below is an upper bound of |x'|, tight to a fraction of a percent
                    const float sx = (sBand[qb * TC_PR + lr] * (1.f / 1.01f) - outl_c2) * outl_inv * ginv;
                    const float gap = cexcl - 1.001f * sx;
                    const float lim = gap > 0.f ? 0.98f * gap * gap : 0.f;
                    if (!(__uint_as_float((unsigned int)(best >> 32)) < lim)) status = ROW_FALLBACK;
                }
                if (status == ROW_OK) {
                    st_rows++;
                    if (p.prev) p.prev[gi] = bj;
                    if (p.out_idx) p.out_idx[gi] = bj;
                    if (need_dist) p.out_dist[gi] = __uint_as_float((unsigned int)(best >> 32));
                } else if (status == ROW_FALLBACK) {
                    st_rows++;
                    st_fb++;
                    const unsigned int pos = atomicAdd(&p.hdr->fb_count, 1u);
                    p.fb_list[pos] = (unsigned int)gi;
                    p.fb_best[pos] = ((unsigned long long)0x7f7fffffu) << 32;          // (FLT_MAX, code 0)
                }
#ifndef C3VQ_TC_EXP_NORED
                if (p.acc) {
                    // scale this warp's 32 staged rows in place (lane <-> column): [w*x (D) | w | 0 ...] = one
                    // accumulator row, then ONE bulk reduce-add per DISTINCT codeword of the warp (TMA engine, f32
                    // adds at the L2): rows that share a codeword are summed into the first of them here.  While a
                    // few codewords own most rows (the first iterations from a uniform-box codebook) this cuts the
                    // reductions on their table rows, which the L2 performs one after the other, by up to 32x.
                    float *xw = sX + (qb * TC_PR + pw * 32) * L::XS;
                    const float *ww = sW + qb * TC_PR + pw * 32;
                    const unsigned int key = status == ROW_OK ? (unsigned int)bj : 0x80000000u + (unsigned int)lane;
                    const unsigned int same = __match_any_sync(0xffffffffu, key);          // lanes (rows) with this codeword
                    const int leader = __ffs(same) - 1;                                   // the lowest of them
                    const unsigned int leaders = __ballot_sync(0xffffffffu, leader == lane);
                    {
                        // every row is scaled in place as float4 (an accumulator row = the first AS / 4 float4 of a staged
                        // row): once the codebook is in use the 32 codewords of a warp are nearly always distinct and
                        // this is all there is to do
                        const int AS4 = (DCONST > 0 ? ((DCONST + 4) & ~3) : p.acc_stride) >> 2;
#pragma unroll 4
                        for (int i = lane; i < 32 * AS4; i += 32) {
                            const int r = i / AS4, c4 = i - r * AS4;
                            float4 *q = reinterpret_cast<float4 *>(xw + r * L::XS) + c4;
                            const float wv = ww[r];
                            float4 v = *q;
                            if (4 * c4 + 3 < D) {
                                v.x = __fmul_rn(v.x, wv); v.y = __fmul_rn(v.y, wv);
                                v.z = __fmul_rn(v.z, wv); v.w = __fmul_rn(v.w, wv);
                            } else {
                                const int c = 4 * c4;
                                v.x = c + 0 < D ? __fmul_rn(v.x, wv) : (c + 0 == D ? wv : 0.f);
                                v.y = c + 1 < D ? __fmul_rn(v.y, wv) : (c + 1 == D ? wv : 0.f);
                                v.z = c + 2 < D ? __fmul_rn(v.z, wv) : (c + 2 == D ? wv : 0.f);
                                v.w = c + 3 < D ? __fmul_rn(v.w, wv) : (c + 3 == D ? wv : 0.f);
                            }
                            *q = v;
                        }
                        // rows that are not the first of their codeword are added into that first one, in row order
                        // (lanes over columns; sources are never destinations)
                        if (leaders != 0xffffffffu) {
                            __syncwarp();
                            for (unsigned int nl = ~leaders; nl; nl &= nl - 1) {
                                const int r = __ffs(nl) - 1;
                                const int dst = __shfl_sync(0xffffffffu, leader, r);
                                // (acc_stride can exceed 64 columns: d = 61 ... 64)
                                for (int col = lane; col < p.acc_stride; col += 32)
                                    xw[dst * L::XS + col] = __fadd_rn(xw[dst * L::XS + col], xw[r * L::XS + col]);
                            }
                        }
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (status == ROW_OK && leader == lane)
                        bulk_reduce_add_f32(p.acc + (size_t)(blockIdx.x % p.acc_parts) * p.acc_part_elems +
                                                (size_t)bj * p.acc_stride,
                                            smem_u32(xw + lane * L::XS), (uint32_t)p.acc_stride * 4u);
                    bulk_commit_group();
                }
#endif
            }
#endif
            __syncwarp();
            if (ptid == 0) sQCount[qb] = 0;       // every thread has read nq (barrier above)
            if (lane == 0) mbar_arrive(Q_EMPTY(qb));
        };

        uint32_t pi = 0;
        int64_t prev_pair = -1;
        // software pipeline with ONE prefetch call site (two would make the prefetch registers spill):
        // iteration -1 only prefetches the first pass
        for (int64_t pair = (int64_t)blockIdx.x - gridDim.x;; pair += gridDim.x) {
          if (pair >= 0) {
            PROF(0);
            // stage the raw rows (exact fp32: the re-rank and the weighted sums of this pass read them back);
            // row stride XS is odd, so both the lane <-> column stores and the lane <-> row loads are conflict free.
            // Buffer pi & 1 was last read by the re-rank of pass pi - 2, which every producer warp has left
            // (its second named barrier) before any warp arrives here; the bulk reduces of that pass read only
            // this warp's own rows and are waited for below.
            float *xs_w = sX + ((pi & 1) * TC_PR + pw * 32) * L::XS;
            bulk_wait_group_read<1>();           // this thread's bulk reduce of pass pi - 2 no longer reads the buffer
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 32; r++) {
#pragma unroll
                for (int h = 0; h < NH; h++)
                    if (lane + 32 * h < D) xs_w[r * L::XS + lane + 32 * h] = xr[r * NH + h];
            }
            sW[(pi & 1) * TC_PR + ptid] = wr;
            __syncwarp();
            // seed: the flag of the previous winner and its codebook row are requested now and used after the conversion
            const bool seed_ok = (unsigned)pj < (unsigned)p.k;
            const float4 *seed_row = reinterpret_cast<const float4 *>(p.cbp) + (seed_ok ? pj : 0);   // code 0 when there is no seed
            uint8_t seed_reg = 0;
            if (seed_ok) seed_reg = p.reg[pj];
            uint32_t av[L::ACOLS];
            float X = 0.f;                       // |x'|^2 of the rounded, centred row (scaled by gam^2) -> error band
            // The row and the centre are read as float4 (rows are 16-byte aligned, stride XS / 4 odd: conflict free).
            // Known widths fuse the seed into the same pass over the row: sa = sum c'(c' - 2x') = |c'|^2 - 2 x'.c' in fp32
            // on the centred operands, the score of the previous winner (see below); its codebook pieces were asked
            // for before the loop.  Columns past D of the staged row hold leftovers of the previous sums and are masked.
            constexpr int NF4 = DCONST > 0 ? (DCONST + 3) / 4 : 1;
            constexpr bool SEED_FUSED = DCONST > 0;
            float4 scv[NF4];
            float sa = 0.f;
            if (SEED_FUSED) {
#pragma unroll
                for (int c4 = 0; c4 < NF4; c4++)
                    scv[c4] = __ldg(seed_row + c4 * (size_t)p.kq);          // unconditional (k >= 1); unused without a seed
            }
            const float4 *xrow4 = reinterpret_cast<const float4 *>(xs_w + lane * L::XS);
            const float4 *mu4 = reinterpret_cast<const float4 *>(sMu);
#pragma unroll
            for (int c4 = 0; c4 < DP / 4; c4++) {
                float xe[4] = {0.f, 0.f, 0.f, 0.f}, me[4] = {0.f, 0.f, 0.f, 0.f}, ce[4] = {0.f, 0.f, 0.f, 0.f};
                if (4 * c4 < D) {
                    const float4 xv = xrow4[c4], mv = mu4[c4];
                    xe[0] = xv.x; xe[1] = xv.y; xe[2] = xv.z; xe[3] = xv.w;
                    me[0] = mv.x; me[1] = mv.y; me[2] = mv.z; me[3] = mv.w;
                    if (SEED_FUSED && c4 < NF4) { ce[0] = scv[c4 < NF4 ? c4 : 0].x; ce[1] = scv[c4 < NF4 ? c4 : 0].y;
                                                  ce[2] = scv[c4 < NF4 ? c4 : 0].z; ce[3] = scv[c4 < NF4 ? c4 : 0].w; }
                }
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const int kk = 4 * c4 + e;
                    if (F16) {
                        // 256 against the three norm pieces, 0 beyond; two elements per 32-bit column, even K in the low half
                        uint32_t hb = (kk < D + 3) ? 0x5C00u : 0u;
                        if (kk < D) {
                            const float xc = xe[e] - me[e];
                            const __half h = __float2half_rn(xc * gam);
                            const float v = __half2float(h);
                            X = fmaf(v, v, X);
                            hb = (uint32_t)__half_as_ushort(h);
                            if (SEED_FUSED) { const float cc = ce[e] - me[e]; sa = fmaf(cc, cc - 2.f * xc, sa); }
                        }
                        if (kk & 1) av[kk >> 1] |= hb << 16; else av[kk >> 1] = hb;
                    } else {
                        float v = (kk < D + NC) ? 1.0f : 0.f;        // 1.0 against the norm pieces, 0 beyond
                        if (kk < D) {
                            const float xc = xe[e] - me[e];
                            v = __uint_as_float(to_tf32(xc));
                            X = fmaf(v, v, X);
                            if (SEED_FUSED) { const float cc = ce[e] - me[e]; sa = fmaf(cc, cc - 2.f * xc, sa); }
                        }
                        av[kk] = __float_as_uint(v);
                    }
                }
            }
            {
                const int64_t gi = pair * TC_PR + ptid;
                // E bounds |score + X - true distance| for every code
                const float C1 = 1.002f * 0.001953125f + 7.62939453125e-6f;   // 2^-9 * 1.002 + 2^-17
                const float C2 = 3.814697265625e-6f + 4.76837158203125e-7f;   // 2^-18 + 2^-21
                float E = C1 * sqrtf(X) * cmax + C2 * cmax * cmax;
                // fp16 only: elements below 2^-14 round with 2^-25 absolute error (against |B| <= 2 gam Cmax, resp.
                // |A| summing to <= sqrt(D X)), and the three norm pieces leave at most 3 * 2^-25 * 256
                if (F16) E += 6.0e-8f * (float)D * cmax + 3.0e-8f * sqrtf((float)D * X) + 3.1e-5f;
                float band = 1.01f * (2.f * E + p.gamma2 * X) + 3.0e-36f;
                if (!(X < (F16 ? TC_F16_XLIM : TC_HUGE2 * 0.0625f)) || force_fb ||
                    !(band < (F16 ? 0.25f * TC_F16_DEAD : TC_BIG)))
                    band = __int_as_float(0x7fc00000);                        // NaN: exact fallback
                if (gi >= p.n) band = -1.f;
                if (DBG && gi < p.n) {
                    p.dbg_norm2[gi] = X * ginv * ginv;
                    p.dbg_band[gi] = band * ginv * ginv;
                }
                sBand[(pi & 1) * TC_PR + ptid] = band;
                // seed of the running minimum: the score of the row's previous winner, evaluated in fp32 on the centred
                // operands (|c'|^2 - 2 x'.c' = sum c'(c' - 2x')), bounds the row's minimum screened score from above once
                // the screening error (<= E) and its own rounding (< E) are added: band >= 2.02 E covers both.  3e38 = none
                float seed = 3.0e38f;
                {
                    if (seed_ok && seed_reg != 0 && gi < p.n) {
                        const float4 *cp = seed_row;
                        const size_t kq = (size_t)p.kq;
                        const float *xg = xs_w + lane * L::XS;
                        float a = sa;
                        auto term = [&](int kk, float cvk) {
                            if (kk < D) {
                                const float mk = sMu[kk], cc = cvk - mk;
                                a = fmaf(cc, cc - 2.f * (xg[kk] - mk), a);
                            }
                        };
                        if (!SEED_FUSED) {
                            for (int c4 = 0; c4 < p.dpc / 4; c4++) {
                                const float4 cv = __ldg(cp + c4 * kq);
                                term(4 * c4 + 0, cv.x);
                                term(4 * c4 + 1, cv.y);
                                term(4 * c4 + 2, cv.z);
                                term(4 * c4 + 3, cv.w);
                            }
                        }
                        const float sd = a * (gam * gam) + band;
                        if (sd < 3.0e38f) seed = sd;             // NaN / inf: no seed
                    }
                }
                sSeed[(pi & 1) * TC_PR + ptid] = seed;
            }
            PROF(2);   // prof[16+2]: producer stage + convert + band
            mbar_wait(A_EMPTY, (pi & 1) ^ 1);          // previous pass' MMAs no longer read the A operand in TMEM
            PROF(1);   // prof[16+1]: producer wait A_EMPTY
            tc_fence_after();
#pragma unroll
            for (int c8 = 0; c8 < L::ACOLS / 8; c8++) {
                const uint32_t r8[8] = {av[c8 * 8 + 0], av[c8 * 8 + 1], av[c8 * 8 + 2], av[c8 * 8 + 3],
                                        av[c8 * 8 + 4], av[c8 * 8 + 5], av[c8 * 8 + 6], av[c8 * 8 + 7]};
                tmem_st8(a_tmem + c8 * 8, r8);
            }
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(A_FULL);
            PROF(7);   // prof[16+7]: A operand -> TMEM
            if (prev_pair >= 0) rerank(prev_pair, pi - 1);
            PROF(3);   // prof[16+3]: re-rank + outputs of the previous pass (includes waiting for its queue)
            prev_pair = pair;
            pi++;
          }
            if (pair + gridDim.x >= p.npairs) break;
            prefetch(pair + gridDim.x);
            PROF(4);   // prof[16+4]: prefetch issue
        }
        if (prev_pair >= 0) rerank(prev_pair, pi - 1);
        bulk_wait_group<0>();                    // every bulk reduce of this thread has been performed
        PROF(5);       // prof[16+5]: last re-rank
        PROF_FLUSH(16, lane == 0 && pw == 0);
        // stats: one atomic per warp (st_cands was counted by every thread of the 256)
        for (int o = 16; o > 0; o >>= 1) {
            st_rows += __shfl_xor_sync(0xffffffffu, st_rows, o);
            st_fb += __shfl_xor_sync(0xffffffffu, st_fb, o);
        }
        if (lane == 0) {
            atomicAdd(&p.hdr->rows, st_rows);
            atomicAdd(&p.hdr->fb_rows, st_fb);
            if (pw == 0) atomicAdd(&p.hdr->cands, st_cands);
        }
    } else {
        // ===================== drain: TMEM scan, running minimum, chunk lists, work queue =====================
        // Warp (tile t, lane quarter q, column half h) reads its half of every accumulator slot of its tile: two
        // threads per row, so a slot is back with the MMA warp after two short tcgen05.ld round trips.
        if (TC_REGS_DRAIN < 72) setmaxnreg_dec<TC_REGS_DRAIN < 72 ? TC_REGS_DRAIN : 64>();
        constexpr int GS = L::GS;                       // codes per candidate group
        constexpr int NGH = L::NG;                      // groups per half job (BN / 2 codes)
        constexpr int GPC = GS == 3 ? 0 : 16 / GS;      // groups per 16-column TMEM load (groups of 3 straddle the loads)
        // the warp index through a shuffle: the compiler then knows that everything derived from it (TMEM address,
        // barrier addresses) is warp-uniform and keeps it in uniform registers instead of recomputing it per step
        const int warp_u = __shfl_sync(0xffffffffu, warp, 0);
        const int dw = warp_u - TC_DRAIN_WARP0;         // 0..15
        const int q = warp_u & 3;                       // TMEM lane quarter this warp may access
        const int t = (dw >> 2) & 1;                    // row tile
        const int h = dw >> 3;                          // column half of every slot served by this warp
        const int lr = t * TC_BM + q * 32 + lane;       // local row 0..255 (shared with the other half's thread)
        const int dl = dw * 32 + lane;                  // drain thread 0..511 (list column)
        // (slot warps: h is the warp's slot, not its column half)
        const uint32_t tsrc0 = ((uint32_t)(q * 32) << 16) + (L::SPT * t) * BN + (L::SLOTWARPS ? 0 : h * (BN / 2));   // TMEM base is 0 (checked)
        uint32_t slot0 = 0;                             // slot warps: slot of the first step of the current pass
        const uint32_t acc_full0 = ACC_FULL(L::SPT * t);     // ACC_EMPTY of the same slot is NSLOT barriers further
        constexpr uint32_t ACC_E_OFF = 8u * L::NSLOT;
        uint32_t *listS = sListS + dl, *listP = sListP + dl;
        const uint32_t lS = smem_u32(listS);            // the payload plane is LIST_PLANE bytes further
        constexpr uint32_t LIST_STRIDE = TC_DT * 4, LIST_END = L::LCAP * LIST_STRIDE, LIST_PLANE = LIST_END;
        uint32_t pi = 0;
        uint32_t wphm = 0;                              // phase parity of the tile's slots (bit = slot)
        PROF_DECL;
        for (int64_t pair = blockIdx.x; pair < p.npairs; pair += gridDim.x, pi++) {
            PROF(0);
            mbar_wait(A_FULL, pi & 1);
            PROF(1);       // prof[24+1]: drain wait A_FULL
            float band = sBand[(pi & 1) * TC_PR + lr];            // computed by the producer of this row
            float m = sSeed[(pi & 1) * TC_PR + lr];               // 3e38, or an upper bound of the row's minimum score
            // unseeded valid rows start at 2^124: above every real score, below the 2^126 of dead / padding codes
            float thr = (band >= 0.f) ? fminf(F16 ? 0.5f * TC_F16_DEAD : 0.25f * TC_BIG, m + p.gamma2 * fabsf(m) + band)
                                      : __int_as_float(0xff800000);
            uint32_t cnt_off = 0;                       // byte offset of the next free list entry
            PROF(2);       // prof[24+2]: drain band

            // group minima of a 16-column chunk (FMNMX3 + FMNMX per group)
            auto groups = [&](const uint32_t (&a)[16], float *gm) {
                if (GS == 4) {
                    gm[0] = min4(a[0], a[1], a[2], a[3]);
                    gm[1] = min4(a[4], a[5], a[6], a[7]);
                    gm[2] = min4(a[8], a[9], a[10], a[11]);
                    gm[3] = min4(a[12], a[13], a[14], a[15]);
                } else if (GS == 8) {
                    gm[0] = min8(a);
                    gm[1] = min8(a + 8);
                } else {
                    gm[0] = min16(a);
                }
            };
            // half-job minimum against the running threshold; the append path is entered per WARP (uniform branch)
            auto process = [&](const float (&gv)[NGH], uint32_t chunk) {
                float cm = gv[0];
#pragma unroll
                for (int i = 1; i + 1 < NGH; i += 2) cm = fminf(fminf(cm, gv[i]), gv[i + 1]);
                if (NGH % 2 == 0) cm = fminf(cm, gv[NGH - 1]);
#ifdef C3VQ_TC_EXP_NOPROC     // experiment: running minimum only, no candidate lists (results are garbage)
                m = fminf(m, cm);
                return;
#endif
                const bool flag = cm <= thr;
                if (__any_sync(0xffffffffu, flag)) {
                    if (flag) {
                        m = fminf(m, cm);
                        thr = m + p.gamma2 * fabsf(m) + band;
                        // group mask without the ALU pipe's compare + select: the sign bit of thr - gv[i] (FADD) is
                        // shifted in by one funnel shift per group; group i ends up at bit NGH-1-i, set = candidate
                        uint32_t ng;
                        if (NGH % 2 != 0 || NGH < 6) {  // few groups: one chain
                            ng = 0;
#pragma unroll
                            for (int i = 0; i < NGH; i++) ng = __funnelshift_l(__float_as_uint(thr - gv[i]), ng, 1);
                        } else {                        // two independent chains of NGH / 2 dependent shifts, joined by one more
                            uint32_t ng0 = 0, ng1 = 0;
#pragma unroll
                            for (int i = 0; i < NGH / 2; i++) {
                                ng0 = __funnelshift_l(__float_as_uint(thr - gv[i]), ng0, 1);
                                ng1 = __funnelshift_l(__float_as_uint(thr - gv[NGH / 2 + i]), ng1, 1);
                            }
                            ng = (ng0 << (NGH / 2)) | ng1;
                        }
                        const uint32_t gm = ~ng & ((1u << NGH) - 1u);
                        if (cnt_off == LIST_END) {      // full: drop entries the tighter threshold excludes
                            uint32_t wpos = 0;
                            for (uint32_t e = 0; e < LIST_END; e += LIST_STRIDE) {
                                const uint32_t es = ld_shared_u32(lS + e), ep = ld_shared_u32(lS + LIST_PLANE + e);
                                if (__uint_as_float(es) <= thr) {
                                    st_shared_u32(lS + wpos, es);
                                    st_shared_u32(lS + LIST_PLANE + wpos, ep);
                                    wpos += LIST_STRIDE;
                                }
                            }
                            cnt_off = wpos;
                        }
                        if (cnt_off < LIST_END) {
                            st_shared_u32(lS + cnt_off, __float_as_uint(cm));
                            st_shared_u32(lS + LIST_PLANE + cnt_off, (chunk << 16) | gm);
                            cnt_off += LIST_STRIDE;
                        } else {
                            band = __int_as_float(0x7fc00000);   // overflow: NaN band = exact fallback for this row
                            thr = band;                          // and nothing is appended any more
                        }
                    }
                }
            };
            auto dump = [&](const uint32_t (&a)[16], int code0) {   // debug build of the kernel only (unscaled scores)
                const float g2inv = F16 ? ldexpf(1.f, -2 * p.hdr->g_exp) : 1.f;
#pragma unroll
                for (int jj = 0; jj < 16; jj++)
                    if (pair * TC_PR + lr < p.n && code0 + jj < p.k)
                        p.dbg_scores[(pair * TC_PR + lr) * p.k + code0 + jj] = __uint_as_float(a[jj]) * g2inv;
            };

            // Tried and dropped (round 2, A/B at 5.8 M x 45, fused sums, mean of iterations 1-7): probing the NEXT
            // step's barrier with a try_wait right after handing this step's slot back, to hide the ~90 cycles a
            // successful wait takes behind the min reduction: 3.63 ms per iteration against 3.46 ms without the probe
            // (the extra try_wait + predicate per step cost more issue slots than the hidden latency gave back)
            if (L::SLOTWARPS) {
            // a drain warp per (tile, lane quarter, SLOT): it takes every other step and reads all 96 columns of it in two
            // batches of 48 (TcSmem::SLOTWARPS).  The MMA warps keep cycling the slots across passes, so the slot of a
            // pass' first step alternates when the step count is odd.
            for (int s = (int)((h - slot0) & 1u); s < nsteps; s += 2) {
                const uint32_t acc_full = acc_full0 + 8u * h, tsrc = tsrc0 + h * BN;
                mbar_wait_hint(acc_full, wphm & 1u);
                wphm ^= 1u;
                tc_fence_after();
#pragma unroll
                for (int bt = 0; bt < 2; bt++) {
                    uint32_t va[16], vb[16], vc[16];
                    float gv[NGH];
                    tmem_ld16_async(tsrc + bt * 48, va);
                    tmem_ld16_async(tsrc + bt * 48 + 16, vb);
                    tmem_ld16_async(tsrc + bt * 48 + 32, vc);
                    tmem_wait_ld16(va);
                    tmem_wait_ld16(vb);
                    tmem_wait_ld16(vc);
                    if (bt == 1) {                      // the whole slot is read: hand it back to the MMA warp
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(acc_full + ACC_E_OFF);
                    }
                    if (DBG) { dump(va, s * BN + bt * 48); dump(vb, s * BN + bt * 48 + 16); dump(vc, s * BN + bt * 48 + 32); }
                    if (GS == 3) {
                        auto colv = [&](int c) {
                            return __uint_as_float(c < 16 ? va[c & 15] : (c < 32 ? vb[c & 15] : vc[c & 15]));
                        };
#pragma unroll
                        for (int g = 0; g < 16; g++) gv[g] = min3f(colv(3 * g), colv(3 * g + 1), colv(3 * g + 2));
                    } else {
                        groups(va, gv);
                        groups(vb, gv + GPC);
                        groups(vc, gv + 2 * GPC);
                    }
                    process(gv, (uint32_t)(2 * s + bt));
                }
            }
            slot0 = (slot0 + (uint32_t)nsteps) & 1u;
            } else
            // unrolled over the tile's slots: slot, TMEM address and barrier address of a step are constants of the body
            for (int s0 = 0; s0 < nsteps; s0 += L::SPT) {
#pragma unroll
              for (int par = 0; par < L::SPT; par++) {
                const int s = s0 + par;
                if (s >= nsteps) break;
                const uint32_t acc_full = acc_full0 + 8u * par, tsrc = tsrc0 + par * BN;
                PROF(4);   // prof[24+4]: drain (LDTM + min + lists)
                mbar_wait_hint(acc_full, (wphm >> par) & 1u);
                wphm ^= 1u << par;
                PROF(3);   // prof[24+3]: drain wait ACC_FULL
                tc_fence_after();
                const int code_s = s * BN + h * (BN / 2);
#ifdef C3VQ_TC_EXP_NODRAIN   // experiment: hand the slot straight back (results are garbage)
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(acc_full + ACC_E_OFF);
                continue;
#endif
                uint32_t va[16], vb[16];
                float gv[NGH];
                if (BN == 96) {
                    // the three loads of the half job, the slot handed back as soon as they have landed, then the minima
                    // (with groups of 4 ptxas keeps all 48 columns in flight; with groups of 8 / 16 it sinks each load
                    // below the reduction of the one before it, which keeps fewer columns live and costs nothing: see TcSmem)
                    uint32_t vc[16];
                    tmem_ld16_async(tsrc, va);
                    tmem_ld16_async(tsrc + 16, vb);
                    tmem_ld16_async(tsrc + 32, vc);
                    tmem_wait_ld16(va);
                    tmem_wait_ld16(vb);
                    tmem_wait_ld16(vc);
                    tc_fence_before();                  // this half of the slot is read: hand it back to the MMA warp
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acc_full + ACC_E_OFF);
                    if (DBG) { dump(va, code_s); dump(vb, code_s + 16); dump(vc, code_s + 32); }
                    if (GS == 3) {                      // a group = three consecutive columns = one FMNMX3
                        auto colv = [&](int c) {
                            return __uint_as_float(c < 16 ? va[c & 15] : (c < 32 ? vb[c & 15] : vc[c & 15]));
                        };
#pragma unroll
                        for (int g = 0; g < 16; g++) gv[g] = min3f(colv(3 * g), colv(3 * g + 1), colv(3 * g + 2));
                    } else {
                        groups(va, gv);
                        groups(vb, gv + GPC);
                        groups(vc, gv + 2 * GPC);
                    }
                    process(gv, (uint32_t)(2 * s + h));
                    continue;
                }
                tmem_ld16_async(tsrc, va);
                tmem_ld16_async(tsrc + 16, vb);
                tmem_wait_ld16(va);
                tmem_wait_ld16(vb);
                if (DBG) { dump(va, code_s); dump(vb, code_s + 16); }
                groups(va, gv);
                if (BN == 112) {                        // 56 columns per thread: 16 + 16 + 16 + 8
                    uint32_t vc[8];
                    tmem_ld16_async(tsrc + 32, va);
                    tmem_ld8_async(tsrc + 48, vc);
                    groups(vb, gv + GPC);
                    tmem_wait_ld16(va);
                    tmem_wait_ld8(vc);
                    tc_fence_before();                  // this half of the slot is read: hand it back to the MMA warp
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acc_full + ACC_E_OFF);
                    if (DBG) {
                        dump(va, code_s + 32);
#pragma unroll
                        for (int jj = 0; jj < 8; jj++)
                            if (pair * TC_PR + lr < p.n && code_s + 48 + jj < p.k)
                                p.dbg_scores[(pair * TC_PR + lr) * p.k + code_s + 48 + jj] =
                                    __uint_as_float(vc[jj]) * (F16 ? ldexpf(1.f, -2 * p.hdr->g_exp) : 1.f);
                    }
                    groups(va, gv + 2 * GPC);
                    gv[12] = min4(vc[0], vc[1], vc[2], vc[3]);
                    gv[13] = min4(vc[4], vc[5], vc[6], vc[7]);
                } else {                                // BN == 64: 32 columns per thread
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acc_full + ACC_E_OFF);
                    groups(vb, gv + GPC);
                }
                process(gv, (uint32_t)(2 * s + h));
              }
            }
            const int cnt = (int)(cnt_off / LIST_STRIDE);
            PROF(4);
            // ---- hand the surviving groups of this pass to the re-rank warps ----
            const int qb = pi & 1;
            mbar_wait(Q_EMPTY(qb), ((pi >> 1) & 1) ^ 1);     // the pass that used this queue slot is finished
            PROF(5);       // prof[24+5]: drain wait Q_EMPTY
            uint8_t status = ROW_OK;
            if (pair * TC_PR + lr >= p.n)
                status = ROW_SKIP;
            else if (!(band >= 0.f))
                status = ROW_FALLBACK;
            // the two threads of a row exchange their running minima: both filter with the row's threshold
            sXchg[(qb * 2 + h) * TC_PR + lr] = m;             // double buffered: one barrier per pass is enough
            named_bar_sync(3 + (dw & 7), 64);                 // only the two warps that share these 32 rows meet
            m = fminf(m, sXchg[(qb * 2 + (h ^ 1)) * TC_PR + lr]);
            // Surviving groups of the row's list, filtered with the final threshold.  Queue slots are allocated per
            // WARP (prefix sum over the lanes, one shared-memory atomic), not per entry: a thousand same-address
            // atomics per pass serialise in the shared-memory atomic unit.
            thr = m + p.gamma2 * fabsf(m) + band;
            uint32_t mk[L::LCAP], gb[L::LCAP], c = 0;
#pragma unroll
            for (int e = 0; e < L::LCAP; e++) {
                mk[e] = 0;
                gb[e] = 0;
                if (status == ROW_OK && e < cnt) {
                    const uint32_t es = ld_shared_u32(lS + e * LIST_STRIDE);
                    if (__uint_as_float(es) <= thr) {
                        const uint32_t ep = ld_shared_u32(lS + LIST_PLANE + e * LIST_STRIDE);
                        uint32_t gm = ep & 0xFFFFu;
                        const uint32_t grp0 = (ep >> 16) * NGH;      // first group (code / GS) of the half job
                        if ((int)(grp0 + NGH) * GS > p.k) {          // the chunk that holds the end of the codebook
                            uint32_t keep = 0;
                            for (int bit = 0; bit < NGH; bit++)
                                if ((int)(grp0 + (NGH - 1 - bit)) * GS < p.k) keep |= 1u << bit;
                            gm &= keep;
                        }
                        mk[e] = gm;
                        gb[e] = grp0;
                        c += __popc(gm);
                    }
                }
            }
            uint32_t incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
            if (total) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(&sQCount[qb], total);
                base = __shfl_sync(0xffffffffu, base, 0);
                uint32_t pos = base + incl - c;
                const uint32_t qaddr = smem_u32(sQueue + qb * TC_QCAP);
#pragma unroll
                for (int e = 0; e < L::LCAP; e++) {
                    uint32_t gm = mk[e];
                    while (gm) {
                        const int bit = __ffs(gm) - 1;
                        gm &= gm - 1;
                        if (pos < (uint32_t)TC_QCAP)
                            st_shared_u32(qaddr + pos * 4, ((uint32_t)lr << 16) | (gb[e] + (NGH - 1 - bit)));
                        else
                            status = ROW_FALLBACK;
                        pos++;
                    }
                }
            }
            sFlag[(qb * 2 + h) * TC_PR + lr] = status;
            __syncwarp();
            if (lane == 0) mbar_arrive(Q_FULL(qb));
            PROF(6);       // prof[24+6]: drain queue push
        }
        PROF_FLUSH(24, tid == TC_DRAIN_WARP0 * 32);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == TC_MISC_WARP0 + 2) {
        tc_fence_after();
        tmem_dealloc(*(volatile uint32_t *)sTmem, 512);
    }
}

// ------------------------------------------------------------------------------------------------
// Exact scan for the rows the tensor-core kernel could not decide (reference arithmetic, lowest index wins ties).
// A work item is (fallback row, block of FB_CODES codewords) handled by one warp, so that even a single
// undecided row (typical: a row whose nearest codewords are hundreds of collapsed dead entries) is spread over
// k / FB_CODES warps instead of costing one warp a serial scan of the whole codebook.  The partial minima meet in
// a 64-bit atomicMin on (distance bits, code): distances are sums of squares from +0, so the unsigned order of
// the bits is the numeric order, NaN / inf sort above FLT_MAX and never win, ties go to the lower code.
constexpr int FB_CODES = 256;
constexpr unsigned long long FB_INIT = ((unsigned long long)0x7f7fffffu) << 32;      // (FLT_MAX, code 0)

__global__ void __launch_bounds__(128)
k_assign_fallback_scan(const float *__restrict__ x, const int64_t *__restrict__ rows, int d,
                       const float *__restrict__ cbp, int dpc, int kq, int k, const TcHeader *hdr,
                       const unsigned int *__restrict__ fb_list, unsigned long long *__restrict__ fb_best)
{
    // codes from the float4-column-major copy (TcParams::cbp): lane l takes the codes blk * 256 + l + 32 t, t = 0..7,
    // eight independent distance chains whose loads are in flight together (each chain in column order, as the
    // reference).  The row-major version walked them one after the other behind dependent loads: 268 us for the
    // 14 k rows the trace-normalised shape codebook leaves undecided, now a fraction of that.
    __shared__ __align__(16) float xs[4][C3VQ_MAX_DIM + 4];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const unsigned int count = hdr->fb_count;
    const unsigned int nblk = (unsigned int)((k + FB_CODES - 1) / FB_CODES);
    const unsigned long long items = (unsigned long long)count * nblk;
    const unsigned int nwarps = gridDim.x * 4;
    const float4 *cb4 = reinterpret_cast<const float4 *>(cbp);
    for (unsigned long long w = blockIdx.x * 4 + wib; w < items; w += nwarps) {
        const unsigned int it = (unsigned int)(w / nblk), blk = (unsigned int)(w % nblk);
        const int64_t gi = fb_list[it];
        const int64_t src = rows ? rows[gi] : gi;
        __syncwarp();
        for (int c = lane; c < dpc; c += 32) xs[wib][c] = c < d ? x[src * d + c] : 0.f;
        __syncwarp();
        const int j0 = (int)blk * FB_CODES + lane;
        float a[8];
#pragma unroll
        for (int t = 0; t < 8; t++) a[t] = 0.f;
        for (int c4 = 0; c4 < dpc / 4; c4++) {
            const float4 xv = reinterpret_cast<const float4 *>(xs[wib])[c4];
            float4 cv[8];
#pragma unroll
            for (int t = 0; t < 8; t++)
                cv[t] = j0 + 32 * t < k ? __ldg(cb4 + (size_t)c4 * kq + j0 + 32 * t) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int t = 0; t < 8; t++) {
                float df;
                if (4 * c4 + 0 < d) { df = xv.x - cv[t].x; a[t] = fmaf(df, df, a[t]); }
                if (4 * c4 + 1 < d) { df = xv.y - cv[t].y; a[t] = fmaf(df, df, a[t]); }
                if (4 * c4 + 2 < d) { df = xv.z - cv[t].z; a[t] = fmaf(df, df, a[t]); }
                if (4 * c4 + 3 < d) { df = xv.w - cv[t].w; a[t] = fmaf(df, df, a[t]); }
            }
        }
        unsigned long long key = ~0ull;
#pragma unroll
        for (int t = 0; t < 8; t++) {
            const int j = j0 + 32 * t;
            const unsigned long long kj = ((unsigned long long)__float_as_uint(a[t]) << 32) | (unsigned int)j;
            if (j < k && kj < key) key = kj;
        }
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long ok = __shfl_xor_sync(0xffffffffu, key, o);
            key = ok < key ? ok : key;
        }
        if (lane == 0 && key < FB_INIT) atomicMin(&fb_best[it], key);
    }
}

// outputs of the fallback rows (and, in the fused mode, their weighted row into the centroid sums)
__global__ void __launch_bounds__(128)
k_assign_fallback_finish(const float *__restrict__ x, const float *__restrict__ w, const int64_t *__restrict__ rows,
                         int d, const TcHeader *hdr, const unsigned int *__restrict__ fb_list,
                         const unsigned long long *__restrict__ fb_best, float *__restrict__ out_dist,
                         int64_t *__restrict__ out_idx, float *__restrict__ acc, int stride, int *__restrict__ prev)
{
    const int lane = threadIdx.x & 31;
    const unsigned int count = hdr->fb_count;
    const unsigned int nwarps = gridDim.x * 4;
    for (unsigned int it = blockIdx.x * 4 + (threadIdx.x >> 5); it < count; it += nwarps) {
        const int64_t gi = fb_list[it];
        const unsigned long long best = fb_best[it];
        const int bj = (int)(best & 0xFFFFFFFFu);            // nothing below FLT_MAX: index 0, distance FLT_MAX
        if (lane == 0) {
            if (out_idx) out_idx[gi] = bj;
            if (out_dist) out_dist[gi] = __uint_as_float((unsigned int)(best >> 32));
            if (prev) prev[gi] = bj;
        }
        if (acc) {
            const int64_t src = rows ? rows[gi] : gi;
            const float wv = w ? w[src] : 1.f;
            float *ap = acc + (size_t)bj * stride;
            for (int c = lane; c <= d; c += 32) atomicAdd(ap + c, c < d ? __fmul_rn(x[src * d + c], wv) : wv);
        }
    }
}

// ------------------------------------------------------------------------------------------------
static inline size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

struct TcLayout {
    size_t off_mu, off_bimg, off_cbp, off_fb, off_fbbest, off_prev, off_reg, off_parts, total;
    int dp, nc, dpc, kq, k_pad, bn;
};

// Screening arithmetic: fp16 operands (kind::f16) by default, C3VQ_SCREEN=tf32 selects kind::tf32 (developer A/B knob;
// the results are identical, both are followed by the exact re-rank).
static bool screen_f16()
{
    const char *e = getenv("C3VQ_SCREEN");               // read per call: tests and A/B tools flip it in-process
    return !(e && (e[0] == 't' || e[0] == 'T'));
}

// C3VQ_SEEDS=0: developer A/B knob, the running minima start unseeded (results are identical either way)
static bool seeds_enabled()
{
    const char *e = getenv("C3VQ_SEEDS");
    return !(e && e[0] == '0');
}

static TcLayout tc_layout(int64_t n, int d, int k, bool f16)
{
    TcLayout l;
    l.dp = tc_dp(d, f16);
    l.nc = tc_nc(d, f16);
    l.dpc = (d + 3) & ~3;
    l.kq = ((k > 0 ? k : 1) + 7) & ~7;
    l.bn = f16 ? tc_bn_f16(l.dp) : tc_bn(l.dp);
    l.k_pad = (k + l.bn - 1) / l.bn * l.bn;
    if (l.k_pad == 0) l.k_pad = l.bn;
    size_t o = al256(sizeof(TcHeader));
    l.off_mu = o;
    o += al256(80 * sizeof(float));
    l.off_bimg = o;
    o += al256((size_t)l.k_pad * l.dp * sizeof(float));      // sized for tf32 elements; the fp16 image needs half
    l.off_cbp = o;
    o += al256((size_t)l.kq * l.dpc * sizeof(float));
    l.off_fb = o;
    o += al256((size_t)(n > 0 ? n : 1) * sizeof(unsigned int));
    l.off_fbbest = o;
    o += al256((size_t)(n > 0 ? n : 1) * sizeof(unsigned long long));
    l.off_prev = o;
    o += al256((size_t)(n > 0 ? n : 1) * sizeof(int));
    l.off_reg = o;
    o += al256((size_t)(k > 0 ? k : 1));
    l.off_parts = o;
    if (n >= TC_ACC_PARTS_MIN_ROWS) o += al256((size_t)TC_ACC_PARTS * (size_t)(k > 0 ? k : 1) * acc_stride(d) * sizeof(float));
    l.total = o;
    return l;
}

size_t tc_workspace_bytes(int64_t n, int d, int k)
{
    if (n < 0 || d < 1 || d > C3VQ_MAX_DIM || k < 0) return 256;
    const size_t a = tc_layout(n, d, k, false).total, b = tc_layout(n, d, k, true).total;
    return a > b ? a : b;
}

template <int DP, int DCONST, bool F16, bool DBG = false>
static int launch_tc_t(const TcParams &p, int grid, cudaStream_t s)
{
    if (!DBG && DCONST == 0 && p.dbg_scores != nullptr) return launch_tc_t<DP, 0, F16, true>(p, grid, s);
    using L = TcSmem<DP, F16>;
    TcParams q = p;
    q.nstages = L::stages();
    if (q.nstages < 1) {
        set_error("tensor-core kernel: no room for a codebook stage at DP=%d", DP);
        return C3VQ_EINVAL;
    }
    const size_t smem = L::bytes(q.nstages);
    C3VQ_CUDA(cudaFuncSetAttribute(k_assign_tc<DP, DCONST, DBG, F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_assign_tc<DP, DCONST, DBG, F16><<<grid, TC_THREADS, smem, s>>>(q);
    C3VQ_LAUNCH_CHECK("k_assign_tc");
    return C3VQ_OK;
}

// acc[i] += pairwise sum over the partial tables (one thread per table element, coalesced over i)
__global__ void __launch_bounds__(256) k_acc_fold(const float *__restrict__ parts, size_t elems, float *__restrict__ acc)
{
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= elems) return;
    float v[TC_ACC_PARTS];
#pragma unroll
    for (int c = 0; c < TC_ACC_PARTS; c++) v[c] = parts[(size_t)c * elems + i];
#pragma unroll
    for (int st = 1; st < TC_ACC_PARTS; st <<= 1)
#pragma unroll
        for (int c = 0; c < TC_ACC_PARTS; c += 2 * st) v[c] = __fadd_rn(v[c], v[c + st]);
    acc[i] = __fadd_rn(acc[i], v[0]);
}

int launch_assign_tc(const float *x, const float *w, const int64_t *rows, int64_t n, int d, const float *codebook,
                     int k, float *out_dist, int64_t *out_idx, float *acc, float *dbg_scores, float *dbg_norm2,
                     float *dbg_band, void *ws, size_t ws_bytes, cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    if (k == 0) return launch_assign_exact(x, rows, n, d, codebook, k, out_dist, out_idx, s);   // defaults only
    C3VQ_CHECK_ARG(n < 0xffffffffLL, "n=%lld too large for the fallback list", (long long)n);
    const bool f16 = screen_f16();
    const TcLayout l = tc_layout(n, d, k, f16);
    if (ws == nullptr || ws_bytes < l.total) {
        set_error("workspace too small: %zu < %zu", ws_bytes, l.total);
        return C3VQ_EWORKSPACE;
    }
    C3VQ_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "workspace must be 256-byte aligned");
    char *base = (char *)ws;
    TcHeader *hdr = (TcHeader *)base;
    float *mu = (float *)(base + l.off_mu);
    float *bimg = (float *)(base + l.off_bimg);
    float *cbp = (float *)(base + l.off_cbp);
    unsigned int *fb = (unsigned int *)(base + l.off_fb);

    C3VQ_CUDA(cudaMemsetAsync(hdr, 0, sizeof(TcHeader), s));
    k_tc_prep_hist<<<(k + TC_PREP_CODES - 1) / TC_PREP_CODES, 256, 0, s>>>(codebook, k, d, x, rows, n, hdr);
    C3VQ_LAUNCH_CHECK("k_tc_prep_hist");
    k_tc_prep_mean<<<(k + TC_PREP_CODES - 1) / TC_PREP_CODES, 256, 0, s>>>(codebook, k, d, hdr);
    C3VQ_LAUNCH_CHECK("k_tc_prep_mean");
    k_tc_prep_image<<<(l.k_pad + 127) / 128, 128, 0, s>>>(codebook, k, l.k_pad, d, l.dp, l.nc, l.dpc, l.bn, f16 ? 1 : 0,
                                                        hdr, mu, bimg, cbp, l.kq, (uint8_t *)(base + l.off_reg));
    C3VQ_LAUNCH_CHECK("k_tc_prep_image");

    TcParams p{};
    p.x = x; p.w = w; p.rows = rows; p.n = n; p.d = d; p.k = k;
    p.nsteps = l.k_pad / l.bn;
    p.npairs = (n + TC_PR - 1) / TC_PR;
    p.bimg = bimg; p.cbp = cbp; p.dpc = l.dpc; p.kq = l.kq; p.mu = mu; p.hdr = hdr;
    p.out_dist = out_dist; p.out_idx = out_idx; p.acc = acc; p.acc_stride = acc_stride(d);
    p.acc_parts = 1; p.acc_part_elems = 0;
    float *parts = nullptr;
    const size_t acc_elems = (size_t)k * acc_stride(d);
    if (acc != nullptr && n >= TC_ACC_PARTS_MIN_ROWS) {
        parts = (float *)(base + l.off_parts);
        C3VQ_CUDA(cudaMemsetAsync(parts, 0, (size_t)TC_ACC_PARTS * acc_elems * sizeof(float), s));
        p.acc = parts; p.acc_parts = TC_ACC_PARTS; p.acc_part_elems = acc_elems;
    }
    p.fb_list = fb;
    p.fb_best = (unsigned long long *)(base + l.off_fbbest);
    p.dbg_scores = dbg_scores; p.dbg_norm2 = dbg_norm2; p.dbg_band = dbg_band;
#ifndef C3VQ_SEEDS_MIN_DP
#define C3VQ_SEEDS_MIN_DP 0
#endif
    p.prev = (rows == nullptr && dbg_scores == nullptr && l.dp > C3VQ_SEEDS_MIN_DP && seeds_enabled()) ? (int *)(base + l.off_prev) : nullptr;
    p.reg = (const uint8_t *)(base + l.off_reg);
    p.gamma2 = 2.1f * (float)(d + 3) * 5.9604644775390625e-8f;      // 2.1 (d+3) 2^-24
    const int sms = sm_count();
    const int grid = (int)(p.npairs < sms ? p.npairs : sms);

    int rc;
#if defined(C3VQ_TC_SINGLE_D7)   // developer build: one instantiation only (fast compile, SASS inspection)
    rc = launch_tc_t<16, 7, true>(p, grid, s);
#elif defined(C3VQ_TC_SINGLE)
    rc = launch_tc_t<48, 45, true>(p, grid, s);
#else
    if (f16) {
        switch (dbg_scores ? -1 : d) {  // the debug build of the kernel exists for the generic widths only
            case 6: rc = launch_tc_t<16, 6, true>(p, grid, s); break;
            case 7: rc = launch_tc_t<16, 7, true>(p, grid, s); break;
            case 45: rc = launch_tc_t<48, 45, true>(p, grid, s); break;
            case 48: rc = launch_tc_t<64, 48, true>(p, grid, s); break;
            default: {
                // generic widths: the image is laid out for DP = tc_dp(d, true); kernels exist for every multiple of 16
                switch (l.dp) {
                    case 16: rc = launch_tc_t<16, 0, true>(p, grid, s); break;
                    case 32: rc = launch_tc_t<32, 0, true>(p, grid, s); break;
                    case 48: rc = launch_tc_t<48, 0, true>(p, grid, s); break;
                    case 64: rc = launch_tc_t<64, 0, true>(p, grid, s); break;
                    default: rc = launch_tc_t<80, 0, true>(p, grid, s); break;
                }
            }
        }
    } else {
        switch (dbg_scores ? -1 : d) {
            case 6: rc = launch_tc_t<8, 6, false>(p, grid, s); break;
            case 7: rc = launch_tc_t<16, 7, false>(p, grid, s); break;
            case 45: rc = launch_tc_t<48, 45, false>(p, grid, s); break;
            case 48: rc = launch_tc_t<56, 48, false>(p, grid, s); break;
            default: {
                // generic widths: the image is laid out for DP = tc_dp(d); kernels exist for every multiple of 8
                switch (l.dp) {
                    case 8: rc = launch_tc_t<8, 0, false>(p, grid, s); break;
                    case 16: rc = launch_tc_t<16, 0, false>(p, grid, s); break;
                    case 24: rc = launch_tc_t<24, 0, false>(p, grid, s); break;
                    case 32: rc = launch_tc_t<32, 0, false>(p, grid, s); break;
                    case 40: rc = launch_tc_t<40, 0, false>(p, grid, s); break;
                    case 48: rc = launch_tc_t<48, 0, false>(p, grid, s); break;
                    case 56: rc = launch_tc_t<56, 0, false>(p, grid, s); break;
                    case 64: rc = launch_tc_t<64, 0, false>(p, grid, s); break;
                    default: rc = launch_tc_t<72, 0, false>(p, grid, s); break;
                }
            }
        }
    }
#endif
    if (rc) return rc;
    if (dbg_scores == nullptr) {
        k_assign_fallback_scan<<<sms * 4, 128, 0, s>>>(x, rows, d, cbp, l.dpc, l.kq, k, hdr, fb, p.fb_best);
        C3VQ_LAUNCH_CHECK("k_assign_fallback_scan");
        k_assign_fallback_finish<<<sms, 128, 0, s>>>(x, w, rows, d, hdr, fb, p.fb_best, out_dist, out_idx, acc,
                                                     acc_stride(d), p.prev);
        C3VQ_LAUNCH_CHECK("k_assign_fallback_finish");
    }
    if (parts != nullptr) {
        k_acc_fold<<<(unsigned)((acc_elems + 255) / 256), 256, 0, s>>>(parts, acc_elems, acc);
        C3VQ_LAUNCH_CHECK("k_acc_fold");
    }
    return C3VQ_OK;
}

int tc_stats(void *ws, int64_t out[3], cudaStream_t s)
{
    TcHeader h;
    C3VQ_CUDA(cudaMemcpyAsync(&h, ws, sizeof(TcHeader), cudaMemcpyDeviceToHost, s));
    C3VQ_CUDA(cudaStreamSynchronize(s));
    out[0] = (int64_t)h.rows;
    out[1] = (int64_t)h.cands;
    out[2] = (int64_t)h.fb_rows;
    return C3VQ_OK;
}

// ------------------------------------------------------------------------------------------------
// c3vq_assign_static: the same search with its scratch (header, codebook image, fallback lists) in a fixed device
// array instead of a caller-provided workspace, for callers that cannot allocate device memory (CUDA graph capture,
// serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library holds the array until exit.
// g_static_scratch names the array in DEVICE code only; on the host the symbol is a stand-in, so the launcher asks
// the runtime for the device address (once per device, with that device current) and hands that to the kernels.
__device__ __align__(256) unsigned char g_static_scratch[C3VQ_STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
static void *g_static_ptr[STATIC_MAX_DEVICES];           // device address of g_static_scratch, per device ordinal
static std::mutex g_static_mutex;

static int static_scratch_ptr(void **out)
{
    int dev = -1;
    C3VQ_CUDA(cudaGetDevice(&dev));
    C3VQ_CHECK_ARG(dev >= 0 && dev < STATIC_MAX_DEVICES, "device ordinal %d outside [0, %d)", dev, STATIC_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (g_static_ptr[dev] == nullptr) {
        void *ptr = nullptr;
        C3VQ_CUDA(cudaGetSymbolAddress(&ptr, g_static_scratch));
        C3VQ_CHECK_ARG(ptr != nullptr && (reinterpret_cast<uintptr_t>(ptr) & 255) == 0,
                       "static scratch address %p is null or misaligned", ptr);
        g_static_ptr[dev] = ptr;
    }
    *out = g_static_ptr[dev];
    return C3VQ_OK;
}

int launch_assign_tc_static(const float *x, const int64_t *rows, int64_t n, int d, const float *codebook, int k,
                            float *out_dist, int64_t *out_idx, cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    const size_t need = tc_workspace_bytes(n, d, k);
    if (need > (size_t)C3VQ_STATIC_SCRATCH_BYTES) {
        set_error("c3vq_assign_static: n=%lld, d=%d, k=%d needs %zu bytes of scratch, the static array holds %zu "
                  "(use c3vq_assign with a caller-provided workspace)", (long long)n, d, k, need,
                  (size_t)C3VQ_STATIC_SCRATCH_BYTES);
        return C3VQ_EWORKSPACE;
    }
    void *ws = nullptr;
    const int rc = static_scratch_ptr(&ws);
    if (rc) return rc;
    // nothing in the array has to survive or start at zero beyond what launch_assign_tc clears itself on every call
    // (the header with its counters; lists and images are rewritten before they are read)
    return launch_assign_tc(x, nullptr, rows, n, d, codebook, k, out_dist, out_idx, nullptr, nullptr, nullptr, nullptr,
                            ws, (size_t)C3VQ_STATIC_SCRATCH_BYTES, s);
}

}  // namespace c3vq
