// assign_tc.cu — tensor-core nearest-codeword search (+ fused weighted centroid sums) for sm_100a.
//
// Replaces distance_reduce (submodules/weighted_distance/weighted_distance.cu:20-44) and, when `acc`
// is given, the two scatter sums of VectorQuantize.update (compression/vq.py:31, :33), reading every
// feature row from HBM exactly once per iteration.
//
// Method (results are bit-identical to the reference's fp32 argmin):
//   1. SCREEN  s_ij = |c'_j|^2 - 2 x'_i . c'_j  with tcgen05.mma kind::tf32 into TMEM, where
//      x' = x - mu, c' = c - mu (mu = codebook mean; a distance-preserving shift that shrinks the
//      cancellation error).  Operands are rounded to TF32 by us (cvt.rna), |c'|^2 rides along as 2-3
//      extra K columns (hi/mid/lo TF32 pieces against a 1.0 in the A operand), so the epilogue is a
//      pure min: one FMNMX3 per two scores plus one compare per 32.
//   2. BAND    every code whose score is within band_i of the row minimum is a candidate, where
//      band_i = 2*E_i + slack bounds the worst-case screening error (TF32 input rounding 2^-11 per
//      operand, tensor-core accumulation, norm split) plus the reference chain's own rounding; the
//      fp32 argmin of the reference is therefore always among the candidates.
//   3. RE-RANK candidates with the reference arithmetic (in-order fma chain, first index wins ties).
//      Rows whose candidate list overflows, or whose magnitudes break the error model, go to an exact
//      warp-per-row scan (k_assign_fallback).
//
// Kernel organisation (persistent, one CTA per SM, 640 threads, warp specialised):
//   warp 0      bulk-copy (TMA engine, cp.async.bulk) producer: codebook image stages -> smem ring
//   warp 1      MMA issuer: per stage 2 x KSTEPS tcgen05.mma (M=128, N=256, K=8), commits to mbarriers
//   warp 2      TMEM allocator (512 columns = two 128x256 fp32 accumulators, one per row tile)
//   warps 4-11  epilogue: tcgen05.ld 32 columns at a time, running min + candidate list per row
//               (one thread owns one row for all K codes), then re-rank + outputs + red.global.add.v4
//   warps 12-19 feature producers: coalesced prefetch of the next 256 rows into registers, centre,
//               round to TF32, store into the UMMA canonical K-major layout, row norms -> error band
// While the tensor core fills the accumulator of row tile 0 the epilogue drains row tile 1 and vice
// versa, so each codebook stage (256 codes) is used for 256 rows: 32 B/cycle/SM of L2 traffic.
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
constexpr int TC_BN = 256;            // codes per stage / per MMA
constexpr int TC_LCAP = 16;           // candidate list capacity per row
constexpr int TC_THREADS = 640;
constexpr int TC_EPI_WARP0 = 4;
constexpr int TC_PROD_WARP0 = 12;
constexpr uint32_t TC_LBO_A = TC_BM * 16 + 16;   // skewed by 16 B: conflict-free scattered stores
constexpr uint32_t TC_LBO_B = TC_BN * 16;
constexpr uint32_t TC_SBO = 128;
constexpr float TC_BIG = 8.507059173023462e37f;  // 2^126: score of dead / padding codes
constexpr float TC_HUGE2 = 1.2676506002282294e30f;  // 2^100: squared-norm limit of the error model
constexpr int TC_SMEM_LIMIT = 227 * 1024;

// Workspace header (device memory, zeroed per call).
struct TcHeader {
    unsigned long long rows;        // stats: rows processed by the tensor-core kernel
    unsigned long long cands;       // stats: candidates that survived pruning (re-rank evaluations)
    unsigned long long fb_rows;     // stats: rows sent to the exact fallback
    unsigned int fb_count;          // fallback list length
    unsigned int force_fallback;    // set by prep when a finite code breaks the magnitude limit
    int cmax_bits;                  // max |c'_j| over finite codes, float bits (non-negative)
    int live;                       // number of finite codes
    float mu_sum[72];               // column sums over finite codes
    unsigned long long prof[32];    // developer build (C3VQ_TC_PROF): cycle counters per role/phase
};

__host__ __device__ inline int tc_dp(int d) { return (d + 2 + 7) & ~7; }          // padded K extent
__host__ __device__ inline int tc_nc(int d) { int r = tc_dp(d) - d; return r < 3 ? r : 3; }  // norm columns

struct TcParams {
    const float *x;
    const float *w;
    const int64_t *rows;
    int64_t n;
    int d, k, nsteps, nstages;
    int64_t npairs;
    const float *bimg;      // [k_pad/256][dp/4][256][4] tf32 image of -2c' and the norm pieces
    const float *cbp;       // [k][dpc] raw codebook, zero padded to a multiple of 4 columns
    int dpc;
    const float *mu;        // [d] centre
    TcHeader *hdr;
    float *out_dist;
    int64_t *out_idx;
    float *acc;
    int acc_stride;
    unsigned int *fb_list;
    float *dbg_scores, *dbg_norm2, *dbg_band;
    float gamma2;           // relative slack for the reference chain's own rounding
};

// ------------------------------------------------------------------------------------------------
// prep 1: column sums over finite codes (for the centre mu)
__global__ void __launch_bounds__(256) k_tc_prep_mean(const float *__restrict__ cb, int k, int d, TcHeader *hdr)
{
    // warp per code (lanes over columns, coalesced row reads), 128 codes per block
    __shared__ float ssum[72];
    __shared__ int slive;
    if (threadIdx.x < 72) ssum[threadIdx.x] = 0.f;
    if (threadIdx.x == 0) slive = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    float s0 = 0.f, s1 = 0.f;
    int live = 0;
    const int j0 = blockIdx.x * 128;
    for (int j = j0 + wib; j < min(k, j0 + 128); j += 8) {
        const float *row = cb + (size_t)j * d;
        const float v0 = lane < d ? row[lane] : 0.f;
        const float v1 = lane + 32 < d ? row[lane + 32] : 0.f;
        const bool fin = __all_sync(0xffffffffu, isfinite(v0) && isfinite(v1));
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
                                TcHeader *hdr, float *__restrict__ mu_out, float *__restrict__ bimg,
                                float *__restrict__ cbp)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k_pad) return;
    const int live = hdr->live;
    const float inv = live > 0 ? 1.f / (float)live : 0.f;
    float *img = bimg + (size_t)(j >> 8) * (TC_BN * dp) + (size_t)(j & 255) * 4;   // + chunk*1024 + (kk&3)
    bool fin = j < k;
    double n2 = 0.0;
    if (j < k) {
        const float *row = cb + (size_t)j * d;
        for (int c = 0; c < d; c++) {
            const float v = row[c];
            fin = fin && isfinite(v);
            const float m = hdr->mu_sum[c] * inv;
            if (j == 0) mu_out[c] = m;
            const float cc = v - m;
            n2 += (double)cc * (double)cc;
            cbp[(size_t)j * dpc + c] = v;
        }
        for (int c = d; c < dpc; c++) cbp[(size_t)j * dpc + c] = 0.f;
    }
    const float nf = (float)n2;
    if (fin && !(nf < TC_HUGE2)) {      // finite code outside the error model: everything goes exact
        atomicOr(&hdr->force_fallback, 1u);
    }
    if (fin) atomicMax(&hdr->cmax_bits, __float_as_int(sqrtf(nf)));
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
        img[(size_t)(c >> 2) * (TC_BN * 4) + (c & 3)] = v;
    }
}

// ------------------------------------------------------------------------------------------------
template <int DP>
struct TcSmem {
    static constexpr int KSTEPS = DP / 8;
    static constexpr uint32_t BSTAGE = TC_BN * DP * 4;
    static constexpr uint32_t ATILE = (DP / 4) * TC_LBO_A;
    static constexpr bool RAW = DP <= 48;                 // raw fp32 rows kept in smem for the re-rank
    static constexpr int RS_MAX = DP - 1;                 // row stride (floats), odd
    static constexpr uint32_t RAW_BYTES = RAW ? (uint32_t)TC_PR * RS_MAX * 4 : 0;
    static constexpr uint32_t LIST_BYTES = TC_PR * TC_LCAP * 4;
    static constexpr uint32_t FIXED = 2 * ATILE + RAW_BYTES + LIST_BYTES + TC_PR * 4 /*rowinfo*/ + 80 * 4 /*mu*/ +
                                      256 /*barriers, tmem ptr*/ + 128 /*alignment*/;
    static int stages()
    {
        int s = (TC_SMEM_LIMIT - (int)FIXED) / (int)BSTAGE;
        return s > 4 ? 4 : s;
    }
    static size_t bytes(int s) { return (size_t)FIXED + (size_t)s * BSTAGE; }
};

__device__ __forceinline__ float min32(const float (&v)[32])
{
    float m0 = fminf(fminf(v[0], v[1]), v[2]);
    float m1 = fminf(fminf(v[3], v[4]), v[5]);
    float m2 = fminf(fminf(v[6], v[7]), v[8]);
    float m3 = fminf(fminf(v[9], v[10]), v[11]);
    m0 = fminf(fminf(m0, v[12]), v[13]);
    m1 = fminf(fminf(m1, v[14]), v[15]);
    m2 = fminf(fminf(m2, v[16]), v[17]);
    m3 = fminf(fminf(m3, v[18]), v[19]);
    m0 = fminf(fminf(m0, v[20]), v[21]);
    m1 = fminf(fminf(m1, v[22]), v[23]);
    m2 = fminf(fminf(m2, v[24]), v[25]);
    m3 = fminf(fminf(m3, v[26]), v[27]);
    m0 = fminf(fminf(m0, v[28]), v[29]);
    m1 = fminf(fminf(m1, v[30]), v[31]);
    return fminf(fminf(m0, m1), fminf(m2, m3));
}

// score rounded DOWN to 16 significant bits, packed with the 16-bit code id
__device__ __forceinline__ uint32_t pack_entry(float s, uint32_t code)
{
    uint32_t b = __float_as_uint(s);
    if (b & 0x80000000u) b += 0xFFFFu;   // negative: grow the magnitude
    return (b & 0xFFFF0000u) | code;
}
__device__ __forceinline__ float entry_score(uint32_t e) { return __uint_as_float(e & 0xFFFF0000u); }

// DCONST > 0: feature width known at compile time (6, 7, 45, 48); DCONST == 0: runtime width p.d <= DP - 2.
template <int DP, int DCONST>
__global__ void __launch_bounds__(TC_THREADS, 1) k_assign_tc(const TcParams p)
{
    using L = TcSmem<DP>;
    constexpr int KSTEPS = L::KSTEPS;
    constexpr bool RAW = L::RAW;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);

    const int D = DCONST > 0 ? DCONST : p.d;
    const int RS = D | 1;                                      // raw row stride (odd -> conflict free)
    const int S = p.nstages;
    uint8_t *sB = smem;                                        // S * BSTAGE
    uint8_t *sA = sB + (size_t)S * L::BSTAGE;                  // 2 * ATILE
    float *sRaw = reinterpret_cast<float *>(sA + 2 * L::ATILE);
    uint32_t *sList = reinterpret_cast<uint32_t *>(reinterpret_cast<uint8_t *>(sRaw) + L::RAW_BYTES);
    float *sInfo = reinterpret_cast<float *>(sList + TC_PR * TC_LCAP);
    float *sMu = sInfo + TC_PR;
    uint64_t *sBar = reinterpret_cast<uint64_t *>(sMu + 80);
    // barrier indices
    const uint32_t bar0 = smem_u32(sBar);
    auto B_FULL = [&](int s) { return bar0 + 8u * s; };
    auto B_EMPTY = [&](int s) { return bar0 + 8u * (4 + s); };
    const uint32_t A_FULL = bar0 + 8u * 8, A_EMPTY = bar0 + 8u * 9;
    auto ACC_FULL = [&](int t) { return bar0 + 8u * (10 + t); };
    auto ACC_EMPTY = [&](int t) { return bar0 + 8u * (12 + t); };
    const uint32_t RAW_FULL = bar0 + 8u * 14, RAW_FREE = bar0 + 8u * 15;
    uint32_t *sTmem = reinterpret_cast<uint32_t *>(sBar + 16);

    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const int lane = tid & 31;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < 4; s++) {
            mbar_init(B_FULL(s), 1);
            mbar_init(B_EMPTY(s), 1);
        }
        mbar_init(A_FULL, TC_PR);
        mbar_init(A_EMPTY, 1);
        for (int t = 0; t < 2; t++) {
            mbar_init(ACC_FULL(t), 1);
            mbar_init(ACC_EMPTY(t), TC_BM);
        }
        mbar_init(RAW_FULL, TC_PR);
        mbar_init(RAW_FREE, TC_PR);
        fence_mbar_init();
    }
    if (warp == 2) {
        tmem_alloc(smem_u32(sTmem), 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *sTmem;

    const int nsteps = p.nsteps;

    if (warp == 0) {
        // ===================== codebook stage producer (TMA engine, 1-D bulk copies) =====================
        if (lane == 0) {
            uint32_t g = 0;
            PROF_DECL;
            for (int64_t pair = blockIdx.x; pair < p.npairs; pair += gridDim.x) {
                for (int s = 0; s < nsteps; s++, g++) {
                    const int stage = g % S;
                    const uint32_t u = g / S;
                    PROF(0);
                    mbar_wait_backoff(B_EMPTY(stage), (u & 1) ^ 1);
                    PROF(1);   // prof[1]: TMA wait B_EMPTY
                    mbar_arrive_expect_tx(B_FULL(stage), L::BSTAGE);
                    bulk_g2s(smem_u32(sB + (size_t)stage * L::BSTAGE), p.bimg + (size_t)s * (TC_BN * DP), L::BSTAGE,
                             B_FULL(stage));
                }
            }
            PROF_FLUSH(0, true);
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t IDESC = idesc_tf32(TC_BM, TC_BN);
            const uint32_t a_addr = smem_u32(sA);
            uint32_t g = 0, pi = 0;
            PROF_DECL;
            for (int64_t pair = blockIdx.x; pair < p.npairs; pair += gridDim.x, pi++) {
                PROF(0);
                mbar_wait(A_FULL, pi & 1);
                PROF(1);   // prof[8+1]: MMA wait A_FULL
                tc_fence_after();
                for (int s = 0; s < nsteps; s++, g++) {
                    const int stage = g % S;
                    const uint32_t u = g / S;
                    PROF(0);
                    mbar_wait(B_FULL(stage), u & 1);
                    PROF(2);   // prof[8+2]: MMA wait B_FULL
                    tc_fence_after();
                    const uint32_t b_addr = smem_u32(sB + (size_t)stage * L::BSTAGE);
#pragma unroll
                    for (int t = 0; t < 2; t++) {
                        PROF(0);
                        mbar_wait(ACC_EMPTY(t), (g & 1) ^ 1);
                        PROF(3);   // prof[8+3]: MMA wait ACC_EMPTY
                        tc_fence_after();
#pragma unroll
                        for (int ks = 0; ks < KSTEPS; ks++) {
                            const uint64_t ad = smem_desc(a_addr + t * L::ATILE + ks * 2 * TC_LBO_A, TC_LBO_A, TC_SBO);
                            const uint64_t bd = smem_desc(b_addr + ks * 2 * TC_LBO_B, TC_LBO_B, TC_SBO);
                            umma_tf32(tmem_base + t * TC_BN, ad, bd, IDESC, ks > 0 ? 1u : 0u);
                        }
                        umma_commit(ACC_FULL(t));
                    }
                    umma_commit(B_EMPTY(stage));
                }
                umma_commit(A_EMPTY);
            }
            PROF(0);       // prof[8+0]: MMA issue + everything else
            PROF_FLUSH(8, true);
        }
    } else if (warp >= TC_PROD_WARP0) {
        // ===================== feature producers (256 threads) =====================
        const int ptid = tid - TC_PROD_WARP0 * 32;
        const int NC = DP - D < 3 ? DP - D : 3;
        // constant padding columns of both A tiles: 1.0 against the norm pieces, 0 elsewhere
        {
            const int t = ptid >> 7, r = ptid & 127;
            uint8_t *base = sA + t * L::ATILE + (r >> 3) * 128 + (r & 7) * 16;
            for (int kk = D; kk < DP; kk++)
                *reinterpret_cast<float *>(base + (kk >> 2) * TC_LBO_A + (kk & 3) * 4) = (kk < D + NC) ? 1.0f : 0.f;
        }
        // Mapping: producer warp pw owns rows pw*32 .. pw*32+31 of the pass, lane <-> column (kk = lane + 32*h).
        // Loads are coalesced per row, every smem offset below is a per-thread base plus an immediate, and
        // with the 16-byte skew of TC_LBO_A the 32 lanes of a store hit 32 different banks.
        constexpr int NH = DP > 32 ? 2 : 1;                       // D <= 30 needs one column group
        const int pw = warp - TC_PROD_WARP0;
        float xr[32 * NH];
        auto prefetch = [&](int64_t pair) {
            const int64_t r0 = pair * TC_PR + pw * 32;
            if (p.rows == nullptr && r0 + 32 <= p.n) {
                // common case: 32 contiguous full rows -> one base pointer, immediate offsets
                const float *xp = p.x + r0 * D + lane;
#pragma unroll
                for (int r = 0; r < 32; r++) {
#pragma unroll
                    for (int h = 0; h < NH; h++) xr[r * NH + h] = (lane + 32 * h < D) ? xp[r * D + 32 * h] : 0.f;
                }
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
            }
        };
        float mu_l[NH];
        uint32_t koff[NH];
#pragma unroll
        for (int h = 0; h < NH; h++) {
            const int kk = lane + 32 * h;
            mu_l[h] = kk < D ? p.mu[kk] : 0.f;
            koff[h] = (kk >> 2) * TC_LBO_A + (kk & 3) * 4;
        }
        uint8_t *a_base = sA + (pw >> 2) * L::ATILE + (pw & 3) * 4 * 128;   // row pw*32 of the pass
        uint32_t pi = 0;
        PROF_DECL;
        // software pipeline with ONE prefetch call site: iteration -1 only prefetches the first pass
        for (int64_t pair = (int64_t)blockIdx.x - gridDim.x;; pair += gridDim.x) {
            if (pair >= 0) {
                PROF(0);   // prof[16+0]: producer prefetch issue
                mbar_wait_backoff(A_EMPTY, (pi & 1) ^ 1);  // previous pass' MMAs no longer read A
                PROF(1);   // prof[16+1]: producer wait A_EMPTY
#pragma unroll
                for (int r = 0; r < 32; r++) {
#pragma unroll
                    for (int h = 0; h < NH; h++) {
                        if (lane + 32 * h < D)
                            *reinterpret_cast<uint32_t *>(a_base + (r >> 3) * 128 + (r & 7) * 16 + koff[h]) =
                                to_tf32(xr[r * NH + h] - mu_l[h]);
                    }
                }
                fence_proxy_async_smem();
                mbar_arrive(A_FULL);
                PROF(2);   // prof[16+2]: producer convert + store A (includes waiting for the prefetched loads)
                if (RAW) {
                    mbar_wait_backoff(RAW_FREE, (pi & 1) ^ 1);   // epilogue finished re-ranking the previous pass
                    PROF(3);   // prof[16+3]: producer wait RAW_FREE
                    float *rp = sRaw + (pw * 32) * RS + lane;
#pragma unroll
                    for (int r = 0; r < 32; r++) {
#pragma unroll
                        for (int h = 0; h < NH; h++)
                            if (lane + 32 * h < D) rp[r * RS + 32 * h] = xr[r * NH + h];
                    }
                    mbar_arrive(RAW_FULL);
                    PROF(4);   // prof[16+4]: producer raw store
                }
                pi++;
            }
            if (pair + gridDim.x >= p.npairs) break;
            prefetch(pair + gridDim.x);
        }
        PROF_FLUSH(16, lane == 0 && pw == 0);
    } else if (warp >= TC_EPI_WARP0) {
        // ===================== epilogue: scan, candidate lists, re-rank, outputs =====================
        const int lr = tid - TC_EPI_WARP0 * 32;        // local row 0..255
        const int t = lr >> 7;
        const int q = warp & 3;                         // TMEM lane quarter this warp may access
        const uint32_t tsrc = tmem_base + ((uint32_t)(q * 32) << 16) + t * TC_BN;
        const bool need_dist = p.out_dist != nullptr;
        const float cmax = __int_as_float(p.hdr->cmax_bits);
        const bool force_fb = p.hdr->force_fallback != 0;
        unsigned long long st_rows = 0, st_cands = 0, st_fb = 0;
        uint32_t g = 0, pi = 0;
        PROF_DECL;
        for (int64_t pair = blockIdx.x; pair < p.npairs; pair += gridDim.x, pi++) {
            const int64_t gi = pair * TC_PR + lr;
            PROF(0);
            mbar_wait(A_FULL, pi & 1);
            PROF(1);       // prof[24+1]: epilogue wait A_FULL
            float band;
            {
                // row norm of the rounded, centred row (read back from the A tile) -> error band
                const int r = lr & 127;
                const uint8_t *base = sA + t * L::ATILE + (r >> 3) * 128 + (r & 7) * 16;
                float X = 0.f;
#pragma unroll
                for (int c4 = 0; c4 < DP / 4; c4++) {
                    const float4 xv = *reinterpret_cast<const float4 *>(base + c4 * TC_LBO_A);
                    if (4 * c4 + 0 < D) X = fmaf(xv.x, xv.x, X);
                    if (4 * c4 + 1 < D) X = fmaf(xv.y, xv.y, X);
                    if (4 * c4 + 2 < D) X = fmaf(xv.z, xv.z, X);
                    if (4 * c4 + 3 < D) X = fmaf(xv.w, xv.w, X);
                }
                // E bounds |score + X - true distance| for every code
                const float C1 = 1.002f * 0.001953125f + 7.62939453125e-6f;   // 2^-9 * 1.002 + 2^-17
                const float C2 = 3.814697265625e-6f + 4.76837158203125e-7f;   // 2^-18 + 2^-21
                const float E = C1 * sqrtf(X) * cmax + C2 * cmax * cmax;
                band = 1.01f * (2.f * E + p.gamma2 * X) + 3.0e-36f;
                if (!(X < TC_HUGE2 * 0.0625f) || force_fb || !(band < TC_BIG))
                    band = __int_as_float(0x7fc00000);                        // NaN: exact fallback
                if (gi >= p.n) band = -1.f;
                if (p.dbg_norm2 && gi < p.n) {
                    p.dbg_norm2[gi] = X;
                    p.dbg_band[gi] = band;
                }
            }
            float m = 3.0e38f;
            float thr = (band >= 0.f) ? __int_as_float(0x7f800000) : __int_as_float(0xff800000);  // +inf / -inf
            int cnt = 0;
            bool ovf = false;
            PROF(2);       // prof[24+2]: epilogue band
            for (int s = 0; s < nsteps; s++, g++) {
                PROF(4);   // prof[24+4]: epilogue drain (LDTM + min + lists)
                mbar_wait(ACC_FULL(t), g & 1);
                PROF(3);   // prof[24+3]: epilogue wait ACC_FULL
                tc_fence_after();
#pragma unroll 1
                for (int c = 0; c < TC_BN / 32; c++) {
                    float v[32];
                    tmem_ld32(tsrc + c * 32, v);
                    if (c == TC_BN / 32 - 1) {          // accumulator fully read: hand it back to the MMA warp
                        tc_fence_before();
                        mbar_arrive(ACC_EMPTY(t));
                    }
                    const int code0 = s * TC_BN + c * 32;
                    if (p.dbg_scores && gi < p.n) {
#pragma unroll
                        for (int j = 0; j < 32; j++)
                            if (code0 + j < p.k) p.dbg_scores[gi * p.k + code0 + j] = v[j];
                    }
                    const float cm = min32(v);
                    if (cm <= thr && cm < 0.5f * TC_BIG) {   // dead / padding codes (score 2^126) never lower the minimum
                        if (cm < m) {
                            m = cm;
                            thr = m + p.gamma2 * fabsf(m) + band;
                        }
                        uint32_t mask = 0;
#pragma unroll
                        for (int j = 0; j < 32; j++) mask |= (v[j] <= thr) ? (1u << j) : 0u;
                        // every entry of this chunk carries the chunk minimum as its (lower-bound) score: enough
                        // to prune it later, and no dynamic indexing of the register tile is needed
                        const uint32_t sc = pack_entry(cm, 0u);
                        while (mask) {
                            const int j = __ffs(mask) - 1;
                            mask &= mask - 1;
                            if (cnt == TC_LCAP) {           // full: drop entries the tighter threshold excludes
                                int wpos = 0;
                                for (int e = 0; e < TC_LCAP; e++) {
                                    const uint32_t ent = sList[e * TC_PR + lr];
                                    if (entry_score(ent) <= thr) sList[(wpos++) * TC_PR + lr] = ent;
                                }
                                cnt = wpos;
                            }
                            if (cnt < TC_LCAP) {
                                sList[cnt * TC_PR + lr] = sc | (uint32_t)(code0 + j);
                                cnt++;
                            } else {
                                ovf = true;
                            }
                        }
                    }
                }
            }
            // ---- re-rank with the reference arithmetic ----
            PROF(4);
            if (RAW) mbar_wait(RAW_FULL, pi & 1);
            PROF(5);       // prof[24+5]: epilogue wait RAW_FULL
            if (gi < p.n) {
                st_rows++;
                const int64_t src = p.rows ? p.rows[gi] : gi;
                const float *xg = p.x + src * D;
                auto getx = [&](int kk) -> float { return RAW ? sRaw[lr * RS + kk] : xg[kk]; };
                const bool fallback = !(band >= 0.f) || ovf;
                if (!fallback) {
                    const float thrF = m + p.gamma2 * fabsf(m) + band;
                    int ns = 0, sole = 0;
                    for (int e = 0; e < cnt; e++) {
                        const uint32_t ent = sList[e * TC_PR + lr];
                        const int j = ent & 0xFFFF;
                        if (entry_score(ent) <= thrF && j < p.k) { ns++; sole = j; }
                    }
                    st_cands += ns;
                    float best = FLT_MAX;
                    int bj = 0;
                    if (ns == 1 && !need_dist) {
                        bj = sole;
                    } else {
                        for (int e = 0; e < cnt; e++) {
                            const uint32_t ent = sList[e * TC_PR + lr];
                            const int j = ent & 0xFFFF;
                            if (!(entry_score(ent) <= thrF && j < p.k)) continue;
                            const float4 *cp = reinterpret_cast<const float4 *>(p.cbp + (size_t)j * p.dpc);
                            float a = 0.f;
                            for (int c4 = 0; c4 < p.dpc / 4; c4++) {
                                const float4 cv = __ldg(cp + c4);
                                float df;
                                if (4 * c4 + 0 < D) { df = getx(4 * c4 + 0) - cv.x; a = fmaf(df, df, a); }
                                if (4 * c4 + 1 < D) { df = getx(4 * c4 + 1) - cv.y; a = fmaf(df, df, a); }
                                if (4 * c4 + 2 < D) { df = getx(4 * c4 + 2) - cv.z; a = fmaf(df, df, a); }
                                if (4 * c4 + 3 < D) { df = getx(4 * c4 + 3) - cv.w; a = fmaf(df, df, a); }
                            }
                            if (a < best || (a == best && j < bj)) { best = a; bj = j; }
                        }
                    }
                    if (p.out_idx) p.out_idx[gi] = bj;
                    if (need_dist) p.out_dist[gi] = best;
                    if (p.acc) {
                        const float wv = p.w ? p.w[src] : 1.f;
                        float *ap = p.acc + (size_t)bj * p.acc_stride;
                        for (int c4 = 0; c4 < p.acc_stride / 4; c4++) {
                            float e4[4];
#pragma unroll
                            for (int i = 0; i < 4; i++) {
                                const int kk = 4 * c4 + i;
                                e4[i] = kk < D ? __fmul_rn(getx(kk), wv) : (kk == D ? wv : 0.f);
                            }
                            red_add_v4(ap + 4 * c4, e4[0], e4[1], e4[2], e4[3]);
                        }
                    }
                } else {
                    st_fb++;
                    const unsigned int pos = atomicAdd(&p.hdr->fb_count, 1u);
                    p.fb_list[pos] = (unsigned int)gi;
                }
            }
            if (RAW) mbar_arrive(RAW_FREE);
            PROF(6);       // prof[24+6]: epilogue re-rank + outputs + scatter
        }
        PROF_FLUSH(24, tid == TC_EPI_WARP0 * 32);
        // stats: one atomic per warp
        for (int o = 16; o > 0; o >>= 1) {
            st_rows += __shfl_xor_sync(0xffffffffu, st_rows, o);
            st_cands += __shfl_xor_sync(0xffffffffu, st_cands, o);
            st_fb += __shfl_xor_sync(0xffffffffu, st_fb, o);
        }
        if (lane == 0) {
            atomicAdd(&p.hdr->rows, st_rows);
            atomicAdd(&p.hdr->cands, st_cands);
            atomicAdd(&p.hdr->fb_rows, st_fb);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ------------------------------------------------------------------------------------------------
// Exact scan for the rows the tensor-core kernel could not decide: one warp per row, lanes stride
// over the codes, reference arithmetic, then a (distance, index) min-reduction with lowest-index ties.
__global__ void __launch_bounds__(128)
k_assign_fallback(const float *__restrict__ x, const float *__restrict__ w, const int64_t *__restrict__ rows, int d,
                  const float *__restrict__ cb, int k, const TcHeader *hdr, const unsigned int *__restrict__ fb_list,
                  float *__restrict__ out_dist, int64_t *__restrict__ out_idx, float *__restrict__ acc, int stride)
{
    __shared__ float xs[4][C3VQ_MAX_DIM];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const unsigned int count = hdr->fb_count;
    const unsigned int nwarps = gridDim.x * 4;
    for (unsigned int it = blockIdx.x * 4 + wib; it < count; it += nwarps) {
        const int64_t gi = fb_list[it];
        const int64_t src = rows ? rows[gi] : gi;
        __syncwarp();
        for (int c = lane; c < d; c += 32) xs[wib][c] = x[src * d + c];
        __syncwarp();
        float best = FLT_MAX;
        int bj = 0x7fffffff;
        for (int j = lane; j < k; j += 32) {
            const float *cr = cb + (size_t)j * d;
            float a = 0.f;
            for (int c = 0; c < d; c++) {
                const float df = xs[wib][c] - cr[c];
                a = fmaf(df, df, a);
            }
            if (a < best) { best = a; bj = j; }
        }
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
            if (ob < best || (ob == best && oj < bj)) { best = ob; bj = oj; }
        }
        if (bj == 0x7fffffff) bj = 0;                    // nothing below FLT_MAX: defined fallback index
        if (lane == 0) {
            if (out_idx) out_idx[gi] = bj;
            if (out_dist) out_dist[gi] = best;
        }
        if (acc) {
            const float wv = w ? w[src] : 1.f;
            float *ap = acc + (size_t)bj * stride;
            for (int c = lane; c <= d; c += 32) atomicAdd(ap + c, c < d ? __fmul_rn(xs[wib][c], wv) : wv);
        }
    }
}

// ------------------------------------------------------------------------------------------------
static inline size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

struct TcLayout {
    size_t off_mu, off_bimg, off_cbp, off_fb, total;
    int dp, nc, dpc, k_pad;
};

static TcLayout tc_layout(int64_t n, int d, int k)
{
    TcLayout l;
    l.dp = tc_dp(d);
    l.nc = tc_nc(d);
    l.dpc = (d + 3) & ~3;
    l.k_pad = (k + TC_BN - 1) / TC_BN * TC_BN;
    if (l.k_pad == 0) l.k_pad = TC_BN;
    size_t o = al256(sizeof(TcHeader));
    l.off_mu = o;
    o += al256(80 * sizeof(float));
    l.off_bimg = o;
    o += al256((size_t)l.k_pad * l.dp * sizeof(float));
    l.off_cbp = o;
    o += al256((size_t)(k > 0 ? k : 1) * l.dpc * sizeof(float));
    l.off_fb = o;
    o += al256((size_t)(n > 0 ? n : 1) * sizeof(unsigned int));
    l.total = o;
    return l;
}

size_t tc_workspace_bytes(int64_t n, int d, int k)
{
    if (n < 0 || d < 1 || d > C3VQ_MAX_DIM || k < 0) return 256;
    return tc_layout(n, d, k).total;
}

template <int DP, int DCONST>
static int launch_tc_t(const TcParams &p, int grid, cudaStream_t s)
{
    using L = TcSmem<DP>;
    TcParams q = p;
    q.nstages = L::stages();
    if (q.nstages < 1) {
        set_error("tensor-core kernel: no room for a codebook stage at DP=%d", DP);
        return C3VQ_EINVAL;
    }
    const size_t smem = L::bytes(q.nstages);
    C3VQ_CUDA(cudaFuncSetAttribute(k_assign_tc<DP, DCONST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_assign_tc<DP, DCONST><<<grid, TC_THREADS, smem, s>>>(q);
    C3VQ_LAUNCH_CHECK("k_assign_tc");
    return C3VQ_OK;
}

int launch_assign_tc(const float *x, const float *w, const int64_t *rows, int64_t n, int d, const float *codebook,
                     int k, float *out_dist, int64_t *out_idx, float *acc, float *dbg_scores, float *dbg_norm2,
                     float *dbg_band, void *ws, size_t ws_bytes, cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    if (k == 0) return launch_assign_exact(x, rows, n, d, codebook, k, out_dist, out_idx, s);   // defaults only
    C3VQ_CHECK_ARG(n < 0xffffffffLL, "n=%lld too large for the fallback list", (long long)n);
    const TcLayout l = tc_layout(n, d, k);
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
    k_tc_prep_mean<<<(k + 127) / 128, 256, 0, s>>>(codebook, k, d, hdr);
    C3VQ_LAUNCH_CHECK("k_tc_prep_mean");
    k_tc_prep_image<<<(l.k_pad + 127) / 128, 128, 0, s>>>(codebook, k, l.k_pad, d, l.dp, l.nc, l.dpc, hdr, mu, bimg,
                                                        cbp);
    C3VQ_LAUNCH_CHECK("k_tc_prep_image");

    TcParams p{};
    p.x = x; p.w = w; p.rows = rows; p.n = n; p.d = d; p.k = k;
    p.nsteps = l.k_pad / TC_BN;
    p.npairs = (n + TC_PR - 1) / TC_PR;
    p.bimg = bimg; p.cbp = cbp; p.dpc = l.dpc; p.mu = mu; p.hdr = hdr;
    p.out_dist = out_dist; p.out_idx = out_idx; p.acc = acc; p.acc_stride = acc_stride(d);
    p.fb_list = fb;
    p.dbg_scores = dbg_scores; p.dbg_norm2 = dbg_norm2; p.dbg_band = dbg_band;
    p.gamma2 = 2.1f * (float)(d + 3) * 5.9604644775390625e-8f;      // 2.1 (d+3) 2^-24
    const int sms = sm_count();
    const int grid = (int)(p.npairs < sms ? p.npairs : sms);

    int rc;
#ifdef C3VQ_TC_SINGLE   // developer build: one instantiation only (fast compile, SASS inspection)
    rc = launch_tc_t<48, 45>(p, grid, s);
#else
    switch (d) {
        case 6: rc = launch_tc_t<8, 6>(p, grid, s); break;
        case 7: rc = launch_tc_t<16, 7>(p, grid, s); break;
        case 45: rc = launch_tc_t<48, 45>(p, grid, s); break;
        case 48: rc = launch_tc_t<56, 48>(p, grid, s); break;
        default: {
            // generic widths: the image is laid out for DP = tc_dp(d); kernels exist for every multiple of 8
            switch (l.dp) {
                case 8: rc = launch_tc_t<8, 0>(p, grid, s); break;
                case 16: rc = launch_tc_t<16, 0>(p, grid, s); break;
                case 24: rc = launch_tc_t<24, 0>(p, grid, s); break;
                case 32: rc = launch_tc_t<32, 0>(p, grid, s); break;
                case 40: rc = launch_tc_t<40, 0>(p, grid, s); break;
                case 48: rc = launch_tc_t<48, 0>(p, grid, s); break;
                case 56: rc = launch_tc_t<56, 0>(p, grid, s); break;
                case 64: rc = launch_tc_t<64, 0>(p, grid, s); break;
                default: rc = launch_tc_t<72, 0>(p, grid, s); break;
            }
        }
    }
#endif
    if (rc) return rc;
    if (dbg_scores == nullptr) {
        k_assign_fallback<<<sms * 4, 128, 0, s>>>(x, w, rows, d, codebook, k, hdr, fb, out_dist, out_idx, acc,
                                                  acc_stride(d));
        C3VQ_LAUNCH_CHECK("k_assign_fallback");
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

}  // namespace c3vq
