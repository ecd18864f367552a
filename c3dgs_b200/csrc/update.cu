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
#include <cstdlib>

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

// ------------------------------------------------------------------------------------------------
// Atomic-free weighted segmented sum ("sorted" path).  The L2 performs fp32 reductions at a few clocks
// per 32-byte sector, so adding 4*(d+1) bytes per row into the accumulator table costs several times
// the HBM time of reading the rows (measured on B200: 5.8 M x 45-D rows, red.global / bulk-reduce ~1.1 ms
// vs 0.17 ms to stream them).  Instead the rows are counting-sorted by codeword (shared-memory privatised
// histogram, one scan, one scatter of row ids + weights), then each warp sums a contiguous run of the
// sorted order in registers and touches the table only at segment boundaries.
//   hist [k], offs [k+1], cursor [k] u32; perm [n] u32 (source row); wperm [n] f32 (its weight)
constexpr int SEG_CH = 128;         // sorted positions per warp work item

__global__ void __launch_bounds__(1024)
k_seg_hist(const int64_t *__restrict__ idx, int64_t n, int k, unsigned int *__restrict__ hist, int use_smem)
{
    extern __shared__ unsigned int sh[];
    if (use_smem) {
        for (int j = threadIdx.x; j < k; j += blockDim.x) sh[j] = 0;
        __syncthreads();
    }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t j = idx[i];
        if ((uint64_t)j >= (uint64_t)k) continue;          // out-of-range code: the row is ignored
        if (use_smem) atomicAdd(&sh[j], 1u); else atomicAdd(&hist[j], 1u);
    }
    if (use_smem) {
        __syncthreads();
        for (int j = threadIdx.x; j < k; j += blockDim.x) {
            const unsigned int c = sh[j];
            if (c) atomicAdd(&hist[j], c);
        }
    }
}

// exclusive scan of hist[k] -> offs[k+1], cursor[j] = offs[j]; one CTA of 1024 threads
__global__ void __launch_bounds__(1024)
k_seg_scan(const unsigned int *__restrict__ hist, int k, unsigned int *__restrict__ offs,
           unsigned int *__restrict__ cursor)
{
    __shared__ unsigned int part[1024];
    const int per = (k + 1023) / 1024;
    const int j0 = threadIdx.x * per;
    unsigned int sum = 0;
    for (int j = j0; j < min(k, j0 + per); j++) sum += hist[j];
    part[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {                     // Hillis-Steele inclusive scan of the partials
        const unsigned int v = threadIdx.x >= o ? part[threadIdx.x - o] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned int run = part[threadIdx.x] - sum;
    for (int j = j0; j < min(k, j0 + per); j++) {
        offs[j] = run;
        cursor[j] = run;
        run += hist[j];
    }
    if (threadIdx.x == 1023) offs[k] = part[1023];
}

// Scatter of (source row, weight) into codeword order.  Each CTA owns a contiguous block of rows: a shared-memory
// histogram of the block, ONE global atomic per (CTA, codeword present) to reserve the codeword's slots for this
// CTA, then a second pass over the block (L2 resident) ranks the rows with shared-memory atomics.  A codeword that
// owns millions of rows therefore sees gridDim.x global atomics, not millions.
__global__ void __launch_bounds__(1024)
k_seg_scatter(const int64_t *__restrict__ idx, const float *__restrict__ w, const int64_t *__restrict__ rows,
              int64_t n, int k, unsigned int *__restrict__ cursor, unsigned int *__restrict__ perm,
              float *__restrict__ wperm)
{
    extern __shared__ unsigned int sh[];                      // [k] local count, then local cursor
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t i0 = (int64_t)blockIdx.x * per, i1 = min(n, i0 + per);
    for (int j = threadIdx.x; j < k; j += blockDim.x) sh[j] = 0;
    __syncthreads();
    for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
        const int64_t j = idx[i];
        if ((uint64_t)j < (uint64_t)k) atomicAdd(&sh[j], 1u);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        const unsigned int c = sh[j];
        sh[j] = c ? atomicAdd(&cursor[j], c) : 0u;
    }
    __syncthreads();
    for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
        const int64_t j = idx[i];
        if ((uint64_t)j >= (uint64_t)k) continue;
        const int64_t src = rows ? rows[i] : i;
        const unsigned int pos = atomicAdd(&sh[j], 1u);
        perm[pos] = (unsigned int)src;
        wperm[pos] = w ? w[src] : 1.f;
    }
}

// Fallback for codebooks whose histogram does not fit in shared memory: one global atomic per row.
__global__ void __launch_bounds__(256)
k_seg_scatter_global(const int64_t *__restrict__ idx, const float *__restrict__ w, const int64_t *__restrict__ rows,
                     int64_t n, int k, unsigned int *__restrict__ cursor, unsigned int *__restrict__ perm,
                     float *__restrict__ wperm)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t j = idx[i];
        if ((uint64_t)j >= (uint64_t)k) continue;
        const int64_t src = rows ? rows[i] : i;
        const unsigned int pos = atomicAdd(&cursor[j], 1u);
        perm[pos] = (unsigned int)src;
        wperm[pos] = w ? w[src] : 1.f;
    }
}

// Each CTA owns ONE contiguous range of the sorted order and walks it 8 * SEG_CH positions at a time, each warp
// SEG_CH of them; lanes over columns (coalesced 4*d-byte row reads).  Partial sums leave the registers at segment
// boundaries only, and go to a shared-memory table of the first SEG_SLOTS codewords of the current window (flushed
// with one global atomic per element when the window moves on), else straight to global.
constexpr int SEG_SLOTS = 16;
#ifndef SEG_UNROLL
#define SEG_UNROLL 16
#endif
constexpr int SEG_U = SEG_UNROLL;   // rows whose loads are in flight per warp
template <int NH>
__global__ void __launch_bounds__(256)
k_seg_sum(const float *__restrict__ x, const unsigned int *__restrict__ perm, const float *__restrict__ wperm,
          const unsigned int *__restrict__ offs, int d, int k, float *__restrict__ acc, int stride, int wcol)
{
    __shared__ float tab[SEG_SLOTS][68];
    __shared__ int s_first;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const unsigned int total = offs[k];
    const bool c0 = lane < d, c1 = NH == 2 && lane + 32 < d;
    const unsigned int per = ((total + gridDim.x - 1) / gridDim.x + 8 * SEG_CH - 1) / (8 * SEG_CH) * (8 * SEG_CH);
    const unsigned int r0 = blockIdx.x * per, r1 = min(total, r0 + per);
    int jf_prev = -1;
    for (int i = threadIdx.x; i < SEG_SLOTS * 68; i += blockDim.x) (&tab[0][0])[i] = 0.f;
    auto flush_table = [&](int jf) {                          // CTA table -> global: slot s is codeword jf + s
        for (int i = threadIdx.x; i < SEG_SLOTS * 65; i += blockDim.x) {
            const int slot = i / 65, c = i - slot * 65;
            const float v = tab[slot][c];
            tab[slot][c] = 0.f;
            const int j = jf + slot;
            if (v != 0.f && j < k) {
                if (c < d) atomicAdd(acc + (size_t)j * stride + c, v);
                else if (c == 64 && wcol >= 0) atomicAdd(acc + (size_t)j * stride + wcol, v);
            }
        }
    };
    for (unsigned int q0 = r0; q0 < r1; q0 += 8 * SEG_CH) {
        if (threadIdx.x == 0) {                               // codeword of the window's first position
            int lo = 0, hi = k;
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (offs[mid] <= q0) lo = mid; else hi = mid;
            }
            s_first = lo;
        }
        __syncthreads();
        const int jf = s_first;
        if (jf != jf_prev) {                                  // the window moved: empty the table first
            if (jf_prev >= 0) flush_table(jf_prev);
            jf_prev = jf;
            __syncthreads();
        }
        const unsigned int p0 = q0 + wib * SEG_CH;
        if (p0 < r1) {
            const unsigned int p1 = min(p0 + (unsigned int)SEG_CH, r1);
            int j = jf;                                       // codeword of position p0: offs[j] <= p0 < offs[j+1]
            {
                int lo = jf, hi = k;
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (offs[mid] <= p0) lo = mid; else hi = mid;
                }
                j = lo;
            }
            unsigned int seg_end = offs[j + 1];
            float a0 = 0.f, a1 = 0.f, aw = 0.f;
            auto flush = [&]() {
                const int slot = j - jf;
                if (slot < SEG_SLOTS) {
                    if (c0) atomicAdd(&tab[slot][lane], a0);
                    if (c1) atomicAdd(&tab[slot][lane + 32], a1);
                    if (lane == 0) atomicAdd(&tab[slot][64], aw);
                } else {
                    float *ap = acc + (size_t)j * stride;
                    if (c0) atomicAdd(ap + lane, a0);
                    if (c1) atomicAdd(ap + lane + 32, a1);
                    if (lane == 0 && wcol >= 0) atomicAdd(ap + wcol, aw);
                }
                a0 = a1 = aw = 0.f;
            };
            for (unsigned int base = p0; base < p1; base += 32) {
                const unsigned int pl = base + lane;
                const unsigned int row_l = pl < p1 ? perm[pl] : 0u;
                const float w_l = pl < p1 ? wperm[pl] : 0.f;
                const int cnt = (int)min(32u, p1 - base);
                if (seg_end >= base + cnt) {                  // no boundary inside: straight-line accumulation
#pragma unroll SEG_U
                    for (int r = 0; r < cnt; r++) {
                        const unsigned int row = __shfl_sync(0xffffffffu, row_l, r);
                        const float wv = __shfl_sync(0xffffffffu, w_l, r);
                        const float *xp = x + (size_t)row * d + lane;
                        const float v0 = c0 ? xp[0] : 0.f;
                        const float v1 = c1 ? xp[32] : 0.f;
                        a0 = __fadd_rn(a0, __fmul_rn(v0, wv));   // product rounded first (the reference's x*w temporary)
                        a1 = __fadd_rn(a1, __fmul_rn(v1, wv));
                        aw = __fadd_rn(aw, wv);
                    }
                } else {
                    for (int r = 0; r < cnt; r++) {
                        while (base + r >= seg_end) {         // leave the finished (or empty) segment
                            flush();
                            j++;
                            seg_end = offs[j + 1];
                        }
                        const unsigned int row = __shfl_sync(0xffffffffu, row_l, r);
                        const float wv = __shfl_sync(0xffffffffu, w_l, r);
                        const float *xp = x + (size_t)row * d + lane;
                        const float v0 = c0 ? xp[0] : 0.f;
                        const float v1 = c1 ? xp[32] : 0.f;
                        a0 = __fadd_rn(a0, __fmul_rn(v0, wv));
                        a1 = __fadd_rn(a1, __fmul_rn(v1, wv));
                        aw = __fadd_rn(aw, wv);
                    }
                }
            }
            flush();
        }
        __syncthreads();
    }
    if (jf_prev >= 0) flush_table(jf_prev);
}

size_t wseg_workspace_bytes(int64_t n, int k);
int launch_accumulate_windowed(const float *x, const float *w, const int64_t *idx, int64_t n, int d, int k, float *acc,
                               int stride, int wcol, void *ws, size_t ws_bytes, cudaStream_t s);

// developer knob: C3VQ_SUMS_WINDOWED=1 selects the windowed sort for rows in storage order (default: the plain
// codeword-sorted gather, which measured faster: profiles/r2a_sums_probe_*.txt)
static bool use_windowed()
{
    const char *e = getenv("C3VQ_SUMS_WINDOWED");       // read per call: tests and A/B tools flip it in-process
    return e && e[0] == '1';
}

size_t seg_workspace_bytes(int64_t n, int k)
{
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t plain = al((size_t)(k + 1) * 4) * 3 + al((size_t)(n > 0 ? n : 1) * 4) * 2;
    const size_t win = wseg_workspace_bytes(n, k);
    return plain > win ? plain : win;
}

int launch_accumulate_sorted(const float *x, const float *w, const int64_t *rows, const int64_t *idx, int64_t n, int d,
                             int k, float *acc, int stride, int wcol, void *ws, size_t ws_bytes, cudaStream_t s)
{
    if (n == 0 || k == 0) return C3VQ_OK;
    C3VQ_CHECK_ARG(n < 0xffffffffLL, "n=%lld too large for 32-bit row ids", (long long)n);
    if (rows == nullptr && k <= 65536 && use_windowed())
        return launch_accumulate_windowed(x, w, idx, n, d, k, acc, stride, wcol, ws, ws_bytes, s);
    if (ws == nullptr || ws_bytes < seg_workspace_bytes(n, k)) {
        set_error("workspace too small: %zu < %zu", ws_bytes, seg_workspace_bytes(n, k));
        return C3VQ_EWORKSPACE;
    }
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    char *p = (char *)ws;
    unsigned int *hist = (unsigned int *)p;   p += al((size_t)(k + 1) * 4);
    unsigned int *offs = (unsigned int *)p;   p += al((size_t)(k + 1) * 4);
    unsigned int *cursor = (unsigned int *)p; p += al((size_t)(k + 1) * 4);
    unsigned int *perm = (unsigned int *)p;   p += al((size_t)n * 4);
    float *wperm = (float *)p;
    const int sms = sm_count();
    C3VQ_CUDA(cudaMemsetAsync(hist, 0, (size_t)k * 4, s));
    const int use_smem = k <= 12288;
    int64_t hb = (n + 1023) / 1024;
    if (hb > sms * 2) hb = sms * 2;
    k_seg_hist<<<(unsigned)hb, 1024, use_smem ? (size_t)k * 4 : 0, s>>>(idx, n, k, hist, use_smem);
    C3VQ_LAUNCH_CHECK("k_seg_hist");
    k_seg_scan<<<1, 1024, 0, s>>>(hist, k, offs, cursor);
    C3VQ_LAUNCH_CHECK("k_seg_scan");
    if (use_smem) {
        int64_t sb = (n + 4095) / 4096;
        if (sb > sms * 2) sb = sms * 2;
        k_seg_scatter<<<(unsigned)sb, 1024, (size_t)k * 4, s>>>(idx, w, rows, n, k, cursor, perm, wperm);
    } else {
        int64_t sb = (n + 255) / 256;
        if (sb > (int64_t)sms * 16) sb = (int64_t)sms * 16;
        k_seg_scatter_global<<<(unsigned)sb, 256, 0, s>>>(idx, w, rows, n, k, cursor, perm, wperm);
    }
    C3VQ_LAUNCH_CHECK("k_seg_scatter");
    int64_t gb = (n + 8 * SEG_CH - 1) / (8 * SEG_CH);
    if (gb > (int64_t)sms * 6) gb = (int64_t)sms * 6;
    if (d <= 32)
        k_seg_sum<1><<<(unsigned)gb, 256, 0, s>>>(x, perm, wperm, offs, d, k, acc, stride, wcol);
    else
        k_seg_sum<2><<<(unsigned)gb, 256, 0, s>>>(x, perm, wperm, offs, d, k, acc, stride, wcol);
    C3VQ_LAUNCH_CHECK("k_seg_sum");
    return C3VQ_OK;
}

// ------------------------------------------------------------------------------------------------
// Windowed variant of the sorted sums (rows == nullptr: x is consumed in storage order).
//
// The codeword-sorted gather above reads 4*d-byte rows from all over x: DRAM moves 64-byte granules and opens a new
// page for almost every row, so it runs at 39 % of the DRAM rate while moving 1.42x the algorithmic bytes.  Here the
// sort key is (window, codeword) with windows of WSEG_WIN = 65536 consecutive rows (11.8 MB at d = 45): the sorted
// order visits the input window by window, the whole grid works inside a band of a few windows at any time, and the
// rows of the band after the current one are pulled into the L2 by sequential prefetches in storage order.  DRAM
// therefore sees one sequential pass over x (traffic = algorithmic bytes) and the row gathers hit the L2.
//   k_wseg_rank   rank of every row inside its (window, codeword) bucket: one warp-aggregated atomicAdd with return
//                 on the bucket counter (which so becomes the histogram); slot[i] = code << 16 | rank
//   k_wseg_scan   per window (independent: window v owns sorted positions [v * WIN, v * WIN + rows of v)):
//                 exclusive scan of its K counters
//   k_wseg_place  perm[offs[bucket] + rank] = (code << 16 | row in window, weight bits): no atomics
//   k_wseg_sum    warps take chunks of WSEG_CH sorted positions in grid-stride order, WSEG_U row loads in flight,
//                 partial sums in registers, one reduction into the table per codeword change
constexpr int WSEG_WIN_LOG = 16;
constexpr int WSEG_WIN = 1 << WSEG_WIN_LOG;
#ifndef C3VQ_WSEG_U
#define C3VQ_WSEG_U 8
#endif
constexpr int WSEG_CH = 32;            // sorted positions per warp work item (one perm entry per lane)
constexpr int WSEG_U = C3VQ_WSEG_U;    // rows whose loads are in flight per warp (divides 32)
constexpr int WSEG_SUM_THREADS = 512;
// Two CTAs per SM: 148 x 32 warps x 32 rows = 151 k rows = 27 MB at d = 45 is the band the grid works on at a time;
// each chunk pulls the storage rows of the chunk WSEG windows-ahead of it into the L2, far enough that a window is
// resident before the first warp of the grid reaches its sorted positions (band + one window).
constexpr int WSEG_CTAS_PER_SM = 2;

__global__ void __launch_bounds__(256)
k_wseg_rank(const int64_t *__restrict__ idx, int64_t n, int k, unsigned int *__restrict__ hist,
            unsigned int *__restrict__ slot)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i0 = warp * 32; i0 < n; i0 += nwarps * 32) {      // warp-uniform trip count
        const int64_t i = i0 + lane;
        const int64_t j = i < n ? idx[i] : -1;
        const bool valid = (uint64_t)j < (uint64_t)k;              // out-of-range code: the row is ignored
        // rows of one warp share a window (i0 is a multiple of 32): the codeword alone tells buckets apart
        const unsigned int key = valid ? (unsigned int)j : 0x80000000u + (unsigned int)lane;
        const unsigned int peers = __match_any_sync(0xffffffffu, key);
        const int leader = __ffs(peers) - 1;
        unsigned int base = 0;
        if (valid && lane == leader)
            base = atomicAdd(&hist[(size_t)(i0 >> WSEG_WIN_LOG) * k + (unsigned int)j], (unsigned int)__popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (i < n) slot[i] = valid ? (((unsigned int)j << 16) | (base + __popc(peers & ((1u << lane) - 1u)))) : 0xffffffffu;
    }
}

// one CTA per window: offs[v][c] = v * WIN + sum_{c' < c} hist[v][c'], wlen[v] = rows of the window with a valid code
__global__ void __launch_bounds__(1024)
k_wseg_scan(const unsigned int *__restrict__ hist, int k, unsigned int *__restrict__ offs, unsigned int *__restrict__ wlen)
{
    __shared__ unsigned int part[1024];
    const unsigned int *h = hist + (size_t)blockIdx.x * k;
    unsigned int *o = offs + (size_t)blockIdx.x * k;
    const int per = (k + 1023) / 1024;
    const int j0 = threadIdx.x * per;
    unsigned int sum = 0;
    for (int j = j0; j < min(k, j0 + per); j++) sum += h[j];
    part[threadIdx.x] = sum;
    __syncthreads();
    for (int st = 1; st < 1024; st <<= 1) {
        const unsigned int v = threadIdx.x >= st ? part[threadIdx.x - st] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned int run = blockIdx.x * (unsigned int)WSEG_WIN + part[threadIdx.x] - sum;
    for (int j = j0; j < min(k, j0 + per); j++) {
        o[j] = run;
        run += h[j];
    }
    if (threadIdx.x == 1023) wlen[blockIdx.x] = part[1023];
}

__global__ void __launch_bounds__(256)
k_wseg_place(const unsigned int *__restrict__ slot, const float *__restrict__ w, int64_t n, int k,
             const unsigned int *__restrict__ offs, uint2 *__restrict__ perm)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned int sl = slot[i];
        if (sl == 0xffffffffu) continue;
        const unsigned int code = sl >> 16;
        const unsigned int pos = offs[(size_t)(i >> WSEG_WIN_LOG) * k + code] + (sl & 0xffffu);
        perm[pos] = make_uint2((code << 16) | (unsigned int)(i & (WSEG_WIN - 1)), __float_as_uint(w ? w[i] : 1.f));
    }
}

template <int NH>
__global__ void __launch_bounds__(WSEG_SUM_THREADS, WSEG_CTAS_PER_SM)
k_wseg_sum(const float *__restrict__ x, const uint2 *__restrict__ perm, const unsigned int *__restrict__ wlen,
           int64_t n, int d, float *__restrict__ acc, int stride, int wcol)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t nchunks = (n + WSEG_CH - 1) / WSEG_CH;            // windows are multiples of the chunk
    constexpr int64_t CPW = WSEG_WIN / WSEG_CH;                     // chunks per window
    const int64_t ahead = ((nwarps + CPW - 1) / CPW + 1) * CPW;     // prefetch distance in chunks: the band + one window
    const bool c0 = lane < d, c1 = NH == 2 && lane + 32 < d;
    const size_t row_bytes = (size_t)d * 4;
    for (int64_t g = warp; g < nchunks; g += nwarps) {
        // pull the storage rows [gn * CH, gn * CH + CH) into the L2, sequentially in memory: chunk g of window v
        // fetches the same slice of window v + ahead / CPW, so every window is fetched exactly once, by the
        // chunks of the window `ahead` before it (the first windows are cold)
        {
            const int64_t gn = g + ahead;
            if (gn < nchunks) {
                const char *pb = reinterpret_cast<const char *>(x) + (size_t)gn * WSEG_CH * row_bytes;
                const size_t bytes = (size_t)min((int64_t)WSEG_CH, n - gn * WSEG_CH) * row_bytes;
                for (size_t o = (size_t)lane * 128; o < bytes; o += 32 * 128)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(pb + o));
            }
        }
        const int64_t p0 = g * WSEG_CH;
        const unsigned int v = (unsigned int)(p0 >> WSEG_WIN_LOG);
        const unsigned int in_win = (unsigned int)(p0 & (WSEG_WIN - 1));
        const unsigned int len = wlen[v];
        if (in_win >= len) continue;                                 // rows with invalid codes left a gap at the window's end
        const int cnt = (int)min((unsigned int)WSEG_CH, len - in_win);
        const float *xw = x + (size_t)v * WSEG_WIN * d + lane;
        const uint2 e = lane < cnt ? perm[p0 + lane] : make_uint2(0u, 0u);
        int cur = -1;
        float a0 = 0.f, a1 = 0.f, aw = 0.f;
        auto flush = [&]() {
            float *ap = acc + (size_t)cur * stride;
            if (c0) atomicAdd(ap + lane, a0);
            if (c1) atomicAdd(ap + lane + 32, a1);
            if (lane == 0 && wcol >= 0) atomicAdd(ap + wcol, aw);
            a0 = a1 = aw = 0.f;
        };
#pragma unroll 1
        for (int r0 = 0; r0 < cnt; r0 += WSEG_U) {
            float v0[WSEG_U], v1[WSEG_U];
            unsigned int key[WSEG_U], wb[WSEG_U];
#pragma unroll
            for (int u = 0; u < WSEG_U; u++) {
                key[u] = __shfl_sync(0xffffffffu, e.x, r0 + u);
                wb[u] = __shfl_sync(0xffffffffu, e.y, r0 + u);
                const float *xp = xw + (size_t)(key[u] & 0xffffu) * d;
                const bool live = r0 + u < cnt;
                v0[u] = (live && c0) ? xp[0] : 0.f;
                v1[u] = (live && c1) ? xp[32] : 0.f;
            }
#pragma unroll
            for (int u = 0; u < WSEG_U; u++) {
                if (r0 + u < cnt) {                                  // warp-uniform
                    const int code = (int)(key[u] >> 16);
                    if (code != cur) {
                        if (cur >= 0) flush();
                        cur = code;
                    }
                    const float wv = __uint_as_float(wb[u]);
                    a0 = __fadd_rn(a0, __fmul_rn(v0[u], wv));        // product rounded first (the reference's x*w temporary)
                    a1 = __fadd_rn(a1, __fmul_rn(v1[u], wv));
                    aw = __fadd_rn(aw, wv);
                }
            }
        }
        if (cur >= 0) flush();
    }
}

static inline size_t wseg_al(size_t v) { return (v + 255) & ~(size_t)255; }

size_t wseg_workspace_bytes(int64_t n, int k)
{
    const size_t nw = (size_t)((n + WSEG_WIN - 1) >> WSEG_WIN_LOG) + 1;
    return 2 * wseg_al(nw * (size_t)k * 4) + wseg_al(nw * 4) + wseg_al((size_t)(n > 0 ? n : 1) * 4) +
           wseg_al((size_t)(n > 0 ? n : 1) * 8);
}

int launch_accumulate_windowed(const float *x, const float *w, const int64_t *idx, int64_t n, int d, int k, float *acc,
                               int stride, int wcol, void *ws, size_t ws_bytes, cudaStream_t s)
{
    if (n == 0 || k == 0) return C3VQ_OK;
    C3VQ_CHECK_ARG(k <= 65536, "k=%d does not fit the 16-bit code field", k);
    C3VQ_CHECK_ARG(n < 0xffffffffLL, "n=%lld too large for 32-bit positions", (long long)n);
    if (ws == nullptr || ws_bytes < wseg_workspace_bytes(n, k)) {
        set_error("workspace too small: %zu < %zu", ws_bytes, wseg_workspace_bytes(n, k));
        return C3VQ_EWORKSPACE;
    }
    const size_t nw = (size_t)((n + WSEG_WIN - 1) >> WSEG_WIN_LOG);
    char *p = (char *)ws;
    unsigned int *hist = (unsigned int *)p; p += wseg_al((nw + 1) * (size_t)k * 4);
    unsigned int *offs = (unsigned int *)p; p += wseg_al((nw + 1) * (size_t)k * 4);
    unsigned int *wlen = (unsigned int *)p; p += wseg_al((nw + 1) * 4);
    unsigned int *slot = (unsigned int *)p; p += wseg_al((size_t)n * 4);
    uint2 *perm = (uint2 *)p;
    const int sms = sm_count();
    C3VQ_CUDA(cudaMemsetAsync(hist, 0, nw * (size_t)k * 4, s));
    int64_t rb = (n + 255) / 256;
    if (rb > (int64_t)sms * 16) rb = (int64_t)sms * 16;
    k_wseg_rank<<<(unsigned)rb, 256, 0, s>>>(idx, n, k, hist, slot);
    C3VQ_LAUNCH_CHECK("k_wseg_rank");
    k_wseg_scan<<<(unsigned)nw, 1024, 0, s>>>(hist, k, offs, wlen);
    C3VQ_LAUNCH_CHECK("k_wseg_scan");
    k_wseg_place<<<(unsigned)rb, 256, 0, s>>>(slot, w, n, k, offs, perm);
    C3VQ_LAUNCH_CHECK("k_wseg_place");
    const int64_t nchunks = (n + WSEG_CH - 1) / WSEG_CH;
    int64_t gb = (nchunks + WSEG_SUM_THREADS / 32 - 1) / (WSEG_SUM_THREADS / 32);
    if (gb > (int64_t)sms * WSEG_CTAS_PER_SM) gb = (int64_t)sms * WSEG_CTAS_PER_SM;
    if (d <= 32)
        k_wseg_sum<1><<<(unsigned)gb, WSEG_SUM_THREADS, 0, s>>>(x, perm, wlen, n, d, acc, stride, wcol);
    else
        k_wseg_sum<2><<<(unsigned)gb, WSEG_SUM_THREADS, 0, s>>>(x, perm, wlen, n, d, acc, stride, wcol);
    C3VQ_LAUNCH_CHECK("k_wseg_sum");
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
