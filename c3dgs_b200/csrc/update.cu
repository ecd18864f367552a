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

// ------------------------------------------------------------------------------------------------
// Atomic-free weighted segmented sum ("sorted" path).  The L2 performs fp32 reductions at a few clocks
// per 32-byte sector, so adding 4*(d+1) bytes per row into the accumulator table costs several times
// the HBM time of reading the rows (measured on B200: 5.8 M x 45-D rows, red.global / bulk-reduce ~1.1 ms
// vs 0.17 ms to stream them).  Instead the rows are counting-sorted by codeword (shared-memory privatised
// histogram, one scan, one scatter of row ids + weights), then each warp sums a contiguous run of the
// sorted order in registers and touches the table only at segment boundaries.
//   hist [k], offs [k+1], cursor [k] u32; perm [n] u32 (source row); wperm [n] f32 (its weight)
// Tried in round 2 and dropped (profiles/r2a_sums_probe_*, r2b_sums_probe_windowed.txt): sorting by (window of 65536
// rows, codeword) with sequential L2 prefetch of the coming windows, so that DRAM sees one sequential pass and the
// gathers hit the L2 — 0.83-0.88 ms against 0.50-0.62 ms for this kernel (twice the sort traffic, four times the
// segment flushes).  What replaced the second pass over x altogether is the fused mode of assign_tc.cu.
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

size_t seg_workspace_bytes(int64_t n, int k)
{
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    return al((size_t)(k + 1) * 4) * 3 + al((size_t)(n > 0 ? n : 1) * 4) * 2;
}

int launch_accumulate_sorted(const float *x, const float *w, const int64_t *rows, const int64_t *idx, int64_t n, int d,
                             int k, float *acc, int stride, int wcol, void *ws, size_t ws_bytes, cudaStream_t s)
{
    if (n == 0 || k == 0) return C3VQ_OK;
    C3VQ_CHECK_ARG(n < 0xffffffffLL, "n=%lld too large for 32-bit row ids", (long long)n);
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
