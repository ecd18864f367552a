// gather.cu — indexed-codebook decode and mini-batch row gather.
//
// Replaces the torch advanced-indexing gathers `codebook[idx]`
// (compression/vq.py:40-42; scene/gaussian_model.py:220, :240, :257, :893-897) and
// `features[batch]` (compression/vq.py:70).  HBM bound: algorithmic bytes per output
// row = 8 (idx, i64) + 4*d (row written); the table is L2 resident for codebooks.
// Rows whose byte length is a multiple of 16 move as float4 with fully coalesced
// stores; other widths are written as float4 too, a warp per span of 32 output rows: from four padded copies of
// the table when the caller lends scratch (k_gather_padded, c3vq_gather_ws), else with 4-byte loads (k_gather_span).
#include "common.cuh"

namespace c3vq {

// output rows are written once and not read by the kernel: streaming stores keep them from evicting the table rows
// (the kept-rows table of a compressed scene is tens of MB and lives in the L2 between its reads)
#ifdef C3VQ_GATHER_PLAIN_ST
#define C3VQ_GATHER_ST(p, v) (*(p) = (v))
#else
#define C3VQ_GATHER_ST(p, v) __stcs((p), (v))
#endif

template <typename T>
__global__ void __launch_bounds__(256)
k_gather(const T *__restrict__ table, int64_t table_rows, int dv, const int64_t *__restrict__ idx, int64_t n,
         T *__restrict__ out)
{
    const int64_t total = n * dv;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = e / dv;
        const int c = (int)(e - i * dv);
        const int64_t j = idx[i];
        T v = T();
        if ((uint64_t)j < (uint64_t)table_rows) v = table[j * dv + c];
        out[e] = v;
    }
}

// 32-bit index arithmetic variant (n*dv < 2^31): the common case, much cheaper division.
template <typename T>
__global__ void __launch_bounds__(256)
k_gather32(const T *__restrict__ table, int64_t table_rows, int dv, const int64_t *__restrict__ idx, int total,
           T *__restrict__ out)
{
    // four independent elements per thread and round: the table rows of a large table (kept rows behind a codebook)
    // come from the L2 or DRAM, and one load in flight per thread leaves the gather latency bound
    const int stride = gridDim.x * blockDim.x;
    for (int e0 = blockIdx.x * blockDim.x + threadIdx.x; e0 < total; e0 += 4 * stride) {
        const T *src[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = e0 + u * stride;
            src[u] = nullptr;
            if (e < total && e >= 0) {
                const int i = e / dv;
                const int64_t j = idx[i];
                if ((uint64_t)j < (uint64_t)table_rows) src[u] = table + j * dv + (e - i * dv);
            }
        }
        T v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) v[u] = src[u] ? __ldg(src[u]) : T();
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = e0 + u * stride;
            if (e < total && e >= 0) out[e] = v[u];
        }
    }
}

// Rows that are not a multiple of 16 bytes (45 floats = 180 B is the SH width): a warp takes 32 consecutive output rows,
// i.e. one contiguous, 16-byte aligned span of 32*d floats, reads its 32 indices with one coalesced load, parks them in
// shared memory and writes the span as float4 — one multiply-shift and one 16-byte store per four elements instead
// of a division, an index load and a 4-byte store per element.
__global__ void __launch_bounds__(256)
k_gather_span(const float *__restrict__ table, int64_t table_rows, int d, unsigned int inv,
              const int64_t *__restrict__ idx, int64_t n, float *__restrict__ out)
{
    __shared__ const float *sj[8][32];                               // source row of each output row (null: zeros)
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t nwarps = (int64_t)gridDim.x * 8;
    for (int64_t r0 = ((int64_t)blockIdx.x * 8 + wib) * 32; r0 < n; r0 += nwarps * 32) {
        const int rows_here = (int)min((int64_t)32, n - r0);
        long long j = -1;
        if (lane < rows_here) {
            j = idx[r0 + lane];
            if ((uint64_t)j >= (uint64_t)table_rows) j = -1;          // out of range: zeros, as before
        }
        __syncwarp();
        sj[wib][lane] = j >= 0 ? table + j * d : nullptr;
        __syncwarp();
        float *dst = out + r0 * d;                                    // 32 * d * 4 bytes per span: 16-byte aligned
        const unsigned int total = (unsigned int)rows_here * (unsigned int)d;
        for (unsigned int e = lane * 4; e < total; e += 128) {
            // row of element e: multiply-shift (inv = ceil(2^20 / d), exact for e < 32 d while d < 181), else divide
            unsigned int r = inv ? (e * inv) >> 20 : e / (unsigned int)d, c = e - r * (unsigned int)d;
            float v[4];
            const float *src = sj[wib][r];                            // reloaded only where the float4 crosses a row end
#pragma unroll
            for (int t = 0; t < 4; t++) {
                v[t] = 0.f;
                if (e + t < total) {
                    if (src != nullptr) v[t] = __ldg(src + c);
                    if (++c == (unsigned int)d) {
                        c = 0;
                        src = sj[wib][(++r) & 31];
                    }
                }
            }
            if (e + 3 < total) {
                C3VQ_GATHER_ST(reinterpret_cast<float4 *>(dst + e), make_float4(v[0], v[1], v[2], v[3]));
            } else {
                for (int t = 0; t < 4 && e + t < total; t++) dst[e + t] = v[t];
            }
        }
    }
}

// (Round 2 tried staging the 32 rows in shared memory with lanes over columns and writing the span out from there:
// a quarter of the instructions, but 0.57 ms against 0.26 ms for this version — the extra shared-memory round trip and
// its warp syncs cost more than the index arithmetic they remove; profiles/r2f_bench.json vs r1x_bench.json.)

// (Also measured in round 2, 5.8 M rows x 45 floats from a 4096-row table: lanes over the columns of one row at a time,
// 4-byte coalesced loads and stores and no index arithmetic at all: 0.50 ms against 0.25 ms for the float4 stores above;
// a branch for the float4 that do not cross a row end: 0.32 ms, the warp runs both paths.  From the same table with
// 48-float rows, i.e. aligned float4 loads, the span copy below reaches 5.95 TB/s: what holds the 180-byte rows back is
// that their source rows are only 4-byte aligned.)

// (End of round 2, ncu on the three kernels at 5.8 M rows, profiles/r5a_ncu_gather_summary.txt: what bounds
// k_gather_span is the L1 data pipe — l1tex__data_pipe_lsu_wavefronts at 89 % of its peak, ~350 wavefronts per span of
// 32 rows (every 4-byte load instruction of a warp touches three or four source rows), where k_gather_span4 needs 171
// and sits at 55 %.  Two more organisations measured against it on the same box, both bit-exact, both dropped:
// (1) the two ALIGNED 16-byte granules around each output float4 loaded as float4 and shifted in registers by
// (address / 4) & 3 with nine selects, the columns past a row end as scalar loads: 0.254 ms against 0.246 ms — half the
// sectors, but each granule load still costs ~10 wavefronts (327 per span) and the instruction count stays ~80 per
// float4; (2) the span assembled in shared memory by 4-byte loads with lanes over consecutive span elements and sent
// out by one cp.async.bulk per span, double buffered per warp, 15 loads in flight per lane: 0.424 ms.)

// Rows of whole float4 (d % 4 == 0; 48 floats = the fork's colour rows): the same span organisation without the
// staging — the 32 * dv float4 of a span are copied lane-linear, four loads in flight per lane, the row of element e
// by a multiply-shift (exact for e < 32 * dv, dv < 128) against row pointers parked in shared memory.
__global__ void __launch_bounds__(256)
k_gather_span4(const float4 *__restrict__ table, int64_t table_rows, int dv, unsigned int inv,
               const int64_t *__restrict__ idx, int64_t n, float4 *__restrict__ out)
{
    __shared__ const float4 *sj[8][32];                              // source row of each output row (null: zeros)
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t nwarps = (int64_t)gridDim.x * 8;
    for (int64_t r0 = ((int64_t)blockIdx.x * 8 + wib) * 32; r0 < n; r0 += nwarps * 32) {
        const int rows_here = (int)min((int64_t)32, n - r0);
        long long j = -1;
        if (lane < rows_here) {
            j = idx[r0 + lane];
            if ((uint64_t)j >= (uint64_t)table_rows) j = -1;
        }
        __syncwarp();
        sj[wib][lane] = j >= 0 ? table + j * dv : nullptr;
        __syncwarp();
        float4 *dst = out + r0 * dv;
        const unsigned int total = (unsigned int)rows_here * (unsigned int)dv;
        for (unsigned int e0 = lane; e0 < total; e0 += 128) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const unsigned int e = e0 + 32 * u;
                v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (e < total) {
                    const unsigned int r = (e * inv) >> 20;
                    const float4 *src = sj[wib][r];
                    if (src != nullptr) v[u] = __ldg(src + (e - r * (unsigned int)dv));
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const unsigned int e = e0 + 32 * u;
                if (e < total) C3VQ_GATHER_ST(dst + e, v[u]);
            }
        }
    }
}

// Rows that are not whole float4, from padded copies of the table (the caller lends scratch: c3vq_gather_ws).  What
// bounds k_gather_span is the L1 data pipe: the source of an output float4 is only 4-byte aligned, so it takes four
// 4-byte loads.  A float4 that starts at column c of its row needs a copy of the table in which column c is 16-byte
// aligned, and c mod 4 takes four values: k_gather_pad_prep writes FOUR copies, copy q holding row j in a slot of P
// floats (P % 4 == 0) behind a pad of (4 - q) & 3 floats, so that the columns c = q (mod 4) are aligned.  For a
// 4096 x 45 codebook that is 3.4 MB, L2 resident.  Then an output float4 inside a row is ONE aligned 16-byte load from
// copy c & 3; where it crosses the row end after m = d - c < 4 columns, the first columns of the next output row come
// from that row's slot in copy 4 - m, whose pad is exactly m floats: the float4 at the slot start holds them at
// positions m .. 3.  Two loads and three selects there, one load elsewhere — the load pattern of k_gather_span4.
// (Writing the crossing float4 in a second sweep, one per lane, so that the main loop has no second load: 0.37 ms
// against 0.23 ms — the output stream then has 16-byte holes that are filled later, and partial-line writes are what
// the L2 / DRAM write path punishes most.)
__global__ void __launch_bounds__(256)
k_gather_pad_prep(const float *__restrict__ table, int rows, int d, int P, float *__restrict__ pad)
{
    const int total = 4 * rows * P;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
        const int slot = e / P, p = e - slot * P;
        const int q = slot / rows, j = slot - q * rows;
        const int col = p - ((4 - q) & 3);
        pad[e] = (col >= 0 && col < d) ? table[(int64_t)j * d + col] : 0.f;
    }
}

// 16-byte load through the read-only path under a predicate (zeros where it is off): no branch around the load
__device__ __forceinline__ float4 ldg_f4_if(const float4 *p, bool on)
{
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "@p ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];\n\t}"
        : "+f"(v.x), "+f"(v.y), "+f"(v.z), "+f"(v.w)
        : "l"(p), "r"((int)on)
        : "memory");                                     // keeps the stores of a round behind all of its loads
    return v;
}

__global__ void __launch_bounds__(256)
k_gather_padded(const float *__restrict__ table, const float4 *__restrict__ pad, int rows, int d, int P4,
                unsigned int inv, const int64_t *__restrict__ idx, int64_t n, float *__restrict__ out)
{
    __shared__ int sj[8][32];                                        // source row of each output row (-1: zeros)
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t nwarps = (int64_t)gridDim.x * 8;
    const unsigned int ud = (unsigned int)d;
    for (int64_t r0 = ((int64_t)blockIdx.x * 8 + wib) * 32; r0 < n; r0 += nwarps * 32) {
        const int rows_here = (int)min((int64_t)32, n - r0);
        long long j = -1;
        if (lane < rows_here) {
            j = idx[r0 + lane];
            if ((uint64_t)j >= (uint64_t)rows) j = -1;
        }
        __syncwarp();
        sj[wib][lane] = (int)j;
        __syncwarp();
        float *dst = out + r0 * d;
        const unsigned int total = (unsigned int)rows_here * ud;
        if (rows_here == 32) {                                       // 32 d floats: whole float4, none crosses the span end
            // four float4 per lane and round, every load issued (predicated, no branches) before the first select
            for (unsigned int e0 = lane * 4; e0 < total; e0 += 512) {
                float4 va[4], vb[4];
                unsigned int m[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const unsigned int e = min(e0 + 128u * u, total - 4u);       // past the end: a harmless repeat
                    const unsigned int r = (e * inv) >> 20, c = e - r * ud, q = c & 3u;
                    m[u] = ud - c;                                   // columns left in output row r (>= 1)
                    const int ja = sj[wib][r], jb = sj[wib][(r + 1u) & 31u];
                    va[u] = ldg_f4_if(pad + ((q * rows + ja) * P4 + ((((4u - q) & 3u) + c) >> 2)), ja >= 0);
                    // columns m .. 3 open output row r + 1: the start of its slot in copy 4 - m
                    vb[u] = ldg_f4_if(pad + (((4u - m[u]) & 3u) * rows + jb) * P4, m[u] < 4u && jb >= 0);
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const unsigned int e = e0 + 128u * u;
                    if (e < total) {
                        float4 v = va[u];
                        if (m[u] <= 1u) v.y = vb[u].y;
                        if (m[u] <= 2u) v.z = vb[u].z;
                        if (m[u] <= 3u) v.w = vb[u].w;
                        C3VQ_GATHER_ST(reinterpret_cast<float4 *>(dst + e), v);
                    }
                }
            }
        } else {                                                     // the last, partial span: from the table itself
            for (unsigned int e = lane; e < total; e += 32) {
                const unsigned int r = e / ud, c = e - r * ud;
                const int js = sj[wib][r];
                dst[e] = js >= 0 ? __ldg(table + (int64_t)js * d + c) : 0.f;
            }
        }
    }
}

constexpr size_t GATHER_PAD_MAX_BYTES = 16u << 20;       // the four copies must stay L2 resident
static inline int gather_pad_pitch(int d) { return (d + 9) & ~3; }   // pad (<= 3) + d + read-past (<= 3), whole float4

size_t gather_workspace_bytes(int64_t table_rows, int d, int64_t n)
{
    if (d < 12 || d >= 181 || d % 4 == 0 || table_rows <= 0) return 0;   // narrow rows: k_gather_span is faster
    const size_t bytes = (size_t)4 * (size_t)table_rows * gather_pad_pitch(d) * sizeof(float);
    if (bytes > GATHER_PAD_MAX_BYTES || n < 8 * table_rows) return 0;   // the copies must pay for themselves
    return bytes;
}

int launch_gather_ws(const float *table, int64_t table_rows, int d, const int64_t *idx, int64_t n, float *out,
                     void *ws, size_t ws_bytes, cudaStream_t s)
{
    const size_t need = gather_workspace_bytes(table_rows, d, n);
    if (need == 0 || ws == nullptr || ws_bytes < need || (reinterpret_cast<uintptr_t>(ws) & 15) != 0 ||
        (reinterpret_cast<uintptr_t>(out) & 15) != 0)
        return launch_gather(table, table_rows, d, idx, n, out, s);
    const int rows = (int)table_rows, P = gather_pad_pitch(d);
    const int total = 4 * rows * P;
    k_gather_pad_prep<<<(total + 255) / 256, 256, 0, s>>>(table, rows, d, P, (float *)ws);
    C3VQ_LAUNCH_CHECK("k_gather_pad_prep");
    int64_t blocks = (n + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    k_gather_padded<<<(unsigned)blocks, 256, 0, s>>>(table, (const float4 *)ws, rows, d, P / 4,
                                                     ((1u << 20) + d - 1) / d, idx, n, out);
    C3VQ_LAUNCH_CHECK("k_gather_padded");
    return C3VQ_OK;
}

template <typename T>
static int launch_gather_t(const T *table, int64_t table_rows, int dv, const int64_t *idx, int64_t n, T *out,
                           cudaStream_t s)
{
    const int64_t total = n * dv;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 64;
    if (blocks > cap) blocks = cap;
    if (total < 0x70000000LL)                 // leaves room for the 4 x stride look-ahead of the 32-bit index
        k_gather32<T><<<(unsigned)blocks, 256, 0, s>>>(table, table_rows, dv, idx, (int)total, out);
    else
        k_gather<T><<<(unsigned)blocks, 256, 0, s>>>(table, table_rows, dv, idx, n, out);
    C3VQ_LAUNCH_CHECK("k_gather");
    return C3VQ_OK;
}

int launch_gather(const float *table, int64_t table_rows, int d, const int64_t *idx, int64_t n, float *out,
                  cudaStream_t s)
{
    if (n == 0 || d == 0) return C3VQ_OK;
    const bool v4 = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(table) & 15) == 0) &&
                    ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
    int64_t blocks = (n + 255) / 256;                             // 8 warps x 32 rows per block and round
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (v4 && d / 4 < 128) {
        const int dv = d / 4;
        k_gather_span4<<<(unsigned)blocks, 256, 0, s>>>(reinterpret_cast<const float4 *>(table), table_rows, dv,
                                                        ((1u << 20) + dv - 1) / dv, idx, n,
                                                        reinterpret_cast<float4 *>(out));
        C3VQ_LAUNCH_CHECK("k_gather_span4");
        return C3VQ_OK;
    }
    if (v4)
        return launch_gather_t<float4>(reinterpret_cast<const float4 *>(table), table_rows, d / 4, idx, n,
                                       reinterpret_cast<float4 *>(out), s);
    if ((reinterpret_cast<uintptr_t>(out) & 15) == 0 && d < (1 << 24)) {
        k_gather_span<<<(unsigned)blocks, 256, 0, s>>>(table, table_rows, d, d < 181 ? ((1u << 20) + d - 1) / d : 0u,
                                                       idx, n, out);
        C3VQ_LAUNCH_CHECK("k_gather_span");
        return C3VQ_OK;
    }
    return launch_gather_t<float>(table, table_rows, d, idx, n, out, s);
}

__global__ void __launch_bounds__(256)
k_gather_w(const float *__restrict__ w, const int64_t *__restrict__ rows, int64_t n, float *__restrict__ out_w)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out_w[i] = w[rows[i]];
}

int launch_gather_rows(const float *x, const float *w, const int64_t *rows, int64_t n, int d, float *out_x,
                       float *out_w, cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    if (out_x) {
        // rows index an unbounded source here; the caller guarantees 0 <= rows[i] < N
        int rc = launch_gather(x, INT64_MAX, d, rows, n, out_x, s);
        if (rc) return rc;
    }
    if (w && out_w) {
        int64_t blocks = (n + 255) / 256;
        const int64_t cap = (int64_t)sm_count() * 32;
        if (blocks > cap) blocks = cap;
        k_gather_w<<<(unsigned)blocks, 256, 0, s>>>(w, rows, n, out_w);
        C3VQ_LAUNCH_CHECK("k_gather_w");
    }
    return C3VQ_OK;
}

// ------------------------------------------------------------------------------------------------
// join_features (compression/vq.py:90-103): the step right after the clustering.
//   out_features = cat(codebook [k, d], all_features[keep_mask])            (boolean-mask compaction, stable)
//   out_indices[i] = k + rank_keep(i)                        if keep_mask[i]
//                  = codebook_indices[i - rank_keep(i)]      otherwise (the i-th non-kept row's codeword)
// One pass over the mask for block counts, a scan of the counts, one pass that ranks every row inside its block
// (warp ballots) and moves the kept rows with lanes over columns.  Integer / copy work: bit-exact.
constexpr int JOIN_ROWS = 1024;     // rows per block (one per thread)

__global__ void __launch_bounds__(JOIN_ROWS)
k_join_count(const uint8_t *__restrict__ keep, int64_t n, unsigned int *__restrict__ block_counts)
{
    const int64_t i = (int64_t)blockIdx.x * JOIN_ROWS + threadIdx.x;
    const int kp = (i < n && keep[i]) ? 1 : 0;
    const int c = __syncthreads_count(kp);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = (unsigned int)c;
}

// exclusive scan of the block counts in place (one CTA); total kept rows -> counts[nb]
__global__ void __launch_bounds__(1024)
k_join_scan(unsigned int *__restrict__ counts, int nb)
{
    __shared__ unsigned int part[1024];
    const int per = (nb + 1023) / 1024;
    const int j0 = threadIdx.x * per;
    unsigned int sum = 0;
    for (int j = j0; j < min(nb, j0 + per); j++) sum += counts[j];
    part[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const unsigned int v = threadIdx.x >= o ? part[threadIdx.x - o] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned int run = part[threadIdx.x] - sum;
    for (int j = j0; j < min(nb, j0 + per); j++) {
        const unsigned int c = counts[j];
        counts[j] = run;
        run += c;
    }
    if (threadIdx.x == 1023) counts[nb] = part[1023];
}

__global__ void __launch_bounds__(JOIN_ROWS)
k_join_write(const float *__restrict__ feats, const uint8_t *__restrict__ keep, int64_t n, int d, int64_t k,
             const int64_t *__restrict__ cb_idx, const unsigned int *__restrict__ block_base,
             float *__restrict__ out_feats, int64_t *__restrict__ out_idx)
{
    __shared__ unsigned int warp_cnt[32];
    __shared__ unsigned int src_row[JOIN_ROWS];             // block-local list of kept rows (in order)
    __shared__ unsigned int s_kept;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * JOIN_ROWS + threadIdx.x;
    const bool kp = i < n && keep[i];
    const unsigned int bal = __ballot_sync(0xffffffffu, kp);
    if (lane == 0) warp_cnt[wib] = __popc(bal);
    __syncthreads();
    if (wib == 0) {                                         // exclusive scan of the 32 warp counts
        unsigned int v = warp_cnt[lane], incl = v;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        warp_cnt[lane] = incl - v;
        if (lane == 31) s_kept = incl;
    }
    __syncthreads();
    const unsigned int local = warp_cnt[wib] + __popc(bal & ((1u << lane) - 1u));   // kept rows before i in the block
    const unsigned int base = block_base[blockIdx.x];       // kept rows before the block
    if (i < n) {
        if (kp) {
            out_idx[i] = k + (int64_t)base + local;
            src_row[local] = threadIdx.x;
        } else {
            out_idx[i] = cb_idx[i - ((int64_t)base + local)];
        }
    }
    __syncthreads();
    // move the kept rows: the block's rows land contiguously at out_feats[(k + base) * d ...]
    const unsigned int kept = s_kept;
    const int64_t total = (int64_t)kept * d;
    float *dst = out_feats + (k + (int64_t)base) * d;
    const float *srcb = feats + (int64_t)blockIdx.x * JOIN_ROWS * d;
    if (d <= (1 << 20)) {                                   // 32-bit index arithmetic (one division per word moved)
        const unsigned int total32 = (unsigned int)total, d32 = (unsigned int)d;
        for (unsigned int e = threadIdx.x; e < total32; e += JOIN_ROWS) {
            const unsigned int r = e / d32;
            dst[e] = srcb[src_row[r] * d32 + (e - r * d32)];
        }
    } else {
        for (int64_t e = threadIdx.x; e < total; e += JOIN_ROWS) {
            const unsigned int r = (unsigned int)(e / d);
            const int c = (int)(e - (int64_t)r * d);
            dst[e] = srcb[(int64_t)src_row[r] * d + c];
        }
    }
}

size_t join_workspace_bytes(int64_t n)
{
    const int64_t nb = (n + JOIN_ROWS - 1) / JOIN_ROWS;
    return (size_t)(nb + 1) * sizeof(unsigned int);
}

int launch_join_count(const uint8_t *keep, int64_t n, void *ws, size_t ws_bytes, cudaStream_t s)
{
    if (ws == nullptr || ws_bytes < join_workspace_bytes(n)) {
        set_error("workspace too small: %zu < %zu", ws_bytes, join_workspace_bytes(n));
        return C3VQ_EWORKSPACE;
    }
    unsigned int *counts = (unsigned int *)ws;
    const int64_t nb = (n + JOIN_ROWS - 1) / JOIN_ROWS;
    C3VQ_CHECK_ARG(nb < 0x7fffffffLL, "n=%lld too large", (long long)n);
    if (nb == 0) {
        C3VQ_CUDA(cudaMemsetAsync(counts, 0, sizeof(unsigned int), s));
        return C3VQ_OK;
    }
    k_join_count<<<(unsigned)nb, JOIN_ROWS, 0, s>>>(keep, n, counts);
    C3VQ_LAUNCH_CHECK("k_join_count");
    k_join_scan<<<1, 1024, 0, s>>>(counts, (int)nb);
    C3VQ_LAUNCH_CHECK("k_join_scan");
    return C3VQ_OK;
}

int launch_join_write(const float *feats, const uint8_t *keep, int64_t n, int d, const float *codebook, int64_t k,
                      const int64_t *cb_idx, float *out_feats, int64_t *out_idx, const void *ws, cudaStream_t s)
{
    if (k > 0 && d > 0)
        C3VQ_CUDA(cudaMemcpyAsync(out_feats, codebook, (size_t)k * d * sizeof(float), cudaMemcpyDeviceToDevice, s));
    const int64_t nb = (n + JOIN_ROWS - 1) / JOIN_ROWS;
    if (nb == 0) return C3VQ_OK;
    k_join_write<<<(unsigned)nb, JOIN_ROWS, 0, s>>>(feats, keep, n, d, k, cb_idx, (const unsigned int *)ws, out_feats,
                                                     out_idx);
    C3VQ_LAUNCH_CHECK("k_join_write");
    return C3VQ_OK;
}

}  // namespace c3vq
