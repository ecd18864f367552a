// morton.cu — Morton (Z-order) permutation of the Gaussians, the reordering the reference applies before it writes
// the compressed scene (GaussianModel._sort_morton, scene/gaussian_model.py:997-1004, with mortonEncode / splitBy3 at
// :1415-1432; called from save_npz, :511).
//
//   pp_min = xyz.min(0); pp_diap = xyz.max(0) - pp_min
//   q      = ((2^21 - 1) * (xyz - pp_min) / pp_diap).long()                    float32 arithmetic, truncation
//   key    = splitBy3(q[a0]) | splitBy3(q[a1]) << 1 | splitBy3(q[a2]) << 2     a = argsort(pp_diap)
//   order  = key.sort().indices
//
// Integer work after three IEEE float32 operations per coordinate, so keys are bit-exact.  The sort is a stable LSD
// radix sort, 8 passes of 8 bits over (key u64, row u32) pairs: equal keys keep their row order (torch's sort makes no
// promise for ties; its CUDA and CPU implementations both happen to be stable).
//
// The key kernel also counts the digit totals of all eight passes; every pass is then ONE kernel
// (k_radix_scatter): tiles by ticket, stable ranking of the tile's keys (every warp owns a contiguous slice and walks it
// 32 keys at a time, digit peers by ballots, so ranks follow the input order without atomics), the tile sorted by digit
// in shared memory, the per-digit offsets of earlier tiles by decoupled look-back, full-sector write-out.
// HBM-bound: 24 bytes per key and pass.
#include "common.cuh"

namespace c3vq {

namespace {

constexpr int RS_THREADS = 512;                       // 16 warps
constexpr int RS_WARPS = RS_THREADS / 32;
#ifndef C3VQ_RS_ROUNDS
#define C3VQ_RS_ROUNDS 8
#endif
constexpr int RS_ROUNDS = C3VQ_RS_ROUNDS;             // 32-key rounds per warp
constexpr int RS_TILE = RS_THREADS * RS_ROUNDS;       // 4096 keys per CTA

// order-preserving float <-> unsigned maps for atomicMin / atomicMax
__device__ __forceinline__ unsigned int f2ord(float f)
{
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned int o)
{
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

// mm[0..2] = ordered min per axis (init 0xffffffff), mm[3..5] = ordered max per axis (init 0)
__global__ void __launch_bounds__(256) k_xyz_minmax(const float *__restrict__ xyz, int64_t n, unsigned int *__restrict__ mm)
{
    unsigned int lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
    // the array as 3 n floats read 12 at a time (three float4 = four points; the cudaMalloc'ed base is 16-byte aligned
    // or the tail loop takes everything): element e belongs to axis e % 3
    const int64_t tot = 3 * n;
    const bool al16 = (reinterpret_cast<uintptr_t>(xyz) & 15) == 0;
    const int64_t n12 = al16 ? tot / 12 : 0;
    const float4 *x4 = reinterpret_cast<const float4 *>(xyz);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n12; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 a = x4[3 * i], b = x4[3 * i + 1], c = x4[3 * i + 2];
        const float v[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
#pragma unroll
        for (int j = 0; j < 12; j++) {
            const unsigned int o = f2ord(v[j]);
            lo[j % 3] = min(lo[j % 3], o);
            hi[j % 3] = max(hi[j % 3], o);
        }
    }
    for (int64_t e = n12 * 12 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < tot; e += (int64_t)gridDim.x * blockDim.x) {
        const unsigned int o = f2ord(xyz[e]);
        const int a = (int)(e % 3);
        if (a == 0) { lo[0] = min(lo[0], o); hi[0] = max(hi[0], o); }
        else if (a == 1) { lo[1] = min(lo[1], o); hi[1] = max(hi[1], o); }
        else { lo[2] = min(lo[2], o); hi[2] = max(hi[2], o); }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        lo[a] = __reduce_min_sync(0xffffffffu, lo[a]);
        hi[a] = __reduce_max_sync(0xffffffffu, hi[a]);
    }
    // one set of atomics per CTA: tens of thousands of same-address atomics were most of this kernel's time
    __shared__ unsigned int smm[6];
    if (threadIdx.x < 6) smm[threadIdx.x] = threadIdx.x < 3 ? 0xffffffffu : 0u;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            atomicMin(&smm[a], lo[a]);
            atomicMax(&smm[3 + a], hi[a]);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            atomicMin(&mm[a], smm[a]);
            atomicMax(&mm[3 + a], smm[3 + a]);
        }
    }
}

// scene/gaussian_model.py:1415-1423
__device__ __forceinline__ unsigned long long split_by_3(unsigned long long a)
{
    unsigned long long x = a & 0x1FFFFFull;
    x = (x | x << 32) & 0x1F00000000FFFFull;
    x = (x | x << 16) & 0x1F0000FF0000FFull;
    x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void __launch_bounds__(256) k_morton_keys(const float *__restrict__ xyz, int64_t n,
                                                     const unsigned int *__restrict__ mm,
                                                     unsigned long long *__restrict__ keys,
                                                     unsigned int *__restrict__ totals)
{
    // totals[p * 256 + d] = number of keys whose digit of pass p is d, for all eight radix passes, counted while the keys
    // are in registers (the multiset of keys does not change from pass to pass)
    __shared__ unsigned int cnt[8 * 256];
    for (int j = threadIdx.x; j < 8 * 256; j += blockDim.x) cnt[j] = 0;
    __syncthreads();
    float mn[3], dp[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        mn[a] = ord2f(mm[a]);
        dp[a] = __fsub_rn(ord2f(mm[3 + a]), mn[a]);                      // :999-1000
    }
    // stable ascending argsort of the three extents (:1003): rank of axis a = number of axes that sort before it
    int rk[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        rk[a] = 0;
#pragma unroll
        for (int b = 0; b < 3; b++) rk[a] += (dp[b] < dp[a] || (dp[b] == dp[a] && b < a)) ? 1 : 0;
    }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        unsigned long long key = 0;
#pragma unroll
        for (int a = 0; a < 3; a++) {
            // :1001, float32 multiply then divide, truncation toward zero; NaN (zero extent) becomes 0 bits
            const float f = __fdiv_rn(__fmul_rn(2097151.f, __fsub_rn(xyz[i * 3 + a], mn[a])), dp[a]);
            const long long q = (f == f) ? (long long)f : 0ll;
            key |= split_by_3((unsigned long long)q) << rk[a];           // axis with the smallest extent -> bit 0
        }
        keys[i] = key;
#pragma unroll
        for (int p = 0; p < 8; p++) atomicAdd(&cnt[p * 256 + ((unsigned int)(key >> (8 * p)) & 255u)], 1u);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < 8 * 256; j += blockDim.x)
        if (cnt[j]) atomicAdd(&totals[j], cnt[j]);
}

// lanes of `act` whose 8-bit digit equals this lane's: eight ballots.  match.any costs a trip per DISTINCT value of the
// warp (the middle passes, whose digits are nearly all distinct, ran at half the speed of the top-byte pass)
__device__ __forceinline__ unsigned int digit_peers(unsigned int act, unsigned int dgt)
{
    unsigned int peers = act;
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const unsigned int m = __ballot_sync(0xffffffffu, (dgt >> b) & 1u);
        peers &= ((dgt >> b) & 1u) ? m : ~m;
    }
    return peers;
}

// One kernel per pass ("one sweep"): a CTA takes the next tile by ticket, counts its digits, publishes the counts,
// and gets the number of keys of each digit in all EARLIER tiles by decoupled look-back over the published words
// (status[tile][digit] = count | AGGREGATE, later prefix-including-this-tile | INCLUSIVE): thread d walks back from
// tile - 1 until it meets an INCLUSIVE word.  Tickets are handed out in tile order, so every tile a CTA waits for has
// started: no deadlock, whatever the hardware's block scheduling order.  Round 2's first version needed a histogram
// kernel, a scan kernel and the scatter per pass (17 + 7 + 75 us at 5.8 M keys).
constexpr unsigned int RS_AGG = 1u << 30, RS_INC = 1u << 31, RS_CNT = RS_AGG - 1u;

// FIRST: the row payload is the position itself (pass 0); LAST: write int64 rows (and optionally the sorted keys).
// The tile is first sorted by digit INSIDE shared memory (stable: ranks follow the input order), then written out in
// that order: consecutive threads then hold consecutive keys of one digit, whose global positions are consecutive, so
// the stores are full-sector runs instead of one 8-byte store per key at a position of its own (round 1's version:
// 1.16 ms for 5.8 M keys, slower than cub on as many pairs).
struct ScatterSmem {
    unsigned long long keys[RS_TILE];                  // 32 KB, digit-sorted tile
    unsigned int rows[RS_TILE];                        // 16 KB
    unsigned int wcnt[RS_WARPS][256];                  // per-warp digit counts, then per-warp running offsets (tile-local)
    unsigned int dbase[256];                           // global position of the tile's first key of each digit
    unsigned int dloc[256];                            // tile-local position of the first key of each digit
    unsigned int tile;                                 // the ticket
};

// CTAs per SM the register allocation leaves room for (78 registers unbounded = one CTA of 16 warps per SM, a quarter
// of the SM's warps, with five CTA-wide barriers and a look-back per tile: latency bound)
#ifndef C3VQ_RS_MINBLOCKS
#define C3VQ_RS_MINBLOCKS 2
#endif
template <bool FIRST, bool LAST>
__global__ void __launch_bounds__(RS_THREADS, C3VQ_RS_MINBLOCKS)
k_radix_scatter(const unsigned long long *__restrict__ keys_in, const unsigned int *__restrict__ rows_in, int64_t n,
                int shift, unsigned int *__restrict__ ticket, unsigned int *__restrict__ status,
                const unsigned int *__restrict__ total,
                unsigned long long *__restrict__ keys_out, unsigned int *__restrict__ rows_out,
                int64_t *__restrict__ order_out)
{
    extern __shared__ __align__(16) unsigned char scatter_smem[];
    ScatterSmem &sm = *reinterpret_cast<ScatterSmem *>(scatter_smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned int lt = (1u << lane) - 1u;
    if (threadIdx.x == 0) sm.tile = atomicAdd(ticket, 1u);
    for (int j = threadIdx.x; j < RS_WARPS * 256; j += RS_THREADS) (&sm.wcnt[0][0])[j] = 0;
    __syncthreads();
    const unsigned int tile = sm.tile;
    // warp w owns keys [base + w * 32 * RS_ROUNDS, + 32 * RS_ROUNDS) of the tile, 32 consecutive keys per round
    const int64_t tbase = (int64_t)tile * RS_TILE;
    const int64_t wbase = tbase + (int64_t)warp * (32 * RS_ROUNDS);
    const int tile_n = (int)((n - tbase) < (int64_t)RS_TILE ? (n - tbase) : (int64_t)RS_TILE);
    unsigned long long key[RS_ROUNDS];
    unsigned int row[RS_ROUNDS];
#pragma unroll
    for (int r = 0; r < RS_ROUNDS; r++) {
        const int64_t i = wbase + r * 32 + lane;
        key[r] = i < n ? keys_in[i] : ~0ull;
        row[r] = FIRST ? (unsigned int)i : (i < n ? rows_in[i] : 0u);
    }
    // rank of every key among the keys of its digit inside the warp's slice (input order), found once
    unsigned int rank[RS_ROUNDS];
#pragma unroll
    for (int r = 0; r < RS_ROUNDS; r++) {
        const int64_t i = wbase + r * 32 + lane;
        const unsigned int dgt = (unsigned int)(key[r] >> shift) & 255u;
        const unsigned int act = __ballot_sync(0xffffffffu, i < n);
        const unsigned int peers = digit_peers(act, dgt);
        unsigned int off = 0;
        if (i < n) off = sm.wcnt[warp][dgt];
        rank[r] = off + __popc(peers & lt);
        __syncwarp();
        if (i < n && (peers & lt) == 0) sm.wcnt[warp][dgt] = off + __popc(peers);   // lowest lane of each digit; rows are warp-private
        __syncwarp();
    }
    __syncthreads();
    // digit d: exclusive scan over the warps, the tile's count, and where the tile's digit-d keys start globally
    unsigned int dcount = 0;
    if (threadIdx.x < 256) {
        const int dgt = threadIdx.x;
        unsigned int run = 0;
        for (int w = 0; w < RS_WARPS; w++) {
            const unsigned int c = sm.wcnt[w][dgt];
            sm.wcnt[w][dgt] = run;
            run += c;
        }
        dcount = run;
        sm.dloc[dgt] = run;                                              // scanned below
        // publish this tile's count now; the counts of the earlier tiles are collected after the local scatter
        volatile unsigned int *st = status + (size_t)tile * 256 + dgt;
        *st = dcount | (tile == 0 ? RS_INC : RS_AGG);
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        // exclusive prefixes over the digits, 8 digits per lane: of the global totals (every CTA recomputes it) and
        // of the tile's own counts
        unsigned int t[8], c[8], s = 0, sc = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            t[j] = total[lane * 8 + j];
            c[j] = sm.dloc[lane * 8 + j];
            s += t[j];
            sc += c[j];
        }
        unsigned int incl = s, inclc = sc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int v = __shfl_up_sync(0xffffffffu, incl, o), vc = __shfl_up_sync(0xffffffffu, inclc, o);
            if (lane >= o) { incl += v; inclc += vc; }
        }
        unsigned int run = incl - s, runc = inclc - sc;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            sm.dbase[lane * 8 + j] = run;                                // + the keys of earlier tiles, below
            sm.dloc[lane * 8 + j] = runc;
            run += t[j];
            runc += c[j];
        }
    }
    __syncthreads();
    // Decoupled look-back, as early as possible (the sooner a tile publishes its inclusive prefix, the shorter the walk of
    // the tiles behind it): thread d adds up digit d's counts of the earlier tiles until it meets an inclusive prefix;
    // the other half of the CTA goes on to the local scatter meanwhile (dbase is not needed before the write-out).
    // Measured and dropped: a warp-parallel look-back (32 earlier tiles x 16 digits per warp and window, one warp
    // reduction per digit): 208 us per pass against 91 us -- the nearest inclusive prefix is one to three tiles back, and
    // the wide windows only multiply the polling traffic.
    if (threadIdx.x < 256 && tile > 0) {
        const int dgt = threadIdx.x;
        unsigned int excl = 0;
        for (int t = (int)tile - 1; t >= 0; t--) {
            volatile unsigned int *pt = status + (size_t)t * 256 + dgt;
            unsigned int v;
            do { v = *pt; } while ((v & (RS_AGG | RS_INC)) == 0u);
            excl += v & RS_CNT;
            if (v & RS_INC) break;
        }
        volatile unsigned int *st = status + (size_t)tile * 256 + dgt;
        *st = (excl + dcount) | RS_INC;
        sm.dbase[dgt] += excl;
    }
    // stable local scatter: tile-local position = start of the digit + keys of earlier warps + rank inside the warp
#pragma unroll
    for (int r = 0; r < RS_ROUNDS; r++) {
        const int64_t i = wbase + r * 32 + lane;
        if (i < n) {
            const unsigned int dgt = (unsigned int)(key[r] >> shift) & 255u;
            const unsigned int lpos = sm.dloc[dgt] + sm.wcnt[warp][dgt] + rank[r];
            sm.keys[lpos] = key[r];
            sm.rows[lpos] = row[r];
        }
    }
    __syncthreads();
    // write-out in digit order: position j of the tile goes to dbase[d] + (j - dloc[d])
#pragma unroll
    for (int r = 0; r < RS_ROUNDS; r++) {
        const int j = r * RS_THREADS + threadIdx.x;
        if (j < tile_n) {
            const unsigned long long k = sm.keys[j];
            const unsigned int dgt = (unsigned int)(k >> shift) & 255u;
            const unsigned int pos = sm.dbase[dgt] + ((unsigned int)j - sm.dloc[dgt]);
            if (LAST) {
                order_out[pos] = (int64_t)sm.rows[j];
                if (keys_out) keys_out[pos] = k;
            } else {
                keys_out[pos] = k;
                rows_out[pos] = sm.rows[j];
            }
        }
    }
}

struct MortonLayout {
    size_t off_mm, off_total, off_ticket, off_status, off_keys0, off_keys1, off_rows0, off_rows1, bytes;
    int nb;
};

MortonLayout morton_layout(int64_t n)
{
    MortonLayout l;
    l.nb = (int)((n + RS_TILE - 1) / RS_TILE);
    if (l.nb < 1) l.nb = 1;
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t o = 0;
    l.off_mm = o, o += 256;
    l.off_total = o, o += up(8 * 256 * 4);                       // digit totals of the eight passes
    l.off_ticket = o, o += 256;                                  // one ticket counter per pass
    l.off_status = o, o += up((size_t)8 * 256 * l.nb * 4);       // look-back words [pass][tile][digit] (zeroed per call)
    l.off_keys0 = o, o += up((size_t)n * 8);
    l.off_keys1 = o, o += up((size_t)n * 8);
    l.off_rows0 = o, o += up((size_t)n * 4);
    l.off_rows1 = o, o += up((size_t)n * 4);
    l.bytes = o;
    return l;
}

}  // namespace

size_t morton_workspace_bytes(int64_t n) { return morton_layout(n).bytes; }

int launch_morton_order(const float *xyz, int64_t n, int64_t *out_order, uint64_t *out_keys, void *ws, size_t ws_bytes,
                        cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    const MortonLayout l = morton_layout(n);
    if (ws_bytes < l.bytes) {
        set_error("workspace too small: %zu < %zu", ws_bytes, l.bytes);
        return C3VQ_EWORKSPACE;
    }
    uint8_t *w = (uint8_t *)ws;
    C3VQ_CHECK_ARG(n < (int64_t)RS_AGG, "n=%lld too large for the 30-bit look-back counters", (long long)n);
    unsigned int *mm = (unsigned int *)(w + l.off_mm), *total = (unsigned int *)(w + l.off_total);
    unsigned int *ticket = (unsigned int *)(w + l.off_ticket), *status = (unsigned int *)(w + l.off_status);
    unsigned long long *keys[2] = {(unsigned long long *)(w + l.off_keys0), (unsigned long long *)(w + l.off_keys1)};
    unsigned int *rows[2] = {(unsigned int *)(w + l.off_rows0), (unsigned int *)(w + l.off_rows1)};

    const unsigned int mm_init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    C3VQ_CUDA(cudaMemcpyAsync(mm, mm_init, sizeof(mm_init), cudaMemcpyHostToDevice, s));
    // totals, tickets and look-back words are contiguous: one memset
    C3VQ_CUDA(cudaMemsetAsync(total, 0, l.off_keys0 - l.off_total, s));
    int64_t gb = (n + 255) / 256;
    if (gb > (int64_t)sm_count() * 8) gb = (int64_t)sm_count() * 8;
    k_xyz_minmax<<<(unsigned)gb, 256, 0, s>>>(xyz, n, mm);
    C3VQ_LAUNCH_CHECK("k_xyz_minmax");
    k_morton_keys<<<(unsigned)gb, 256, 0, s>>>(xyz, n, mm, keys[0], total);
    C3VQ_LAUNCH_CHECK("k_morton_keys");
    constexpr int PASSES = 8;
    const size_t ssm = sizeof(ScatterSmem);
    C3VQ_CUDA(cudaFuncSetAttribute(k_radix_scatter<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm));
    C3VQ_CUDA(cudaFuncSetAttribute(k_radix_scatter<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm));
    C3VQ_CUDA(cudaFuncSetAttribute(k_radix_scatter<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm));
    for (int p = 0; p < PASSES; p++) {
        const int in = p & 1, out = in ^ 1, shift = 8 * p;
        unsigned int *tk = ticket + p, *st = status + (size_t)p * 256 * l.nb;
        const unsigned int *tot = total + p * 256;
        if (p == 0)
            k_radix_scatter<true, false><<<l.nb, RS_THREADS, ssm, s>>>(keys[in], nullptr, n, shift, tk, st, tot,
                                                                    keys[out], rows[out], nullptr);
        else if (p == PASSES - 1)
            k_radix_scatter<false, true><<<l.nb, RS_THREADS, ssm, s>>>(keys[in], rows[in], n, shift, tk, st, tot,
                                                                    (unsigned long long *)out_keys, nullptr, out_order);
        else
            k_radix_scatter<false, false><<<l.nb, RS_THREADS, ssm, s>>>(keys[in], rows[in], n, shift, tk, st, tot,
                                                                     keys[out], rows[out], nullptr);
        C3VQ_LAUNCH_CHECK("k_radix_scatter");
    }
    return C3VQ_OK;
}

}  // namespace c3vq
