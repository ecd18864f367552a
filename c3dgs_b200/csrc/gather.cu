// gather.cu — indexed-codebook decode and mini-batch row gather.
//
// Replaces the torch advanced-indexing gathers `codebook[idx]`
// (compression/vq.py:40-42; scene/gaussian_model.py:220, :240, :257, :893-897) and
// `features[batch]` (compression/vq.py:70).  HBM bound: algorithmic bytes per output
// row = 8 (idx, i64) + 4*d (row written); the table is L2 resident for codebooks.
// Rows whose byte length is a multiple of 16 move as float4 with fully coalesced
// stores; other widths move one float per thread, still with coalesced stores.
#include "common.cuh"

namespace c3vq {

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
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
        const int i = e / dv;
        const int c = e - i * dv;
        const int64_t j = idx[i];
        T v = T();
        if ((uint64_t)j < (uint64_t)table_rows) v = table[j * dv + c];
        out[e] = v;
    }
}

template <typename T>
static int launch_gather_t(const T *table, int64_t table_rows, int dv, const int64_t *idx, int64_t n, T *out,
                           cudaStream_t s)
{
    const int64_t total = n * dv;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 64;
    if (blocks > cap) blocks = cap;
    if (total < 0x7fffffffLL)
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
    if (v4)
        return launch_gather_t<float4>(reinterpret_cast<const float4 *>(table), table_rows, d / 4, idx, n,
                                       reinterpret_cast<float4 *>(out), s);
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

}  // namespace c3vq
