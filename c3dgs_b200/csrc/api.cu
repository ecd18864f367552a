// api.cu — the extern "C" boundary of libc3vq.so (declared in include/c3vq.h).
// No torch types cross this file; the Python host passes raw device pointers and
// the current CUDA stream.
#include <cstdarg>
#include <cstring>
#include <atomic>

#include "common.cuh"

namespace c3vq {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

const char *get_error() { return g_err; }

static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int sm_count()
{
    static thread_local int cached_dev = -1, cached = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        cached = v;
        cached_dev = dev;
    }
    return cached;
}

static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

static int check_common(const void *x, int64_t n, int d, const void *cb, int k)
{
    C3VQ_CHECK_ARG(n >= 0, "n=%lld is negative", (long long)n);
    C3VQ_CHECK_ARG(d >= 1 && d <= C3VQ_MAX_DIM, "d=%d outside [1, %d]", d, C3VQ_MAX_DIM);
    C3VQ_CHECK_ARG(k >= 0 && k <= C3VQ_MAX_CODES, "k=%d outside [0, %d]", k, C3VQ_MAX_CODES);
    C3VQ_CHECK_ARG(n == 0 || x != nullptr, "x is null");
    C3VQ_CHECK_ARG(k == 0 || cb != nullptr, "codebook is null");
    return C3VQ_OK;
}

}  // namespace c3vq

using namespace c3vq;

#pragma GCC visibility push(default)
extern "C" {

int c3vq_version(void) { return 100; }

const char *c3vq_last_error(void) { return get_error(); }

int64_t c3vq_launch_count(void) { return (int64_t)g_launches.load(std::memory_order_relaxed); }

int c3vq_acc_stride(int d) { return acc_stride(d); }

size_t c3vq_assign_workspace_bytes(int64_t n, int d, int k) { return tc_workspace_bytes(n, d, k); }

int c3vq_assign(const float *x, const int64_t *rows, int64_t n, int d, const float *codebook, int k,
                float *out_dist, int64_t *out_idx, void *ws, size_t ws_bytes, c3vq_stream_t stream)
{
    int rc = check_common(x, n, d, codebook, k);
    if (rc) return rc;
    C3VQ_CHECK_ARG(n == 0 || out_idx != nullptr, "out_idx is null");
    if (n == 0) return C3VQ_OK;
    return launch_assign_tc(x, nullptr, rows, n, d, codebook, k, out_dist, out_idx, nullptr, nullptr, nullptr,
                            nullptr, ws, ws_bytes, (cudaStream_t)stream);
}

int c3vq_assign_static(const float *x, const int64_t *rows, int64_t n, int d, const float *codebook, int k,
                       float *out_dist, int64_t *out_idx, c3vq_stream_t stream)
{
    int rc = check_common(x, n, d, codebook, k);
    if (rc) return rc;
    C3VQ_CHECK_ARG(n == 0 || out_idx != nullptr, "out_idx is null");
    if (n == 0) return C3VQ_OK;
    return launch_assign_tc_static(x, rows, n, d, codebook, k, out_dist, out_idx, (cudaStream_t)stream);
}

int c3vq_assign_exact(const float *x, const int64_t *rows, int64_t n, int d, const float *codebook, int k,
                      float *out_dist, int64_t *out_idx, c3vq_stream_t stream)
{
    int rc = check_common(x, n, d, codebook, k);
    if (rc) return rc;
    C3VQ_CHECK_ARG(n == 0 || out_idx != nullptr, "out_idx is null");
    return launch_assign_exact(x, rows, n, d, codebook, k, out_dist, out_idx, (cudaStream_t)stream);
}

int c3vq_accumulate(const float *x, const float *w, const int64_t *rows, const int64_t *idx, int64_t n, int d,
                    int k, float *acc, c3vq_stream_t stream)
{
    int rc = check_common(x, n, d, acc, k);
    if (rc) return rc;
    C3VQ_CHECK_ARG(n == 0 || idx != nullptr, "idx is null");
    return launch_accumulate(x, w, rows, idx, n, d, k, acc, acc_stride(d), d, (cudaStream_t)stream);
}

size_t c3vq_accumulate_sorted_workspace_bytes(int64_t n, int k) { return n < 0 || k < 0 ? 0 : seg_workspace_bytes(n, k); }

int c3vq_accumulate_sorted(const float *x, const float *w, const int64_t *rows, const int64_t *idx, int64_t n, int d,
                           int k, float *acc, void *ws, size_t ws_bytes, c3vq_stream_t stream)
{
    int rc = check_common(x, n, d, acc, k);
    if (rc) return rc;
    C3VQ_CHECK_ARG(n == 0 || idx != nullptr, "idx is null");
    return launch_accumulate_sorted(x, w, rows, idx, n, d, k, acc, acc_stride(d), d, ws, ws_bytes,
                                    (cudaStream_t)stream);
}

int c3vq_ema(float *codebook, float *entry_importance, const float *acc, int k, int d, double decay, double eps,
             int scale_normalize, c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(d >= 1 && d <= C3VQ_MAX_DIM, "d=%d outside [1, %d]", d, C3VQ_MAX_DIM);
    C3VQ_CHECK_ARG(k >= 0, "k=%d is negative", k);
    C3VQ_CHECK_ARG(k == 0 || (codebook && entry_importance && acc), "null pointer");
    C3VQ_CHECK_ARG(!scale_normalize || d >= 6, "scale_normalize divides by columns 0 + 3 + 5 (vq.py:76-77): d=%d < 6", d);
    return launch_ema(codebook, entry_importance, acc, k, d, decay, eps, scale_normalize, (cudaStream_t)stream);
}

// Workspace layout shared by c3vq_update / c3vq_assign_accumulate:
//   [tensor-core workspace (stats first)] [acc: k*stride floats] [idx: n int64] [sorted-sum workspace]
size_t c3vq_update_workspace_bytes(int64_t n, int d, int k)
{
    if (n < 0 || d < 1 || k < 0) return 0;
    return align256(tc_workspace_bytes(n, d, k)) + align256((size_t)k * acc_stride(d) * sizeof(float)) +
           align256((size_t)n * sizeof(int64_t)) + align256(seg_workspace_bytes(n, k));
}

static int assign_accumulate_impl(const float *x, const float *w, const int64_t *rows, int64_t n, int d,
                                  const float *codebook, int k, int use_tensor_cores, float *acc,
                                  float *out_dist, int64_t *out_idx, void *tc_ws, size_t tc_bytes,
                                  int64_t *idx_buf, cudaStream_t s)
{
    if (n == 0 || k == 0) return C3VQ_OK;
    int64_t *idx = out_idx ? out_idx : idx_buf;
    int rc;
    if (use_tensor_cores == 2) {
        // fused: the assignment kernel adds w*x of every finished row itself (one bulk reduce per row); reads x
        // once, but is bounded by the L2's fp32 reduction rate for wide rows
        rc = launch_assign_tc(x, w, rows, n, d, codebook, k, out_dist, idx, acc, nullptr, nullptr, nullptr, tc_ws,
                              tc_bytes, s);
    } else if (use_tensor_cores) {
        // tensor-core assignment, then the atomic-free sorted segmented sum (second pass over x)
        rc = launch_assign_tc(x, nullptr, rows, n, d, codebook, k, out_dist, idx, nullptr, nullptr, nullptr, nullptr,
                              tc_ws, tc_bytes, s);
        if (rc) return rc;
        void *seg_ws = (char *)idx_buf + align256((size_t)n * sizeof(int64_t));
        rc = launch_accumulate_sorted(x, w, rows, idx, n, d, k, acc, acc_stride(d), d, seg_ws,
                                      seg_workspace_bytes(n, k), s);
    } else {
        rc = launch_assign_exact(x, rows, n, d, codebook, k, out_dist, idx, s);
        if (rc) return rc;
        rc = launch_accumulate(x, w, rows, idx, n, d, k, acc, acc_stride(d), d, s);
    }
    return rc;
}

int c3vq_assign_accumulate(const float *x, const float *w, const int64_t *rows, int64_t n, int d,
                           const float *codebook, int k, int use_tensor_cores, float *acc, float *out_dist,
                           int64_t *out_idx, void *ws, size_t ws_bytes, c3vq_stream_t stream)
{
    int rc = check_common(x, n, d, codebook, k);
    if (rc) return rc;
    C3VQ_CHECK_ARG(n == 0 || w != nullptr, "w is null");
    C3VQ_CHECK_ARG(k == 0 || acc != nullptr, "acc is null");
    if (ws_bytes < c3vq_update_workspace_bytes(n, d, k) || (ws == nullptr && ws_bytes > 0)) {
        set_error("workspace too small: %zu < %zu", ws_bytes, c3vq_update_workspace_bytes(n, d, k));
        return C3VQ_EWORKSPACE;
    }
    char *p = (char *)ws;
    const size_t tc_bytes = tc_workspace_bytes(n, d, k);
    void *tc_ws = p;
    p += align256(tc_bytes) + align256((size_t)k * acc_stride(d) * sizeof(float));
    return assign_accumulate_impl(x, w, rows, n, d, codebook, k, use_tensor_cores, acc, out_dist, out_idx, tc_ws,
                                  tc_bytes, (int64_t *)p, (cudaStream_t)stream);
}

int c3vq_update(const float *x, const float *w, const int64_t *rows, int64_t n, int d, float *codebook,
                float *entry_importance, int k, double decay, double eps, int scale_normalize,
                int use_tensor_cores, float *out_dist, int64_t *out_idx, void *ws, size_t ws_bytes,
                c3vq_stream_t stream)
{
    int rc = check_common(x, n, d, codebook, k);
    if (rc) return rc;
    C3VQ_CHECK_ARG(k == 0 || entry_importance != nullptr, "entry_importance is null");
    C3VQ_CHECK_ARG(n == 0 || w != nullptr, "w is null");
    C3VQ_CHECK_ARG(!scale_normalize || d >= 6, "scale_normalize divides by columns 0 + 3 + 5 (vq.py:76-77): d=%d < 6", d);
    if (ws_bytes < c3vq_update_workspace_bytes(n, d, k) || (ws == nullptr && ws_bytes > 0)) {
        set_error("workspace too small: %zu < %zu", ws_bytes, c3vq_update_workspace_bytes(n, d, k));
        return C3VQ_EWORKSPACE;
    }
    if (k == 0) return C3VQ_OK;
    // n == 0 is not a no-op: the reference's update on an empty batch scatters nothing and then applies both EMAs
    // (vq.py:31-34), i.e. every codeword and entry_importance decay by `decay`; the same happens here.
    cudaStream_t s = (cudaStream_t)stream;
    char *p = (char *)ws;
    const size_t tc_bytes = tc_workspace_bytes(n, d, k);
    void *tc_ws = p;
    p += align256(tc_bytes);
    const size_t acc_bytes = (size_t)k * acc_stride(d) * sizeof(float);
    float *acc = (float *)p;
    p += align256(acc_bytes);
    int64_t *idx_buf = (int64_t *)p;

    C3VQ_CUDA(cudaMemsetAsync(acc, 0, acc_bytes, s));
    rc = assign_accumulate_impl(x, w, rows, n, d, codebook, k, use_tensor_cores, acc, out_dist, out_idx, tc_ws,
                                tc_bytes, idx_buf, s);
    if (rc) return rc;
    return launch_ema(codebook, entry_importance, acc, k, d, decay, eps, scale_normalize, s);
}

int c3vq_gather(const float *table, int64_t table_rows, int d, const int64_t *idx, int64_t n, float *out,
                c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0 && d >= 0 && table_rows >= 0, "negative size");
    C3VQ_CHECK_ARG(n == 0 || d == 0 || (table && idx && out), "null pointer");
    return launch_gather(table, table_rows, d, idx, n, out, (cudaStream_t)stream);
}

size_t c3vq_gather_workspace_bytes(int64_t table_rows, int d, int64_t n)
{
    return (table_rows < 0 || d < 0 || n < 0) ? 0 : gather_workspace_bytes(table_rows, d, n);
}

int c3vq_gather_ws(const float *table, int64_t table_rows, int d, const int64_t *idx, int64_t n, float *out, void *ws,
                   size_t ws_bytes, c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0 && d >= 0 && table_rows >= 0, "negative size");
    C3VQ_CHECK_ARG(n == 0 || d == 0 || (table && idx && out), "null pointer");
    return launch_gather_ws(table, table_rows, d, idx, n, out, ws, ws_bytes, (cudaStream_t)stream);
}

int c3vq_scatter_add(const float *grad_out, const int64_t *idx, int64_t n, int d, float *grad_table,
                     int64_t table_rows, c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0 && d >= 1 && table_rows >= 0, "bad size");
    C3VQ_CHECK_ARG(n == 0 || (grad_out && idx && grad_table), "null pointer");
    return launch_accumulate(grad_out, nullptr, nullptr, idx, n, d, table_rows, grad_table, d, -1,
                             (cudaStream_t)stream);
}

size_t c3vq_join_workspace_bytes(int64_t n) { return n < 0 ? 0 : join_workspace_bytes(n); }

int c3vq_join_count(const uint8_t *keep_mask, int64_t n, int64_t *out_kept, void *ws, size_t ws_bytes,
                    c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0, "n=%lld is negative", (long long)n);
    C3VQ_CHECK_ARG(n == 0 || keep_mask != nullptr, "keep_mask is null");
    C3VQ_CHECK_ARG(out_kept != nullptr, "out_kept is null");
    cudaStream_t s = (cudaStream_t)stream;
    int rc = launch_join_count(keep_mask, n, ws, ws_bytes, s);
    if (rc) return rc;
    const int64_t nb = (n + 1023) / 1024;
    unsigned int kept = 0;
    C3VQ_CUDA(cudaMemcpyAsync(&kept, (const unsigned int *)ws + nb, sizeof(kept), cudaMemcpyDeviceToHost, s));
    C3VQ_CUDA(cudaStreamSynchronize(s));
    *out_kept = (int64_t)kept;
    return C3VQ_OK;
}

int c3vq_join_write(const float *all_features, const uint8_t *keep_mask, int64_t n, int d, const float *codebook,
                    int64_t k, const int64_t *codebook_indices, float *out_features, int64_t *out_indices,
                    const void *ws, c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0 && d >= 0 && k >= 0, "negative size");
    C3VQ_CHECK_ARG(n == 0 || (keep_mask && out_indices && ws), "null pointer");
    return launch_join_write(all_features, keep_mask, n, d, codebook, k, codebook_indices, out_features, out_indices,
                             ws, (cudaStream_t)stream);
}

int c3vq_extract_rot_scale(const float *cov, int64_t n, int full, float *out_rot, float *out_scale,
                           c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0, "n=%lld is negative", (long long)n);
    C3VQ_CHECK_ARG(n == 0 || (cov && out_rot && out_scale), "null pointer");
    C3VQ_CHECK_ARG(((uintptr_t)cov & 7) == 0 && ((uintptr_t)out_rot & 15) == 0,
                   "cov must be 8-byte and out_rot 16-byte aligned");
    return launch_rot_scale(cov, n, full, out_rot, out_scale, (cudaStream_t)stream);
}

size_t c3vq_morton_workspace_bytes(int64_t n) { return n < 0 ? 0 : morton_workspace_bytes(n); }

int c3vq_morton_order(const float *xyz, int64_t n, int64_t *out_order, uint64_t *out_keys, void *ws, size_t ws_bytes,
                      c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0 && n < ((int64_t)1 << 32), "n=%lld out of range (rows are tracked in 32 bits)", (long long)n);
    C3VQ_CHECK_ARG(n == 0 || (xyz && out_order && ws), "null pointer");
    return launch_morton_order(xyz, n, out_order, out_keys, ws, ws_bytes, (cudaStream_t)stream);
}

int c3vq_quantize_i8(const float *x, int64_t n, double scale, int64_t zero_point, int8_t *out, c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0, "n=%lld is negative", (long long)n);
    C3VQ_CHECK_ARG(n == 0 || (x && out), "null pointer");
    C3VQ_CHECK_ARG(scale > 0.0 && scale == scale, "scale must be positive");
    C3VQ_CHECK_ARG(((uintptr_t)x & 15) == 0 && ((uintptr_t)out & 3) == 0, "x must be 16-byte and out 4-byte aligned");
    return launch_quantize_i8(x, n, scale, zero_point, out, (cudaStream_t)stream);
}

int c3vq_dequantize_i8(const int8_t *q, int64_t n, float scale, int32_t zero_point, float *out, c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0, "n=%lld is negative", (long long)n);
    C3VQ_CHECK_ARG(n == 0 || (q && out), "null pointer");
    C3VQ_CHECK_ARG(((uintptr_t)q & 3) == 0 && ((uintptr_t)out & 15) == 0, "q must be 4-byte and out 16-byte aligned");
    return launch_dequantize_i8(q, n, scale, zero_point, out, (cudaStream_t)stream);
}

int c3vq_convert(int kind, const void *src, int64_t n, void *dst, c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0, "n=%lld is negative", (long long)n);
    C3VQ_CHECK_ARG(n == 0 || (src && dst), "null pointer");
    C3VQ_CHECK_ARG(((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0, "src and dst must be 16-byte aligned");
    return launch_convert(kind, src, n, dst, (cudaStream_t)stream);
}

int c3vq_build_covariance(const float *scaling, const float *rotation, int64_t n, float scaling_modifier,
                          int activate, float *out_cov, c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0, "n=%lld is negative", (long long)n);
    C3VQ_CHECK_ARG(n == 0 || (scaling && rotation && out_cov), "null pointer");
    return launch_build_cov(scaling, rotation, n, scaling_modifier, activate, out_cov, (cudaStream_t)stream);
}

int c3vq_compact_rows(const void *src, const uint8_t *keep_mask, int64_t n, int64_t row_bytes, void *out,
                      const void *ws, c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0 && row_bytes >= 0, "negative size");
    C3VQ_CHECK_ARG((row_bytes & 3) == 0 && row_bytes < ((int64_t)1 << 31), "row_bytes=%lld must be a multiple of 4",
                   (long long)row_bytes);
    C3VQ_CHECK_ARG(n == 0 || row_bytes == 0 || (src && keep_mask && out && ws), "null pointer");
    C3VQ_CHECK_ARG(((uintptr_t)src & 3) == 0 && ((uintptr_t)out & 3) == 0, "src and out must be 4-byte aligned");
    return launch_compact_rows(src, keep_mask, n, (int)(row_bytes / 4), out, ws, (cudaStream_t)stream);
}

int c3vq_preprocess_indexed(int64_t p, int sh_degree, int sh_coeffs, const float *means3d, const float *scales,
                            const float *scale_factors, float scale_modifier, const float *rotations,
                            const float *opacities, const float *shs, const float *viewmatrix, const float *projmatrix,
                            const float *cam_pos, int width, int height, float focal_x, float focal_y, float tan_fovx,
                            float tan_fovy, int prefiltered, const int64_t *sh_indices, const int64_t *g_indices,
                            int clamp_color, int32_t *out_radii, float *out_xy, float *out_depths, float *out_cov3d,
                            float *out_rgb, float *out_conic_opacity, uint32_t *out_tiles_touched, uint8_t *out_clamped,
                            c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(p >= 0, "p=%lld is negative", (long long)p);
    C3VQ_CHECK_ARG(sh_degree >= 0 && sh_degree <= 3 && sh_coeffs >= (sh_degree + 1) * (sh_degree + 1),
                   "sh_degree=%d needs %d coefficients per row, got %d", sh_degree, (sh_degree + 1) * (sh_degree + 1),
                   sh_coeffs);
    C3VQ_CHECK_ARG(width > 0 && height > 0, "bad image size %d x %d", width, height);
    C3VQ_CHECK_ARG(p == 0 || (means3d && scales && scale_factors && rotations && opacities && shs && viewmatrix &&
                              projmatrix && cam_pos && sh_indices && g_indices),
                   "null input pointer");
    C3VQ_CHECK_ARG(p == 0 || (out_radii && out_xy && out_depths && out_cov3d && out_rgb && out_conic_opacity &&
                              out_tiles_touched && out_clamped),
                   "null output pointer");
    C3VQ_CHECK_ARG(((uintptr_t)rotations & 15) == 0 && ((uintptr_t)out_conic_opacity & 15) == 0 &&
                       ((uintptr_t)out_xy & 7) == 0 && ((uintptr_t)out_cov3d & 7) == 0,
                   "rotations / out_conic_opacity must be 16-byte and out_xy / out_cov3d 8-byte aligned");
    return launch_preprocess_indexed(p, sh_degree, sh_coeffs, means3d, scales, scale_factors, scale_modifier, rotations,
                                     opacities, shs, viewmatrix, projmatrix, cam_pos, width, height, focal_x, focal_y,
                                     tan_fovx, tan_fovy, prefiltered, sh_indices, g_indices, clamp_color, out_radii,
                                     out_xy, out_depths, out_cov3d, out_rgb, out_conic_opacity, out_tiles_touched,
                                     out_clamped, (cudaStream_t)stream);
}

int c3vq_preprocess_indexed_backward(int64_t p, int sh_degree, int sh_coeffs, const float *means3d,
                                     const int32_t *radii, const float *shs, const uint8_t *clamped,
                                     const float *scales, const float *scale_factors, const float *rotations,
                                     float scale_modifier, const float *cov3ds, const float *viewmatrix,
                                     const float *projmatrix, float focal_x, float focal_y, float tan_fovx,
                                     float tan_fovy, const float *cam_pos, const float *dl_dmean2d,
                                     const float *dl_dconic, const float *dl_dcolor, float *dl_dmeans,
                                     float *dl_dcov3d, float *dl_dsh, float *dl_dscale, float *dl_dscale_factor,
                                     float *dl_drot, const int64_t *sh_indices, const int64_t *g_indices,
                                     c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(p >= 0, "p=%lld is negative", (long long)p);
    C3VQ_CHECK_ARG(sh_degree >= 0 && sh_degree <= 3 && sh_coeffs >= (sh_degree + 1) * (sh_degree + 1),
                   "sh_degree=%d needs %d coefficients per row, got %d", sh_degree, (sh_degree + 1) * (sh_degree + 1),
                   sh_coeffs);
    C3VQ_CHECK_ARG(p == 0 || (means3d && radii && cov3ds && viewmatrix && projmatrix && cam_pos && dl_dmean2d &&
                              dl_dconic && scale_factors),
                   "null input pointer");
    C3VQ_CHECK_ARG(p == 0 || shs == nullptr || (clamped && dl_dcolor && dl_dsh && sh_indices),
                   "shs given without clamped / dl_dcolor / dl_dsh / sh_indices");
    C3VQ_CHECK_ARG(p == 0 || scales == nullptr || (rotations && dl_dscale && dl_drot && g_indices),
                   "scales given without rotations / dl_dscale / dl_drot / g_indices");
    C3VQ_CHECK_ARG(p == 0 || (dl_dmeans && dl_dcov3d && dl_dscale_factor), "null output pointer");
    C3VQ_CHECK_ARG(((uintptr_t)rotations & 15) == 0 && ((uintptr_t)dl_drot & 15) == 0 && ((uintptr_t)dl_dconic & 15) == 0,
                   "rotations / dl_drot / dl_dconic must be 16-byte aligned");
    return launch_preprocess_indexed_bwd(p, sh_degree, sh_coeffs, means3d, radii, shs, clamped, scales, scale_factors,
                                         rotations, scale_modifier, cov3ds, viewmatrix, projmatrix, focal_x, focal_y,
                                         tan_fovx, tan_fovy, cam_pos, dl_dmean2d, dl_dconic, dl_dcolor, dl_dmeans,
                                         dl_dcov3d, dl_dsh, dl_dscale, dl_dscale_factor, dl_drot, sh_indices, g_indices,
                                         (cudaStream_t)stream);
}

int c3vq_gather_rows(const float *x, const float *w, const int64_t *rows, int64_t n, int d, float *out_x,
                     float *out_w, c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(n >= 0 && d >= 1, "bad size");
    C3VQ_CHECK_ARG(n == 0 || (x && rows), "null pointer");
    return launch_gather_rows(x, w, rows, n, d, out_x, out_w, (cudaStream_t)stream);
}

int c3vq_debug_scores(const float *x, int64_t n, int d, const float *codebook, int k, float *scores,
                      float *row_norm2, float *row_band, void *ws, size_t ws_bytes, c3vq_stream_t stream)
{
    int rc = check_common(x, n, d, codebook, k);
    if (rc) return rc;
    C3VQ_CHECK_ARG(scores != nullptr, "scores is null");
    if (n == 0 || k == 0) return C3VQ_OK;
    return launch_assign_tc(x, nullptr, nullptr, n, d, codebook, k, nullptr, nullptr, nullptr, scores, row_norm2,
                            row_band, ws, ws_bytes, (cudaStream_t)stream);
}

int c3vq_assign_stats(void *ws, int64_t out[3], c3vq_stream_t stream)
{
    C3VQ_CHECK_ARG(ws && out, "null pointer");
    return tc_stats(ws, out, (cudaStream_t)stream);
}

}  // extern "C"
#pragma GCC visibility pop
