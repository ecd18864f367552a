// common.cuh — shared helpers for libc3vq (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cfloat>

#include "../../include/c3vq.h"

namespace c3vq {

// thread-local last-error text (c3vq_last_error)
void set_error(const char *fmt, ...);
const char *get_error();

int sm_count();
void count_launch();  // c3vq_launch_count bookkeeping (one call per kernel launched)

#define C3VQ_CHECK_ARG(cond, ...)                 \
    do {                                          \
        if (!(cond)) {                            \
            ::c3vq::set_error(__VA_ARGS__);       \
            return C3VQ_EINVAL;                   \
        }                                         \
    } while (0)

#define C3VQ_CUDA(expr)                                                                        \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            ::c3vq::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                              __LINE__);                                                       \
            return C3VQ_ECUDA;                                                                 \
        }                                                                                      \
    } while (0)

#define C3VQ_LAUNCH_CHECK(name)                                                          \
    do {                                                                                 \
        cudaError_t _e = cudaGetLastError();                                             \
        if (_e != cudaSuccess) {                                                         \
            ::c3vq::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));  \
            return C3VQ_ECUDA;                                                           \
        }                                                                                \
        ::c3vq::count_launch();                                                          \
    } while (0)

// Accumulator row stride: d sums + 1 weight, rounded up to a multiple of 4 floats so rows are 16-byte aligned.
__host__ __device__ inline int acc_stride(int d) { return (d + 1 + 3) & ~3; }

// Padded feature width of the SIMT exact kernel (zero padding leaves the fma chain bit-identical).
inline int exact_dp(int d) { return d <= 8 ? 8 : d <= 16 ? 16 : d <= 32 ? 32 : d <= 48 ? 48 : 64; }

// ---- launchers implemented in the .cu files ----
int launch_assign_exact(const float *x, const int64_t *rows, int64_t n, int d, const float *codebook, int k,
                        float *out_dist, int64_t *out_idx, cudaStream_t s);
// wcol >= 0: column of acc that receives the weight; wcol < 0: no weight column (decode backward)
int launch_accumulate(const float *x, const float *w, const int64_t *rows, const int64_t *idx, int64_t n, int d,
                      int64_t k, float *acc, int stride, int wcol, cudaStream_t s);
// atomic-free variant: counting sort by codeword + segmented sum; ws = seg_workspace_bytes(n, k)
size_t seg_workspace_bytes(int64_t n, int k);
int launch_accumulate_sorted(const float *x, const float *w, const int64_t *rows, const int64_t *idx, int64_t n, int d,
                             int k, float *acc, int stride, int wcol, void *ws, size_t ws_bytes, cudaStream_t s);
int launch_ema(float *codebook, float *entry_importance, const float *acc, int k, int d, double decay, double eps,
               int scale_normalize, cudaStream_t s);
int launch_gather(const float *table, int64_t table_rows, int d, const int64_t *idx, int64_t n, float *out,
                  cudaStream_t s);
size_t gather_workspace_bytes(int64_t table_rows, int d, int64_t n);
int launch_gather_ws(const float *table, int64_t table_rows, int d, const int64_t *idx, int64_t n, float *out,
                     void *ws, size_t ws_bytes, cudaStream_t s);
int launch_gather_rows(const float *x, const float *w, const int64_t *rows, int64_t n, int d, float *out_x,
                       float *out_w, cudaStream_t s);

// extract_rot_scale (rot_scale.cu): cov rows (6 stored entries, or 3x3 when full != 0) -> quaternion [n,4], scaling [n,3]
int launch_rot_scale(const float *cov, int64_t n, int full, float *rot, float *scale, cudaStream_t s);

// Morton order (morton.cu): xyz [n,3] -> stable permutation sorted by the 63-bit Morton key (and the sorted keys)
size_t morton_workspace_bytes(int64_t n);
int launch_morton_order(const float *xyz, int64_t n, int64_t *out_order, uint64_t *out_keys, void *ws, size_t ws_bytes,
                        cudaStream_t s);

// join_features (gather.cu): count + scan of the keep mask (kept total at ((unsigned*)ws)[blocks]), then the write pass
size_t join_workspace_bytes(int64_t n);
int launch_join_count(const uint8_t *keep, int64_t n, void *ws, size_t ws_bytes, cudaStream_t s);
int launch_join_write(const float *feats, const uint8_t *keep, int64_t n, int d, const float *codebook, int64_t k,
                      const int64_t *cb_idx, float *out_feats, int64_t *out_idx, const void *ws, cudaStream_t s);

// wire formats and pipeline glue (wire.cu)
int launch_quantize_i8(const float *x, int64_t n, double scale, int64_t zero_point, int8_t *out, cudaStream_t s);
int launch_dequantize_i8(const int8_t *q, int64_t n, float scale, int zero_point, float *out, cudaStream_t s);
int launch_convert(int kind, const void *src, int64_t n, void *dst, cudaStream_t s);
int launch_build_cov(const float *scaling, const float *rotation, int64_t n, float modifier, int activate, float *cov,
                     cudaStream_t s);
int launch_compact_rows(const void *src, const uint8_t *keep, int64_t n, int words, void *out, const void *ws,
                        cudaStream_t s);

// indexed rasteriser preprocess (preprocess.cu)
int launch_preprocess_indexed(int64_t P, int deg, int M, const float *means, const float *scales, const float *factors,
                              float scale_modifier, const float *rots, const float *opac, const float *shs,
                              const float *view, const float *proj, const float *campos, int W, int H, float focal_x,
                              float focal_y, float tan_fovx, float tan_fovy, int prefiltered, const int64_t *sh_idx,
                              const int64_t *g_idx, int clamp_color, int *radii, float *xy, float *depths, float *cov3d,
                              float *rgb, float *conic, unsigned int *tiles, uint8_t *clamped, cudaStream_t s);

// backward of the indexed preprocess (preprocess_bwd.cu); dL_dsh / dL_dscale / dL_drot are accumulated into
int launch_preprocess_indexed_bwd(int64_t P, int deg, int M, const float *means, const int *radii, const float *shs,
                                  const uint8_t *clamped, const float *scales, const float *factors, const float *rots,
                                  float scale_modifier, const float *cov3d, const float *view, const float *proj,
                                  float focal_x, float focal_y, float tan_fovx, float tan_fovy, const float *campos,
                                  const float *dL_dmean2d, const float *dL_dconic, const float *dL_dcolor,
                                  float *dL_dmeans, float *dL_dcov3d, float *dL_dsh, float *dL_dscale, float *dL_dfactor,
                                  float *dL_drot, const int64_t *sh_idx, const int64_t *g_idx, cudaStream_t s);

// tensor-core assignment (assign_tc.cu)
size_t tc_workspace_bytes(int64_t n, int d, int k);
int launch_assign_tc(const float *x, const float *w, const int64_t *rows, int64_t n, int d, const float *codebook,
                     int k, float *out_dist, int64_t *out_idx, float *acc /* null: no fused sums */,
                     float *dbg_scores, float *dbg_norm2, float *dbg_band, void *ws, size_t ws_bytes,
                     cudaStream_t s);
int tc_stats(void *ws, int64_t out[3], cudaStream_t s);
int launch_assign_tc_static(const float *x, const int64_t *rows, int64_t n, int d, const float *codebook, int k,
                            float *out_dist, int64_t *out_idx, cudaStream_t s);

}  // namespace c3vq
