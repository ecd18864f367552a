/*
 * c3vq.h — C ABI of libc3vq.so, the B200 (sm_100a) implementation of c3dgs's
 * sensitivity-aware vector clustering hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types.  Every
 * pointer is a DEVICE pointer on the current CUDA device unless it says "host".
 * Every entry point returns 0 on success or a negative C3VQ_E* code, never throws,
 * and records a message readable through c3vq_last_error() (thread-local).
 * `stream` is a cudaStream_t passed as void*; work is enqueued, not synchronised.
 *
 * Reference interfaces replaced (paths relative to the c3dgs repository root):
 *   c3vq_assign        <- weightedDistanceCUDA           submodules/weighted_distance/weighted_distance.cu:60-93
 *                         (decl weighted_distance.h:3-5, binding ext.cpp:4-6, kernel :20-44)
 *   c3vq_accumulate    <- torch_scatter.scatter x2       compression/vq.py:31, :33
 *   c3vq_ema           <- ema_inplace x2 + divide        compression/vq.py:32, :34, :45-46; trace normalise :76-77
 *   c3vq_update        <- VectorQuantize.update          compression/vq.py:28-35
 *   c3vq_assign_accumulate <- vq.py:30-31,33 (update without the EMAs; multi-GPU building block)
 *   c3vq_gather        <- codebook[idx] decode           compression/vq.py:40-42; scene/gaussian_model.py:220, :240, :257, :893-897
 *   c3vq_gather_ws     <- the same decode with scratch   (aligned loads from padded copies of the table)
 *   c3vq_scatter_add   <- backward of that gather        (autograd index_put_(accumulate=True))
 *   c3vq_gather_rows   <- features[batch]                compression/vq.py:70
 */
#ifndef C3VQ_H
#define C3VQ_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define C3VQ_OK 0
#define C3VQ_EINVAL (-1)     /* bad argument (shape, null pointer, unsupported size) */
#define C3VQ_EWORKSPACE (-2) /* workspace too small */
#define C3VQ_ECUDA (-3)      /* CUDA runtime / launch error (message holds cudaGetErrorString) */

#define C3VQ_MAX_DIM 64      /* widest feature row the kernels accept (SH colour is 48) */
#define C3VQ_MAX_CODES 65536 /* candidate lists store code ids in 16 bits */

typedef void *c3vq_stream_t; /* cudaStream_t */

/* Library version (major*100 + minor) and last error text of the calling thread. */
int c3vq_version(void);
const char *c3vq_last_error(void);
/* Number of CUDA kernels this library has launched in this process (monotonic; for benchmarks). */
int64_t c3vq_launch_count(void);

/*
 * Nearest-codeword search: for each of the n rows of x[n,d] (row i is
 * x + rows[i]*d when `rows` is non-null, else x + i*d) find the codeword of
 * codebook[k,d] with the smallest squared L2 distance, where a distance is the fp32
 * chain acc = fma(x_c - cb_c, x_c - cb_c, acc), c = 0..d-1, and the FIRST strict
 * minimum in ascending codeword order wins (NaN never wins).  Results are identical,
 * bit for bit, to the reference kernel; rows with no distance below FLT_MAX give
 * index 0 / FLT_MAX (the reference leaves that index uninitialised).
 *
 * c3vq_assign runs the tensor-core path (TF32 tcgen05 screening + exact fp32
 * re-rank of every candidate inside a proven error band); c3vq_assign_exact runs
 * the SIMT in-order kernel over all k codewords.  Both give the same output.
 * out_dist may be null.  ws: c3vq_assign_workspace_bytes(n, d, k) bytes; its
 * contents need no initialisation.  The call leaves each row's winner in it and a
 * later call on the same workspace uses that as a starting bound (re-evaluated
 * against that call's codebook: whatever the workspace holds, the results are the
 * same; reusing one workspace across the iterations of a clustering is faster).
 */
size_t c3vq_assign_workspace_bytes(int64_t n, int d, int k);
int c3vq_assign(const float *x, const int64_t *rows, int64_t n, int d,
                const float *codebook, int k,
                float *out_dist, int64_t *out_idx,
                void *ws, size_t ws_bytes, c3vq_stream_t stream);
int c3vq_assign_exact(const float *x, const int64_t *rows, int64_t n, int d,
                      const float *codebook, int k,
                      float *out_dist, int64_t *out_idx, c3vq_stream_t stream);

/*
 * c3vq_assign without any device allocation on either side of the call.  No
 * counterpart in the reference: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same search, same bit-exact results as c3vq_assign;
 * the scratch lives in a fixed __device__ array of C3VQ_STATIC_SCRATCH_BYTES that
 * every GPU which loads the library holds until the process exits.  The call
 * itself runs no cudaMalloc / cudaFree / allocating library routine.
 * Limit: c3vq_assign_workspace_bytes(n, d, k) <= C3VQ_STATIC_SCRATCH_BYTES, which
 * covers n <= 65536 rows against k <= 4096 codewords at every d <= C3VQ_MAX_DIM
 * (more rows at smaller d or k); beyond it the call returns C3VQ_EWORKSPACE with
 * a message that names the sizes, and nothing is launched.
 * All calls on a device share the one array: calls that overlap on the GPU
 * (different streams, or threads that do not order their launches) race.  Calls
 * on one stream are ordered and safe; the array needs no clearing between calls
 * (each call resets the counters it reads).
 */
#define C3VQ_STATIC_SCRATCH_BYTES (4 << 20)
int c3vq_assign_static(const float *x, const int64_t *rows, int64_t n, int d,
                       const float *codebook, int k,
                       float *out_dist, int64_t *out_idx, c3vq_stream_t stream);

/*
 * Sensitivity-weighted centroid sums.  acc is [k][c3vq_acc_stride(d)] fp32:
 * columns 0..d-1 accumulate w_i * x_i, column d accumulates w_i, the rest is
 * padding.  The call ADDS into acc (zero it first; a multi-GPU caller all-reduces
 * it between c3vq_accumulate and c3vq_ema).  w == null means unit weights.
 * `rows` as above (applies to both x and w).
 */
int c3vq_acc_stride(int d);
int c3vq_accumulate(const float *x, const float *w, const int64_t *rows, const int64_t *idx,
                    int64_t n, int d, int k, float *acc, c3vq_stream_t stream);

/*
 * Same sums without per-row atomics: rows are counting-sorted by codeword, then summed segment by
 * segment in registers (the table is touched only at segment boundaries).  Several times faster for
 * wide rows, where the L2's fp32 reduction rate, not HBM, bounds c3vq_accumulate.
 * ws: c3vq_accumulate_sorted_workspace_bytes(n, k) bytes.
 */
size_t c3vq_accumulate_sorted_workspace_bytes(int64_t n, int k);
int c3vq_accumulate_sorted(const float *x, const float *w, const int64_t *rows, const int64_t *idx,
                           int64_t n, int d, int k, float *acc,
                           void *ws, size_t ws_bytes, c3vq_stream_t stream);

/*
 * EMA epilogue on the k codewords:
 *   entry_importance = decay*entry_importance + (1-decay)*sum_w
 *   codebook         = decay*codebook + (1-decay) * sum_x / (sum_w + eps)
 * then, if scale_normalize, codebook[j,:] /= codebook[j,0]+codebook[j,3]+codebook[j,5].
 * decay and eps are the Python doubles of the reference; they are narrowed to fp32
 * exactly as torch narrows them ((float)decay, (float)(1.0-decay), (float)eps).
 * scale_normalize needs d >= 6 (columns 0, 3, 5 are the trace of the stored covariance; the reference raises an
 * index error for narrower rows): C3VQ_EINVAL otherwise.
 */
int c3vq_ema(float *codebook, float *entry_importance, const float *acc,
             int k, int d, double decay, double eps, int scale_normalize, c3vq_stream_t stream);

/*
 * One VectorQuantize.update: assignment, weighted sums and EMA in one call, reading
 * x once.  out_dist / out_idx receive the per-row minimum distance and codeword
 * (either may be null).  ws: c3vq_update_workspace_bytes(n, d, k) bytes.
 * use_tensor_cores: 0 = SIMT exact assignment + atomic sums; 1 = tensor-core assignment, then the sorted
 * segmented sums (second pass over x, no per-row atomics); 2 = tensor-core assignment with the sums fused
 * into the kernel (x read from HBM once: rows of a warp that share a codeword are combined in shared memory, one
 * bulk reduce-add per distinct codeword, into 32 partial tables that are folded pairwise when n > 2^18; the host's
 * default).  The tensor-core screening runs on fp16 operands scaled by a power of two (kind::f16; environment
 * C3VQ_SCREEN=tf32 selects kind::tf32); either way the results are the exact fp32 re-rank's, bit-identical to the
 * reference kernel.
 * n == 0 is not a no-op: like the reference's update on an empty batch (vq.py:31-34: nothing scattered, both EMAs
 * applied) every codeword and entry_importance decay by `decay`.  `rows`, when given, must lie in [0, rows of x):
 * the kernels do not check them (the Python host does unless the caller vouches for them).
 * One workspace serves one stream at a time: calls that may run concurrently need their own.
 */
size_t c3vq_update_workspace_bytes(int64_t n, int d, int k);
int c3vq_update(const float *x, const float *w, const int64_t *rows, int64_t n, int d,
                float *codebook, float *entry_importance, int k,
                double decay, double eps, int scale_normalize, int use_tensor_cores,
                float *out_dist, int64_t *out_idx,
                void *ws, size_t ws_bytes, c3vq_stream_t stream);

/*
 * First half of c3vq_update for row-sharded (multi-GPU) clustering: assignment plus
 * weighted sums ADDED into acc (layout of c3vq_accumulate), no EMA.  The caller zeroes
 * acc, all-reduces it across ranks (one NCCL all-reduce of k*c3vq_acc_stride(d) floats)
 * and then calls c3vq_ema on every rank.  ws: c3vq_update_workspace_bytes(n, d, k).
 */
int c3vq_assign_accumulate(const float *x, const float *w, const int64_t *rows, int64_t n, int d,
                           const float *codebook, int k, int use_tensor_cores,
                           float *acc, float *out_dist, int64_t *out_idx,
                           void *ws, size_t ws_bytes, c3vq_stream_t stream);

/* Indexed-codebook decode: out[i,:] = table[idx[i],:], rows of d floats. */
int c3vq_gather(const float *table, int64_t table_rows, int d,
                const int64_t *idx, int64_t n, float *out, c3vq_stream_t stream);

/*
 * The same decode with scratch lent by the caller.  Rows that are not whole float4 (the 45-float SH rows) are only
 * 4-byte aligned in the table, which costs the plain call four 4-byte loads per 16 bytes written; with
 * c3vq_gather_workspace_bytes(table_rows, d, n) bytes of scratch (16-byte aligned) the call first writes four padded
 * copies of the table, one per alignment, and then decodes with aligned 16-byte loads.  The size is 0 when the
 * copies would not pay (whole-float4 rows, rows under 12 floats, tables above 16 MB of copies, n < 8 table_rows): then, and with ws null
 * or too small, the call is c3vq_gather.  Same results bit for bit; out-of-range indices decode to zeros.
 */
size_t c3vq_gather_workspace_bytes(int64_t table_rows, int d, int64_t n);
int c3vq_gather_ws(const float *table, int64_t table_rows, int d,
                   const int64_t *idx, int64_t n, float *out,
                   void *ws, size_t ws_bytes, c3vq_stream_t stream);

/* Backward of the decode: grad_table[idx[i],:] += grad_out[i,:] (grad_table zeroed by caller). */
int c3vq_scatter_add(const float *grad_out, const int64_t *idx, int64_t n, int d,
                     float *grad_table, int64_t table_rows, c3vq_stream_t stream);

/*
 * join_features (compression/vq.py:90-103), the step right after the clustering:
 *   out_features [k + kept, d] = cat(codebook [k, d], all_features[keep_mask])     (stable compaction)
 *   out_indices [n] (int64)    = k + rank among the kept rows                 where keep_mask[i] != 0
 *                              = codebook_indices[rank among the other rows]  elsewhere
 * Two calls because the output size depends on the data: c3vq_join_count scans the mask (one byte per row,
 * torch.bool layout) into ws and returns the number of kept rows (host value; synchronises `stream`), the
 * caller allocates out_features and calls c3vq_join_write with the same ws.  codebook_indices has n - kept
 * entries.  ws: c3vq_join_workspace_bytes(n) bytes.  Integer / copy work: results are bit-exact.
 */
size_t c3vq_join_workspace_bytes(int64_t n);
int c3vq_join_count(const uint8_t *keep_mask, int64_t n, int64_t *out_kept,
                    void *ws, size_t ws_bytes, c3vq_stream_t stream);
int c3vq_join_write(const float *all_features, const uint8_t *keep_mask, int64_t n, int d,
                    const float *codebook, int64_t k, const int64_t *codebook_indices,
                    float *out_features, int64_t *out_indices,
                    const void *ws, c3vq_stream_t stream);

/*
 * extract_rot_scale(to_full_cov(cov)) (utils/splats.py:7-36; called on the joined covariance table at
 * compression/vq.py:186): per row, add 1e-8 to the diagonal, eigendecompose the symmetric 3x3 (float32, upper
 * triangle, eigenvalues ascending), out_scale = sqrt(eigenvalues) with NaN -> 1e-6, out_rot = unit quaternion
 * (real part first) of eigenvectors * det.  cov rows are the 6 stored entries xx, xy, xz, yy, yz, zz when
 * full == 0 (to_full_cov is fused away) or row-major 3x3 when full != 0.  Eigenvalues and R diag(s^2) R^T agree
 * with the reference to float32 rounding (a few 1e-7 of the largest entry); the signs of individual
 * eigenvectors are an implementation detail of LAPACK and are not reproduced.
 */
int c3vq_extract_rot_scale(const float *cov, int64_t n, int full, float *out_rot /* [n,4] */,
                           float *out_scale /* [n,3] */, c3vq_stream_t stream);

/*
 * Morton (Z-order) permutation, GaussianModel._sort_morton (scene/gaussian_model.py:997-1004; mortonEncode and
 * splitBy3 at :1415-1432), applied by the reference before the compressed scene is written (:511):
 * q = ((2^21 - 1) * (xyz - min) / (max - min)).long() per axis in float32, key = bit-interleave with the axis of
 * smallest extent in bit 0, out_order = argsort(key).  Keys are bit-exact; the sort is stable (equal keys keep
 * their row order; so are the argsort of the three extents).  out_keys (optional) receives the sorted keys.
 * ws: c3vq_morton_workspace_bytes(n) bytes.  n < 2^32.
 */
size_t c3vq_morton_workspace_bytes(int64_t n);
int c3vq_morton_order(const float *xyz /* [n,3] */, int64_t n, int64_t *out_order /* [n] */,
                      uint64_t *out_keys /* [n] or null */, void *ws, size_t ws_bytes, c3vq_stream_t stream);

/*
 * Per-tensor int8 affine quantisation of the compressed scene file, GaussianModel.save_npz
 * (scene/gaussian_model.py:528-620: torch.quantize_per_tensor(x, scale, zero_point, torch.qint8).int_repr()):
 *   out[i] = clamp((int64)(nearbyint((double)x[i] / scale) + zero_point), -128, 127)     (round half to even)
 * which is the arithmetic of torch's CUDA quantiser; results are bit-exact with it.  x 16-byte aligned.
 */
int c3vq_quantize_i8(const float *x, int64_t n, double scale, int64_t zero_point, int8_t *out, c3vq_stream_t stream);

/* Its inverse, GaussianModel.load_npz (scene/gaussian_model.py:644-716): out[i] = (float)(q[i] - zero_point) * scale. */
int c3vq_dequantize_i8(const int8_t *q, int64_t n, float scale, int32_t zero_point, float *out, c3vq_stream_t stream);

/*
 * Element-type conversions of the same file format: positions are stored as float16 (gaussian_model.py:522,
 * read back :639), index arrays as int32 (:566-569, read back :723-735).
 */
#define C3VQ_F32_TO_F16 0   /* round to nearest even (Tensor.half()) */
#define C3VQ_F16_TO_F32 1
#define C3VQ_I64_TO_I32 2   /* keeps the low 32 bits (Tensor.int()) */
#define C3VQ_I32_TO_I64 3
int c3vq_convert(int kind, const void *src, int64_t n, void *dst, c3vq_stream_t stream);

/*
 * The normalised covariance the shape codebook is learnt on, GaussianModel.get_normalized_covariance(strip_sym=True)
 * (scene/gaussian_model.py:274-278 -> :55-63; utils/general_utils.py:68-119): q <- rotation / |rotation|,
 * L = R(q) diag(scaling_modifier * scaling), out_cov [n,6] = upper triangle (xx, xy, xz, yy, yz, zz) of L L^T.
 * activate != 0 applies the model's activations first (scaling <- normalize(relu(scaling)), rotation <-
 * normalize(rotation); gaussian_model.py:66, :77), i.e. the call takes the raw parameters.  float32; agrees with
 * the reference's batched matmul to rounding (sum order of three products).
 */
int c3vq_build_covariance(const float *scaling /* [n,3] */, const float *rotation /* [n,4] */, int64_t n,
                          float scaling_modifier, int activate, float *out_cov /* [n,6] */, c3vq_stream_t stream);

/*
 * Boolean-mask row compaction, t[mask] for every per-Gaussian tensor in GaussianModel.mask_splats
 * (scene/gaussian_model.py:1027-1046; the pruning step of compress_gaussians, compression/vq.py:206-210):
 * out[rank(i), :] = src[i, :] where keep_mask[i] != 0, rows of row_bytes bytes (a multiple of 4).  ws is the
 * workspace c3vq_join_count filled for the same mask (one count serves all tensors of the model).
 */
int c3vq_compact_rows(const void *src, const uint8_t *keep_mask, int64_t n, int64_t row_bytes, void *out,
                      const void *ws, c3vq_stream_t stream);

/*
 * Per-Gaussian preprocess of the indexed rasteriser — the consumer that decodes the codebooks inside the kernel
 * (submodules/diff-gaussian-rasterization/cuda_rasterizer/forward_indexed.cu:166-272, FORWARD::preprocess_indexed
 * declared at forward.h:53-81; helpers forward_indexed.cu:20-159 and auxiliary.h:42-152).  Same arguments, in the
 * reference's order, minus the two "precomputed" inputs the indexed path passes as null:
 *   scales [G,3] / rotations [G,4] are the shape table read through g_indices [P]; shs [F, sh_coeffs, 3] is the colour
 *   table read through sh_indices [P]; scale_factors, opacities, means3d are per Gaussian; viewmatrix / projmatrix
 *   (column-major 4x4) and cam_pos are device arrays like in the reference.
 * Outputs as in the reference: radii (0 = culled), screen position, view depth, world covariance (6), rgb, conic +
 * opacity, number of 16 x 16 tiles covered, per-channel clamp flags.  Every row of every output is written: a Gaussian
 * culled at the near plane (depth <= 0.01) gets zeros everywhere, one with a singular 2D covariance or no covered tile
 * gets its cov3d and zeros elsewhere (the reference leaves those rows untouched).
 * float32 throughout; agrees with the reference kernel to rounding (sum order inside the 3x3 products), see
 * tests/test_gpu_preprocess.py for the stated tolerances.  Alignment: rotations and out_conic_opacity 16 bytes,
 * out_xy and out_cov3d 8 bytes (C3VQ_EINVAL otherwise).
 */
int c3vq_preprocess_indexed(int64_t p, int sh_degree, int sh_coeffs, const float *means3d, const float *scales,
                            const float *scale_factors, float scale_modifier, const float *rotations,
                            const float *opacities, const float *shs, const float *viewmatrix, const float *projmatrix,
                            const float *cam_pos, int width, int height, float focal_x, float focal_y, float tan_fovx,
                            float tan_fovy, int prefiltered, const int64_t *sh_indices, const int64_t *g_indices,
                            int clamp_color, int32_t *out_radii, float *out_xy /* [P,2] */, float *out_depths,
                            float *out_cov3d /* [P,6] */, float *out_rgb /* [P,3] */,
                            float *out_conic_opacity /* [P,4] */, uint32_t *out_tiles_touched,
                            uint8_t *out_clamped /* [P,3] */, c3vq_stream_t stream);

/*
 * Backward of the same preprocess — where finetuning writes gradients into the two codebooks
 * (cuda_rasterizer/backward_indexed.cu:469-550 BACKWARD::preprocess_indexed, declared at backward.h:63-92: the kernels
 * computeCov2D_indexedCUDA :346-465 and preprocess_indexedCUDA :288-341 with computeColorFromSH_indexed :20-205 and
 * computeCov3D_indexed :210-283).  Arguments in the reference's order:
 *   inputs  radii / clamped / cov3ds are the forward pass' outputs; dl_dmean2d [P,3] (x, y used), dl_dconic [P,2,2]
 *           (entries 0, 1, 3 used) and dl_dcolor [P,3] come from the render backward.
 *   outputs dl_dmeans [P,3], dl_dcov3d [P,6], dl_dscale_factor [P] are written for every Gaussian (zeros where
 *           radii <= 0; the reference leaves those rows at the zeros its wrapper allocated);
 *           dl_dsh [F, sh_coeffs, 3], dl_dscale [G,3], dl_drot [G,4] are the TABLE gradients: every visible Gaussian
 *           ADDS into the rows sh_indices / g_indices name, so the caller zeroes them first
 *           (rasterize_points_indexed.cu:172-175 does the same with torch::zeros).
 * One kernel instead of the reference's two; colour rows of 16 coefficients and rotation rows are added with 16-byte
 * vector reductions.  float32; the sums over Gaussians are order dependent here as in the reference (atomics), so the
 * table gradients agree to rounding, not bit for bit (tests/test_gpu_preprocess.py states the tolerances).
 */
int c3vq_preprocess_indexed_backward(int64_t p, int sh_degree, int sh_coeffs, const float *means3d,
                                     const int32_t *radii, const float *shs, const uint8_t *clamped /* [P,3] */,
                                     const float *scales, const float *scale_factors, const float *rotations,
                                     float scale_modifier, const float *cov3ds /* [P,6] */, const float *viewmatrix,
                                     const float *projmatrix, float focal_x, float focal_y, float tan_fovx,
                                     float tan_fovy, const float *cam_pos, const float *dl_dmean2d,
                                     const float *dl_dconic, const float *dl_dcolor, float *dl_dmeans,
                                     float *dl_dcov3d, float *dl_dsh, float *dl_dscale, float *dl_dscale_factor,
                                     float *dl_drot, const int64_t *sh_indices, const int64_t *g_indices,
                                     c3vq_stream_t stream);

/* Mini-batch row gather: out_x[i,:] = x[rows[i],:], out_w[i] = w[rows[i]] (w/out_w may be null). */
int c3vq_gather_rows(const float *x, const float *w, const int64_t *rows, int64_t n, int d,
                     float *out_x, float *out_w, c3vq_stream_t stream);

/*
 * Debug / evidence: the screening scores the tensor-core path computes,
 * score[i][j] ~= |cb_j - mu|^2 - 2 (x_i - mu).(cb_j - mu), and the per-row error
 * band used to select re-rank candidates.  Small inputs only (n*k floats).
 */
int c3vq_debug_scores(const float *x, int64_t n, int d, const float *codebook, int k,
                      float *scores, float *row_norm2, float *row_band,
                      void *ws, size_t ws_bytes, c3vq_stream_t stream);

/* Counters of the last tensor-core assignment on this thread's device (host call, synchronises `stream`):
 * out[0] = rows processed, out[1] = total re-rank candidates, out[2] = rows sent to the exact fallback. */
int c3vq_assign_stats(void *ws, int64_t out[3], c3vq_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* C3VQ_H */
