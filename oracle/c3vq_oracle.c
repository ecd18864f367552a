/*
 * c3vq_oracle.c — CPU restatement of the c3dgs vector-clustering hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (c3dgs_b200/) may
 * import, link or call this file.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it, as the checker or as
 * the reported CPU baseline.
 *
 * Parity status: PINNED against (a) golden vectors produced by importing the
 * reference's own compression/vq.py in the build container (tests/golden/,
 * generator oracle/make_golden.py) and (b) the reference CUDA kernel compiled
 * unmodified for sm_100 (oracle/_ref/, recipe oracle/Makefile) on the GPU box.
 * The torch_scatter boundary (un-vendored, unpinned dependency,
 * environment.yml:15) is restated as zero-init + in-order fp32 index-add.
 *
 * Every function cites the reference lines it follows (paths relative to the
 * reference repository root).
 *
 * Build: see oracle/Makefile (gcc -O2 -mavx2 -mfma -ffp-contract=off -fopenmp).
 * fmaf() is written explicitly; -ffp-contract=off keeps the compiler from
 * fusing anything else, -mfma makes fmaf() a single instruction.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

/* Number of codewords scanned together.  Each lane keeps its own in-order
 * accumulator, so results are identical to a scalar scan. */
#define LANES 16

ORACLE_API int c3vq_oracle_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

ORACLE_API void c3vq_oracle_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/*
 * Squared-L2 distance of one point to one codeword.
 * submodules/weighted_distance/weighted_distance.cu:9-18 — `result += diff*diff`
 * in index order; nvcc's default --fmad=true contracts it to one FFMA per
 * element (setup.py:16 passes no flags), hence fmaf().
 */
ORACLE_API float c3vq_oracle_distance(int d, const float *coefs, const float *codeword)
{
    float result = 0.f;
    for (int i = 0; i < d; i++) {
        float diff = coefs[i] - codeword[i];
        result = fmaf(diff, diff, result);
    }
    return result;
}

/*
 * Nearest-codeword search.
 * weighted_distance.cu:20-44: start at FLT_MAX, scan codewords in ascending
 * order, replace on strict `<` (first index wins ties, NaN never wins).
 * weighted_distance.cu:78-82: outputs are f32 distances and i64 indices.
 * The reference leaves min_index uninitialised when no distance is below
 * FLT_MAX (weighted_distance.cu:32); this restatement (and the CUDA product)
 * define that case as index 0, distance FLT_MAX.
 *
 * The codebook is transposed once so LANES codewords are scanned together;
 * each lane's arithmetic is the scalar chain above.
 */
ORACLE_API void c3vq_oracle_assign(const float *x, int64_t n, int d,
                                   const float *codebook, int k,
                                   float *out_dist, int64_t *out_idx)
{
    if (n <= 0) return;
    int kp = (k + LANES - 1) / LANES * LANES;
    float *cbt = NULL;
    if (k > 0) {
        cbt = (float *)aligned_alloc(64, (size_t)kp * (d > 0 ? d : 1) * sizeof(float));
        for (int c = 0; c < d; c++)
            for (int j = 0; j < kp; j++)
                cbt[(size_t)c * kp + j] = (j < k) ? codebook[(size_t)j * d + c] : 0.f;
    }
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        const float *xi = x + (size_t)i * d;
        float best = FLT_MAX;
        int64_t best_j = 0;
        for (int j0 = 0; j0 < k; j0 += LANES) {
            float acc[LANES];
            for (int l = 0; l < LANES; l++) acc[l] = 0.f;
            for (int c = 0; c < d; c++) {
                const float xc = xi[c];
                const float *col = cbt + (size_t)c * kp + j0;
#pragma omp simd
                for (int l = 0; l < LANES; l++) {
                    float diff = xc - col[l];
                    acc[l] = fmaf(diff, diff, acc[l]);
                }
            }
            int lim = (k - j0 < LANES) ? (k - j0) : LANES;
            for (int l = 0; l < lim; l++) {
                if (acc[l] < best) {
                    best = acc[l];
                    best_j = j0 + l;
                }
            }
        }
        out_dist[i] = best;
        out_idx[i] = best_j;
    }
    free(cbt);
}

/*
 * Sensitivity-weighted centroid sums.
 * compression/vq.py:31 — scatter(importance, idx, 0, "sum", dim_size=K)
 * compression/vq.py:33 — scatter(x * importance[:, None], idx, 0, "sum", dim_size=K)
 * torch_scatter semantics: zero-initialised output + scatter_add_ along dim 0.
 * Summation here is fp32 in row order (what a sequential index_add_ does); the
 * reference's CUDA atomics have no defined order, which is why codebook parity
 * is tolerance based.  `w == NULL` means unit weights (decode backward).
 */
ORACLE_API void c3vq_oracle_scatter_sum(const float *x, const float *w, const int64_t *idx,
                                        int64_t n, int d, int k,
                                        float *sum_x, float *sum_w)
{
    memset(sum_x, 0, (size_t)k * d * sizeof(float));
    if (sum_w) memset(sum_w, 0, (size_t)k * sizeof(float));
    for (int64_t i = 0; i < n; i++) {
        int64_t j = idx[i];
        float wi = w ? w[i] : 1.f;
        if (sum_w) sum_w[j] += wi;
        float *row = sum_x + (size_t)j * d;
        const float *xi = x + (size_t)i * d;
        for (int c = 0; c < d; c++) row[c] += xi[c] * wi; /* product rounded, then added */
    }
}

/* Same sums accumulated in fp64 (used by tests to bound the fp32 tolerance). */
ORACLE_API void c3vq_oracle_scatter_sum_f64(const float *x, const float *w, const int64_t *idx,
                                            int64_t n, int d, int k,
                                            double *sum_x, double *sum_w)
{
    memset(sum_x, 0, (size_t)k * d * sizeof(double));
    memset(sum_w, 0, (size_t)k * sizeof(double));
    for (int64_t i = 0; i < n; i++) {
        int64_t j = idx[i];
        double wi = w ? (double)w[i] : 1.0;
        sum_w[j] += wi;
        for (int c = 0; c < d; c++)
            sum_x[(size_t)j * d + c] += (double)(x[(size_t)i * d + c] * (w ? w[i] : 1.f));
    }
}

/*
 * Exponential moving average and centroid normalisation.
 * compression/vq.py:45-46 — moving_avg.mul_(decay).add_(new, alpha=1-decay)
 * compression/vq.py:32   — entry_importance EMA first,
 * compression/vq.py:34   — then codebook EMA with new = sum_x / (sum_w + eps).
 * decay and (1-decay) are Python doubles narrowed to f32 by torch.  torch's
 * add_(new, alpha) evaluates fma(alpha, new, self) on both its CPU (vectorised
 * fmadd) and CUDA (--fmad) kernels; checked bit-for-bit against the golden
 * fixtures, hence fmaf() here.
 */
ORACLE_API void c3vq_oracle_ema(float *codebook, float *entry_importance,
                                const float *sum_x, const float *sum_w,
                                int k, int d, double decay, double eps)
{
    const float fdecay = (float)decay;
    const float falpha = (float)(1.0 - decay);
    const float feps = (float)eps;
    for (int j = 0; j < k; j++) {
        entry_importance[j] = fmaf(falpha, sum_w[j], entry_importance[j] * fdecay);
        float denom = sum_w[j] + feps;
        for (int c = 0; c < d; c++) {
            float nw = sum_x[(size_t)j * d + c] / denom;
            float *m = codebook + (size_t)j * d + c;
            *m = fmaf(falpha, nw, (*m) * fdecay);
        }
    }
}

/*
 * Trace normalisation of 6-D upper-triangular covariances.
 * compression/vq.py:76-77 — codebook /= codebook[:, [0,3,5]].sum(-1)[:, None]
 * (torch sums the three gathered columns left to right).
 */
ORACLE_API void c3vq_oracle_trace_normalize(float *codebook, int k, int d)
{
    if (d < 6) return;
    for (int j = 0; j < k; j++) {
        float *r = codebook + (size_t)j * d;
        float tr = (r[0] + r[3]) + r[5];
        for (int c = 0; c < d; c++) r[c] = r[c] / tr;
    }
}

/*
 * One VectorQuantize.update step.  compression/vq.py:28-35.
 * Returns the per-row minimum distances like the reference does (:35).
 * scratch: caller-provided k*d + k floats.
 */
ORACLE_API void c3vq_oracle_update(const float *x, const float *w, int64_t n, int d,
                                   float *codebook, float *entry_importance, int k,
                                   double decay, double eps,
                                   float *out_dist, int64_t *out_idx, float *scratch)
{
    float *sum_x = scratch;
    float *sum_w = scratch + (size_t)k * d;
    c3vq_oracle_assign(x, n, d, codebook, k, out_dist, out_idx);
    c3vq_oracle_scatter_sum(x, w, out_idx, n, d, k, sum_x, sum_w);
    c3vq_oracle_ema(codebook, entry_importance, sum_x, sum_w, k, d, decay, eps);
}

/*
 * Indexed-codebook decode: out[i, :] = table[idx[i], :].
 * compression/vq.py:40-42 (codebook[idx]); scene/gaussian_model.py:220, :240,
 * :257, :893-897 (scaling / rotation / SH feature decode).
 */
ORACLE_API void c3vq_oracle_gather(const float *table, const int64_t *idx, int64_t n, int d,
                                   float *out)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++)
        memcpy(out + (size_t)i * d, table + (size_t)idx[i] * d, (size_t)d * sizeof(float));
}
