// wire.cu — the data formats either side of the clustering path, sm_100a.
//
//   * per-tensor int8 affine quantisation of the compressed scene file (scene/gaussian_model.py:528-620:
//     torch.quantize_per_tensor(x, scale, zero_point, qint8).int_repr()) and its inverse on load (:644-716:
//     (q.int() - zero_point) * scale);
//   * float32 <-> float16 positions (:522, :639) and int64 <-> int32 index arrays (:566-569, :723-735);
//   * the normalised covariance the shape codebook is learnt on (gaussian_model.py:55-63, :274-278;
//     utils/general_utils.py:68-119): cov = (R diag(s)) (R diag(s))^T, upper triangle;
//   * boolean-mask row compaction (GaussianModel.mask_splats, :1027-1046), sharing join_features' block scan.
//
// All of it is streaming byte/element work bound by HBM: 16-byte loads per thread where the element type allows.
#include "common.cuh"

#include <cuda_fp16.h>

namespace c3vq {

namespace {

constexpr int WIRE_THREADS = 256;

inline unsigned wire_grid(int64_t work_items)
{
    const int64_t blocks = (work_items + WIRE_THREADS - 1) / WIRE_THREADS;
    const int64_t cap = (int64_t)sm_count() * 16;            // grid-stride: 16 CTAs of 256 threads per SM
    return (unsigned)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

// ---- int8 affine quantisation --------------------------------------------------------------------------------
// torch's CUDA kernel computes, per element, in double:  q = (int64)(nearbyint((double)x / scale) + zero_point),
// clamped to [-128, 127].  The float product t = x * fl(1 / (float)scale) is within 2^-23 |t| of the double quotient
// when the scale is a float32 value (it is: FakeQuantize keeps it in a float32 tensor), so rintf(t) is the same
// integer unless t lies within that distance of a rounding boundary; only those elements (a few per 100 k),
// non-finite ones and non-float32 scales take the double expression itself.
__device__ __forceinline__ int quantize_slow(float x, double scale, long long zp)
{
    long long q = (long long)(nearbyint((double)x / scale) + (double)zp);
    q = q < -128 ? -128 : q;
    q = q > 127 ? 127 : q;
    return (int)q;
}

__device__ __forceinline__ int quantize_one(float x, float inv_scale_f, double scale, int zp, bool fast_ok)
{
    const float t = __fmul_rn(x, inv_scale_f);
    const float r = rintf(t);
    const float f = fabsf(t - r);                            // exact for |t| < 2^23
    const float at = fabsf(t);
    // at < 1024 also rejects NaN/Inf; the margin (2^-22 |t|) is twice the product's distance from the quotient
    const bool safe = fast_ok && at < 1024.0f && (0.5f - f) > at * 2.3841858e-7f;
    if (!safe) return quantize_slow(x, scale, (long long)zp);
    int q = (int)r + zp;
    q = max(q, -128);
    return min(q, 127);
}

__global__ void __launch_bounds__(WIRE_THREADS)
k_quantize_i8(const float *__restrict__ x, int64_t n, float inv_scale_f, double scale, int zp, int fast_ok,
              int8_t *__restrict__ out)
{
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * WIRE_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * WIRE_THREADS + threadIdx.x; i < n4; i += stride) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(x) + i);
        const int q0 = quantize_one(v.x, inv_scale_f, scale, zp, fast_ok);
        const int q1 = quantize_one(v.y, inv_scale_f, scale, zp, fast_ok);
        const int q2 = quantize_one(v.z, inv_scale_f, scale, zp, fast_ok);
        const int q3 = quantize_one(v.w, inv_scale_f, scale, zp, fast_ok);
        const unsigned packed = (unsigned)(q0 & 0xff) | ((unsigned)(q1 & 0xff) << 8) | ((unsigned)(q2 & 0xff) << 16) |
                                ((unsigned)(q3 & 0xff) << 24);
        reinterpret_cast<unsigned *>(out)[i] = packed;
    }
    // tail (n not a multiple of 4)
    const int64_t t = (n4 << 2) + (int64_t)blockIdx.x * WIRE_THREADS + threadIdx.x;
    if (t < n) out[t] = (int8_t)quantize_one(x[t], inv_scale_f, scale, zp, fast_ok);
}

__global__ void __launch_bounds__(WIRE_THREADS)
k_dequantize_i8(const int8_t *__restrict__ q, int64_t n, float scale, int zp, float *__restrict__ out)
{
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * WIRE_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * WIRE_THREADS + threadIdx.x; i < n4; i += stride) {
        const unsigned p = __ldg(reinterpret_cast<const unsigned *>(q) + i);
        float4 v;
        v.x = (float)((int)(int8_t)(p & 0xff) - zp) * scale;
        v.y = (float)((int)(int8_t)((p >> 8) & 0xff) - zp) * scale;
        v.z = (float)((int)(int8_t)((p >> 16) & 0xff) - zp) * scale;
        v.w = (float)((int)(int8_t)(p >> 24) - zp) * scale;
        reinterpret_cast<float4 *>(out)[i] = v;
    }
    const int64_t t = (n4 << 2) + (int64_t)blockIdx.x * WIRE_THREADS + threadIdx.x;
    if (t < n) out[t] = (float)((int)q[t] - zp) * scale;
}

// ---- element-type conversions ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(WIRE_THREADS)
k_f32_to_f16(const float *__restrict__ x, int64_t n, __half *__restrict__ out)
{
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * WIRE_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * WIRE_THREADS + threadIdx.x; i < n4; i += stride) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(x) + i);
        const __half2 lo = __floats2half2_rn(v.x, v.y), hi = __floats2half2_rn(v.z, v.w);
        uint2 p;
        p.x = *reinterpret_cast<const unsigned *>(&lo);
        p.y = *reinterpret_cast<const unsigned *>(&hi);
        reinterpret_cast<uint2 *>(out)[i] = p;
    }
    const int64_t t = (n4 << 2) + (int64_t)blockIdx.x * WIRE_THREADS + threadIdx.x;
    if (t < n) out[t] = __float2half_rn(x[t]);
}

__global__ void __launch_bounds__(WIRE_THREADS)
k_f16_to_f32(const __half *__restrict__ x, int64_t n, float *__restrict__ out)
{
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * WIRE_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * WIRE_THREADS + threadIdx.x; i < n4; i += stride) {
        const uint2 p = __ldg(reinterpret_cast<const uint2 *>(x) + i);
        const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&p.x));
        const float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&p.y));
        reinterpret_cast<float4 *>(out)[i] = make_float4(a.x, a.y, b.x, b.y);
    }
    const int64_t t = (n4 << 2) + (int64_t)blockIdx.x * WIRE_THREADS + threadIdx.x;
    if (t < n) out[t] = __half2float(x[t]);
}

__global__ void __launch_bounds__(WIRE_THREADS)
k_i64_to_i32(const int64_t *__restrict__ x, int64_t n, int32_t *__restrict__ out)
{
    const int64_t n2 = n >> 1;
    const int64_t stride = (int64_t)gridDim.x * WIRE_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * WIRE_THREADS + threadIdx.x; i < n2; i += stride) {
        const longlong2 v = __ldg(reinterpret_cast<const longlong2 *>(x) + i);
        reinterpret_cast<int2 *>(out)[i] = make_int2((int)v.x, (int)v.y);      // .int(): keeps the low 32 bits
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && (n & 1)) out[n - 1] = (int32_t)x[n - 1];
}

__global__ void __launch_bounds__(WIRE_THREADS)
k_i32_to_i64(const int32_t *__restrict__ x, int64_t n, int64_t *__restrict__ out)
{
    const int64_t n2 = n >> 1;
    const int64_t stride = (int64_t)gridDim.x * WIRE_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * WIRE_THREADS + threadIdx.x; i < n2; i += stride) {
        const int2 v = __ldg(reinterpret_cast<const int2 *>(x) + i);
        longlong2 o;
        o.x = v.x;
        o.y = v.y;
        reinterpret_cast<longlong2 *>(out)[i] = o;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && (n & 1)) out[n - 1] = x[n - 1];
}

// ---- normalised covariance ---------------------------------------------------------------------------------------
// One thread per Gaussian; the 7 inputs and 6 outputs of 32 consecutive rows are staged through shared memory so
// the global accesses of a warp are contiguous (row strides of 12/16/24 bytes would otherwise split every access).
constexpr int COV_ROWS = 256;

__global__ void __launch_bounds__(COV_ROWS)
k_build_cov(const float *__restrict__ scaling, const float *__restrict__ rotation, int64_t n, float modifier,
            int activate, float *__restrict__ cov)
{
    __shared__ float s_in[COV_ROWS * 7];
    __shared__ float s_out[COV_ROWS * 6];
    const int64_t row0 = (int64_t)blockIdx.x * COV_ROWS;
    const int rows = (int)min((int64_t)COV_ROWS, n - row0);
    for (int e = threadIdx.x; e < rows * 3; e += COV_ROWS) s_in[e] = scaling[row0 * 3 + e];
    for (int e = threadIdx.x; e < rows * 4; e += COV_ROWS) s_in[COV_ROWS * 3 + e] = rotation[row0 * 4 + e];
    __syncthreads();
    if ((int)threadIdx.x < rows) {
        float s0 = s_in[threadIdx.x * 3 + 0], s1 = s_in[threadIdx.x * 3 + 1], s2 = s_in[threadIdx.x * 3 + 2];
        float qr = s_in[COV_ROWS * 3 + threadIdx.x * 4 + 0], qx = s_in[COV_ROWS * 3 + threadIdx.x * 4 + 1];
        float qy = s_in[COV_ROWS * 3 + threadIdx.x * 4 + 2], qz = s_in[COV_ROWS * 3 + threadIdx.x * 4 + 3];
        if (activate) {
            // scaling_activation = normalize(relu(s)), rotation_activation = normalize(q)   (gaussian_model.py:66, :77)
            s0 = fmaxf(s0, 0.0f), s1 = fmaxf(s1, 0.0f), s2 = fmaxf(s2, 0.0f);
            const float sn = fmaxf(sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(s0, s0), __fmul_rn(s1, s1)), __fmul_rn(s2, s2))), 1e-12f);
            s0 = s0 / sn, s1 = s1 / sn, s2 = s2 / sn;
            const float qn = fmaxf(sqrtf(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(qr, qr), __fmul_rn(qx, qx)), __fmul_rn(qy, qy)),
                                                   __fmul_rn(qz, qz))), 1e-12f);
            qr = qr / qn, qx = qx / qn, qy = qy / qn, qz = qz / qn;
        }
        s0 *= modifier, s1 *= modifier, s2 *= modifier;
        // build_rotation (general_utils.py:84-107) divides by the norm once more
        const float nrm = sqrtf(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(qr, qr), __fmul_rn(qx, qx)), __fmul_rn(qy, qy)),
                                          __fmul_rn(qz, qz)));
        qr = qr / nrm, qx = qx / nrm, qy = qy / nrm, qz = qz / nrm;
        // every product / sum rounded on its own, as the reference's elementwise torch ops do (no fma contraction)
#define MUL(a, b) __fmul_rn(a, b)
#define ADD(a, b) __fadd_rn(a, b)
#define SUB(a, b) __fsub_rn(a, b)
        const float r00 = SUB(1.f, MUL(2.f, ADD(MUL(qy, qy), MUL(qz, qz))));
        const float r01 = MUL(2.f, SUB(MUL(qx, qy), MUL(qr, qz)));
        const float r02 = MUL(2.f, ADD(MUL(qx, qz), MUL(qr, qy)));
        const float r10 = MUL(2.f, ADD(MUL(qx, qy), MUL(qr, qz)));
        const float r11 = SUB(1.f, MUL(2.f, ADD(MUL(qx, qx), MUL(qz, qz))));
        const float r12 = MUL(2.f, SUB(MUL(qy, qz), MUL(qr, qx)));
        const float r20 = MUL(2.f, SUB(MUL(qx, qz), MUL(qr, qy)));
        const float r21 = MUL(2.f, ADD(MUL(qy, qz), MUL(qr, qx)));
        const float r22 = SUB(1.f, MUL(2.f, ADD(MUL(qx, qx), MUL(qy, qy))));
#undef MUL
#undef ADD
#undef SUB
        // L = R diag(s); cov = L L^T, upper triangle (strip_symmetric)
        const float l00 = r00 * s0, l01 = r01 * s1, l02 = r02 * s2;
        const float l10 = r10 * s0, l11 = r11 * s1, l12 = r12 * s2;
        const float l20 = r20 * s0, l21 = r21 * s1, l22 = r22 * s2;
        float *o = s_out + threadIdx.x * 6;
        o[0] = fmaf(l02, l02, fmaf(l01, l01, l00 * l00));
        o[1] = fmaf(l02, l12, fmaf(l01, l11, l00 * l10));
        o[2] = fmaf(l02, l22, fmaf(l01, l21, l00 * l20));
        o[3] = fmaf(l12, l12, fmaf(l11, l11, l10 * l10));
        o[4] = fmaf(l12, l22, fmaf(l11, l21, l10 * l20));
        o[5] = fmaf(l22, l22, fmaf(l21, l21, l20 * l20));
    }
    __syncthreads();
    for (int e = threadIdx.x; e < rows * 6; e += COV_ROWS) cov[row0 * 6 + e] = s_out[e];
}

// ---- boolean-mask row compaction ---------------------------------------------------------------------------------
// out[rank(i), :] = src[i, :] for mask[i] != 0, rows of `words` 32-bit words; block_base = the scanned block counts
// c3vq_join_count left in ws (same 1024-row blocks as k_join_write).
constexpr int COMPACT_ROWS = 1024;

__global__ void __launch_bounds__(COMPACT_ROWS)
k_compact_rows(const uint32_t *__restrict__ src, const uint8_t *__restrict__ keep, int64_t n, int words,
               const unsigned int *__restrict__ block_base, uint32_t *__restrict__ out)
{
    __shared__ unsigned int warp_cnt[32];
    __shared__ unsigned int src_row[COMPACT_ROWS];
    __shared__ unsigned int s_kept;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * COMPACT_ROWS + threadIdx.x;
    const bool kp = i < n && keep[i];
    const unsigned int bal = __ballot_sync(0xffffffffu, kp);
    if (lane == 0) warp_cnt[wib] = __popc(bal);
    __syncthreads();
    if (wib == 0) {
        unsigned int v = warp_cnt[lane], incl = v;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        warp_cnt[lane] = incl - v;
        if (lane == 31) s_kept = incl;
    }
    __syncthreads();
    if (kp) src_row[warp_cnt[wib] + __popc(bal & ((1u << lane) - 1u))] = threadIdx.x;
    __syncthreads();
    const int64_t total = (int64_t)s_kept * words;
    uint32_t *dst = out + (int64_t)block_base[blockIdx.x] * words;
    const uint32_t *srcb = src + (int64_t)blockIdx.x * COMPACT_ROWS * words;
    if (words <= (1 << 20)) {                               // 32-bit index arithmetic (one division per word moved)
        const unsigned int total32 = (unsigned int)total, w32 = (unsigned int)words;
        for (unsigned int e = threadIdx.x; e < total32; e += COMPACT_ROWS) {
            const unsigned int r = e / w32;
            dst[e] = srcb[src_row[r] * w32 + (e - r * w32)];
        }
    } else {
        for (int64_t e = threadIdx.x; e < total; e += COMPACT_ROWS) {
            const unsigned int r = (unsigned int)(e / words);
            const int c = (int)(e - (int64_t)r * words);
            dst[e] = srcb[(int64_t)src_row[r] * words + c];
        }
    }
}

}  // namespace

int launch_quantize_i8(const float *x, int64_t n, double scale, int64_t zero_point, int8_t *out, cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    const float scale_f = (float)scale;
    const float inv_scale_f = 1.0f / scale_f;
    // the fast path needs the scale to be a float32 value, the zero point a small integer and 16-byte alignment
    const int fast_ok = ((double)scale_f == scale) && zero_point >= -1024 && zero_point <= 1024;
    const int zp = (int)(zero_point < -100000 ? -100000 : (zero_point > 100000 ? 100000 : zero_point));
    if (!fast_ok && (zero_point < -100000 || zero_point > 100000)) {
        set_error("zero_point=%lld out of range for int8 quantisation", (long long)zero_point);
        return C3VQ_EINVAL;
    }
    k_quantize_i8<<<wire_grid(n / 4 + 3), WIRE_THREADS, 0, s>>>(x, n, inv_scale_f, scale, zp, fast_ok, out);
    C3VQ_LAUNCH_CHECK("k_quantize_i8");
    return C3VQ_OK;
}

int launch_dequantize_i8(const int8_t *q, int64_t n, float scale, int zero_point, float *out, cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    k_dequantize_i8<<<wire_grid(n / 4 + 3), WIRE_THREADS, 0, s>>>(q, n, scale, zero_point, out);
    C3VQ_LAUNCH_CHECK("k_dequantize_i8");
    return C3VQ_OK;
}

int launch_convert(int kind, const void *src, int64_t n, void *dst, cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    switch (kind) {
    case 0:
        k_f32_to_f16<<<wire_grid(n / 4 + 3), WIRE_THREADS, 0, s>>>((const float *)src, n, (__half *)dst);
        break;
    case 1:
        k_f16_to_f32<<<wire_grid(n / 4 + 3), WIRE_THREADS, 0, s>>>((const __half *)src, n, (float *)dst);
        break;
    case 2:
        k_i64_to_i32<<<wire_grid(n / 2 + 1), WIRE_THREADS, 0, s>>>((const int64_t *)src, n, (int32_t *)dst);
        break;
    case 3:
        k_i32_to_i64<<<wire_grid(n / 2 + 1), WIRE_THREADS, 0, s>>>((const int32_t *)src, n, (int64_t *)dst);
        break;
    default:
        set_error("unknown conversion %d", kind);
        return C3VQ_EINVAL;
    }
    C3VQ_LAUNCH_CHECK("k_convert");
    return C3VQ_OK;
}

int launch_build_cov(const float *scaling, const float *rotation, int64_t n, float modifier, int activate, float *cov,
                     cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    const int64_t nb = (n + COV_ROWS - 1) / COV_ROWS;
    C3VQ_CHECK_ARG(nb < 0x7fffffffLL, "n=%lld too large", (long long)n);
    k_build_cov<<<(unsigned)nb, COV_ROWS, 0, s>>>(scaling, rotation, n, modifier, activate, cov);
    C3VQ_LAUNCH_CHECK("k_build_cov");
    return C3VQ_OK;
}

int launch_compact_rows(const void *src, const uint8_t *keep, int64_t n, int words, void *out, const void *ws,
                        cudaStream_t s)
{
    const int64_t nb = (n + COMPACT_ROWS - 1) / COMPACT_ROWS;
    if (nb == 0 || words == 0) return C3VQ_OK;
    k_compact_rows<<<(unsigned)nb, COMPACT_ROWS, 0, s>>>((const uint32_t *)src, keep, n, words,
                                                         (const unsigned int *)ws, (uint32_t *)out);
    C3VQ_LAUNCH_CHECK("k_compact_rows");
    return C3VQ_OK;
}

}  // namespace c3vq
