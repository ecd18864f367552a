// rot_scale.cu — covariance codebook rows -> (unit quaternion, scaling): the reference's
// extract_rot_scale(to_full_cov(cov)) (utils/splats.py:7-36, called at compression/vq.py:186).
//
// The reference runs a batched LAPACK / cuSOLVER `eigh` on [n,3,3] and then ~25 small torch kernels for the
// matrix -> quaternion conversion.  Here one thread owns one row: 6 floats in, a cyclic Jacobi eigen-solver on the
// 3x3 held in registers (float32, like the reference's eigh), the quaternion conversion and 7 floats out.  The work
// per row is a few hundred flops against 52 bytes, so the kernel is bound by HBM traffic and launch latency.
//
// What is and is not determined: eigenvalues (ascending, as eigh returns them) and the product R diag(s^2) R^T are;
// the signs of individual eigenvectors are not (LAPACK's are an implementation detail), so the quaternion can differ
// from the reference's by the sign choices that leave that product unchanged.  tests/ compare through the product.
#include "common.cuh"

namespace c3vq {

namespace {

// one Jacobi rotation in the (p,q) plane; r is the remaining index.  app/aqq/apq are the 2x2 block, arp/arq the
// couplings to r, (v0p,v0q) (v1p,v1q) (v2p,v2q) the two eigenvector columns.
__device__ __forceinline__ void jacobi_rotate(float &app, float &aqq, float &apq, float &arp, float &arq, float &v0p,
                                              float &v0q, float &v1p, float &v1q, float &v2p, float &v2q)
{
    if (apq == 0.f) return;
    const float theta = (aqq - app) / (2.f * apq);
    // tan of the smaller rotation angle; theta^2 may overflow to inf, which gives t = 0 (nothing to rotate)
    const float t = copysignf(1.f, theta) / (fabsf(theta) + sqrtf(fmaf(theta, theta, 1.f)));
    const float c = 1.f / sqrtf(fmaf(t, t, 1.f));
    const float s = t * c;
    app = fmaf(-t, apq, app);
    aqq = fmaf(t, apq, aqq);
    apq = 0.f;
    float x = arp, y = arq;
    arp = c * x - s * y;
    arq = s * x + c * y;
    x = v0p, y = v0q, v0p = c * x - s * y, v0q = s * x + c * y;
    x = v1p, y = v1q, v1p = c * x - s * y, v1q = s * x + c * y;
    x = v2p, y = v2q, v2p = c * x - s * y, v2q = s * x + c * y;
}

__device__ __forceinline__ void swap_col(float &la, float &lb, float &v0a, float &v0b, float &v1a, float &v1b,
                                         float &v2a, float &v2b)
{
    if (lb < la) {
        float t;
        t = la, la = lb, lb = t;
        t = v0a, v0a = v0b, v0b = t;
        t = v1a, v1a = v1b, v1b = t;
        t = v2a, v2a = v2b, v2b = t;
    }
}

__device__ __forceinline__ float sqrt_pos(float x) { return x > 0.f ? sqrtf(x) : 0.f; }   // splats.py:103-110

constexpr int ROT_SWEEPS = 12;   // cyclic Jacobi converges quadratically; 3x3 float32 needs 4-6 sweeps

// FULL: rows are 3x3 row-major (9 floats), only the upper triangle is read (eigh UPLO="U");
// else rows are the 6 stored entries xx, xy, xz, yy, yz, zz (strip_symmetric, utils/general_utils.py:68-81)
template <bool FULL>
__global__ void __launch_bounds__(256) k_rot_scale(const float *__restrict__ cov, int64_t n, float *__restrict__ rot,
                                                   float *__restrict__ scale)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float a00, a01, a02, a11, a12, a22;
        if (FULL) {
            const float *c = cov + i * 9;
            a00 = c[0], a01 = c[1], a02 = c[2], a11 = c[4], a12 = c[5], a22 = c[8];
        } else {
            const float2 *c = reinterpret_cast<const float2 *>(cov + i * 6);
            const float2 u = c[0], v = c[1], w = c[2];
            a00 = u.x, a01 = u.y, a02 = v.x, a11 = v.y, a12 = w.x, a22 = w.y;
        }
        // splats.py:30: cov + eye * 1e-8, a float32 addition
        a00 = __fadd_rn(a00, 1e-8f), a11 = __fadd_rn(a11, 1e-8f), a22 = __fadd_rn(a22, 1e-8f);

        // scale by a power of two (exact) so that theta^2 and the products below stay inside float32's range
        const float amax = fmaxf(fmaxf(fmaxf(fabsf(a00), fabsf(a11)), fabsf(a22)),
                                 fmaxf(fmaxf(fabsf(a01), fabsf(a02)), fabsf(a12)));
        int e = 0;
        if (amax > 0.f && amax <= FLT_MAX) frexpf(amax, &e);
        const float down = ldexpf(1.f, -e);
        a00 *= down, a01 *= down, a02 *= down, a11 *= down, a12 *= down, a22 *= down;

        float v00 = 1.f, v01 = 0.f, v02 = 0.f, v10 = 0.f, v11 = 1.f, v12 = 0.f, v20 = 0.f, v21 = 0.f, v22 = 1.f;
#pragma unroll 1
        for (int sweep = 0; sweep < ROT_SWEEPS; sweep++) {
            // an off-diagonal entry that no longer changes either diagonal neighbour is numerically zero
            const float d0 = fabsf(a00), d1 = fabsf(a11), d2 = fabsf(a22);
            if (d0 + fabsf(a01) == d0 && d1 + fabsf(a01) == d1) a01 = 0.f;
            if (d0 + fabsf(a02) == d0 && d2 + fabsf(a02) == d2) a02 = 0.f;
            if (d1 + fabsf(a12) == d1 && d2 + fabsf(a12) == d2) a12 = 0.f;
            if (a01 == 0.f && a02 == 0.f && a12 == 0.f) break;
            jacobi_rotate(a00, a11, a01, a02, a12, v00, v01, v10, v11, v20, v21);   // (0,1), r = 2
            jacobi_rotate(a00, a22, a02, a01, a12, v00, v02, v10, v12, v20, v22);   // (0,2), r = 1
            jacobi_rotate(a11, a22, a12, a01, a02, v01, v02, v11, v12, v21, v22);   // (1,2), r = 0
        }
        // ascending eigenvalues, columns follow (eigh's order)
        float l0 = a00, l1 = a11, l2 = a22;
        swap_col(l0, l1, v00, v01, v10, v11, v20, v21);
        swap_col(l1, l2, v01, v02, v11, v12, v21, v22);
        swap_col(l0, l1, v00, v01, v10, v11, v20, v21);
        const float up = ldexpf(1.f, e);
        l0 *= up, l1 *= up, l2 *= up;

        // splats.py:31: sqrt, NaN (negative eigenvalue) -> 1e-6
        float s0 = sqrtf(l0), s1 = sqrtf(l1), s2 = sqrtf(l2);
        s0 = (s0 != s0) ? 1e-6f : s0;
        s1 = (s1 != s1) ? 1e-6f : s1;
        s2 = (s2 != s2) ? 1e-6f : s2;
        scale[i * 3 + 0] = s0, scale[i * 3 + 1] = s1, scale[i * 3 + 2] = s2;

        // splats.py:35: R * det(R) so that the matrix handed to the quaternion conversion has determinant +1
        const float det = v00 * (v11 * v22 - v12 * v21) - v01 * (v10 * v22 - v12 * v20) + v02 * (v10 * v21 - v11 * v20);
        const float m00 = v00 * det, m01 = v01 * det, m02 = v02 * det, m10 = v10 * det, m11 = v11 * det,
                    m12 = v12 * det, m20 = v20 * det, m21 = v21 * det, m22 = v22 * det;
        // splats.py:59-101: four candidates, take the one with the largest denominator (first maximum)
        const float q0 = sqrt_pos(1.f + m00 + m11 + m22), q1 = sqrt_pos(1.f + m00 - m11 - m22),
                    q2 = sqrt_pos(1.f - m00 + m11 - m22), q3 = sqrt_pos(1.f - m00 - m11 + m22);
        float qr, qi, qj, qk, den;
        if (q0 >= q1 && q0 >= q2 && q0 >= q3) {
            qr = q0 * q0, qi = m21 - m12, qj = m02 - m20, qk = m10 - m01, den = q0;
        } else if (q1 >= q2 && q1 >= q3) {
            qr = m21 - m12, qi = q1 * q1, qj = m10 + m01, qk = m02 + m20, den = q1;
        } else if (q2 >= q3) {
            qr = m02 - m20, qi = m10 + m01, qj = q2 * q2, qk = m12 + m21, den = q2;
        } else {
            qr = m10 - m01, qi = m20 + m02, qj = m21 + m12, qk = q3 * q3, den = q3;
        }
        const float inv = 1.f / (2.f * fmaxf(den, 0.1f));
        qr *= inv, qi *= inv, qj *= inv, qk *= inv;
        // F.normalize: x / max(|x|, 1e-12)
        const float nrm = fmaxf(sqrtf(qr * qr + qi * qi + qj * qj + qk * qk), 1e-12f);
        reinterpret_cast<float4 *>(rot)[i] = make_float4(qr / nrm, qi / nrm, qj / nrm, qk / nrm);
    }
}

}  // namespace

int launch_rot_scale(const float *cov, int64_t n, int full, float *rot, float *scale, cudaStream_t s)
{
    if (n == 0) return C3VQ_OK;
    const int64_t blocks = (n + 255) / 256;
    const int grid = (int)(blocks < (int64_t)sm_count() * 8 ? blocks : (int64_t)sm_count() * 8);
    if (full)
        k_rot_scale<true><<<grid, 256, 0, s>>>(cov, n, rot, scale);
    else
        k_rot_scale<false><<<grid, 256, 0, s>>>(cov, n, rot, scale);
    C3VQ_LAUNCH_CHECK("k_rot_scale");
    return C3VQ_OK;
}

}  // namespace c3vq
