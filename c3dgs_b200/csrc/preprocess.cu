// preprocess.cu — the other consumer of the codebooks: the per-Gaussian preprocess of the
// reference's indexed rasteriser, which decodes scale / rotation / SH coefficients through the codebook indices
// inside the kernel (submodules/diff-gaussian-rasterization/cuda_rasterizer/forward_indexed.cu:166-272 with the
// helpers at :20-159 and auxiliary.h:42-152).  Per Gaussian: near-plane cull, projection, 3D covariance from the
// indexed (scaling, rotation) row and the per-Gaussian size factor, EWA 2D covariance + 0.3 px low-pass, conic,
// 3-sigma radius, covered tile rectangle (16 x 16 tiles), view-dependent colour from the indexed SH row, depth.
//
// One thread per Gaussian.  The per-Gaussian streams (position, factor, opacity, two indices in; 75 bytes out) are
// HBM traffic; the table rows (28 B shape, 12 * M B colour) are gathers that mostly hit the L2 and are read with
// 16-byte loads where the row alignment allows.  No graphics API is involved: this is the arithmetic that feeds the
// tile binning.
#include "common.cuh"

namespace c3vq {

namespace {

constexpr int PRE_THREADS = 256;
constexpr int PRE_TILE = 16;            // BLOCK_X = BLOCK_Y of the reference (config.h)

__constant__ float kSH_C0 = 0.28209479177387814f;
__constant__ float kSH_C1 = 0.4886025119029199f;
__constant__ float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                -1.0925484305920792f, 0.5462742152960396f};
__constant__ float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                                -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

struct PreParams {
    int64_t P;
    int deg, M;
    const float *means, *scales, *factors, *rots, *opac, *shs, *view, *proj, *campos;
    const int64_t *sh_idx, *g_idx;
    float scale_modifier, focal_x, focal_y, tan_fovx, tan_fovy;
    int W, H, prefiltered, clamp_color, sh_vec;
    unsigned int grid_x, grid_y;
    int *radii;
    float *xy, *depths, *cov3d, *rgb, *conic;
    unsigned int *tiles;
    uint8_t *clamped;
};

struct V3 { float x, y, z; };

__device__ __forceinline__ V3 sh_row(const float *__restrict__ shs, int64_t row, int M, int c)
{
    const float *p = shs + (row * M + c) * 3;
    return {__ldg(p), __ldg(p + 1), __ldg(p + 2)};
}

__global__ void __launch_bounds__(PRE_THREADS, 4) k_preprocess_indexed(const PreParams q)
{
    __shared__ float sv[16], sp[16], scam[3];
    if (threadIdx.x < 16) {
        sv[threadIdx.x] = q.view[threadIdx.x];
        sp[threadIdx.x] = q.proj[threadIdx.x];
    }
    if (threadIdx.x < 3) scam[threadIdx.x] = q.campos[threadIdx.x];
    __syncthreads();
    const int64_t idx = (int64_t)blockIdx.x * PRE_THREADS + threadIdx.x;
    if (idx >= q.P) return;

    // Every output of every Gaussian is written exactly once (the reference leaves the rows of culled Gaussians
    // untouched, i.e. whatever the caller's buffer held): zeros for whatever a culled Gaussian does not reach.
    auto zero_rest = [&]() {
        q.radii[idx] = 0;
        q.tiles[idx] = 0;
        q.depths[idx] = 0.f;
        reinterpret_cast<float2 *>(q.xy)[idx] = make_float2(0.f, 0.f);
        reinterpret_cast<float4 *>(q.conic)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        q.rgb[3 * idx] = 0.f, q.rgb[3 * idx + 1] = 0.f, q.rgb[3 * idx + 2] = 0.f;
        q.clamped[3 * idx] = 0, q.clamped[3 * idx + 1] = 0, q.clamped[3 * idx + 2] = 0;
    };

    const float px = q.means[3 * idx], py = q.means[3 * idx + 1], pz = q.means[3 * idx + 2];
    // near culling (auxiliary.h:139-152): view-space depth must exceed 0.01
    const V3 pv = {sv[0] * px + sv[4] * py + sv[8] * pz + sv[12], sv[1] * px + sv[5] * py + sv[9] * pz + sv[13],
                   sv[2] * px + sv[6] * py + sv[10] * pz + sv[14]};
    if (!q.prefiltered && pv.z <= 0.01f) {
        float2 *cz = reinterpret_cast<float2 *>(q.cov3d + idx * 6);    // 24-byte rows: 8-byte aligned
        cz[0] = make_float2(0.f, 0.f), cz[1] = make_float2(0.f, 0.f), cz[2] = make_float2(0.f, 0.f);
        zero_rest();
        return;
    }

    // projection (forward_indexed.cu:211-214)
    const float hx = sp[0] * px + sp[4] * py + sp[8] * pz + sp[12], hy = sp[1] * px + sp[5] * py + sp[9] * pz + sp[13];
    const float hw = sp[3] * px + sp[7] * py + sp[11] * pz + sp[15];
    const float p_w = 1.0f / (hw + 0.0000001f);
    const float projx = hx * p_w, projy = hy * p_w;

    // 3D covariance from the indexed shape row (forward_indexed.cu:123-159): the quaternion is used as stored
    const int64_t g = q.g_idx[idx];
    const float mod = q.factors[idx] * q.scale_modifier;
    const float s0 = mod * __ldg(q.scales + 3 * g), s1 = mod * __ldg(q.scales + 3 * g + 1), s2 = mod * __ldg(q.scales + 3 * g + 2);
    const float4 rq = __ldg(reinterpret_cast<const float4 *>(q.rots) + g);
    const float r = rq.x, x = rq.y, y = rq.z, z = rq.w;
    const float R00 = 1.f - 2.f * (y * y + z * z), R01 = 2.f * (x * y - r * z), R02 = 2.f * (x * z + r * y);
    const float R10 = 2.f * (x * y + r * z), R11 = 1.f - 2.f * (x * x + z * z), R12 = 2.f * (y * z - r * x);
    const float R20 = 2.f * (x * z - r * y), R21 = 2.f * (y * z + r * x), R22 = 1.f - 2.f * (x * x + y * y);
    // M = S R^T (row k scaled by s_k); Sigma = M^T M, summed over k = 0, 1, 2 in that order
    const float m00 = s0 * R00, m01 = s0 * R10, m02 = s0 * R20;
    const float m10 = s1 * R01, m11 = s1 * R11, m12 = s1 * R21;
    const float m20 = s2 * R02, m21 = s2 * R12, m22 = s2 * R22;
    const float c0 = m00 * m00 + m10 * m10 + m20 * m20, c1 = m00 * m01 + m10 * m11 + m20 * m21;
    const float c2 = m00 * m02 + m10 * m12 + m20 * m22, c3 = m01 * m01 + m11 * m11 + m21 * m21;
    const float c4 = m01 * m02 + m11 * m12 + m21 * m22, c5 = m02 * m02 + m12 * m12 + m22 * m22;
    float2 *cov3 = reinterpret_cast<float2 *>(q.cov3d + idx * 6);     // 24-byte rows: 8-byte aligned
    cov3[0] = make_float2(c0, c1), cov3[1] = make_float2(c2, c3), cov3[2] = make_float2(c4, c5);

    // EWA 2D covariance (forward_indexed.cu:79-121)
    const float limx = 1.3f * q.tan_fovx, limy = 1.3f * q.tan_fovy;
    const float tz = pv.z;
    const float tx = fminf(limx, fmaxf(-limx, pv.x / tz)) * tz, ty = fminf(limy, fmaxf(-limy, pv.y / tz)) * tz;
    const float j00 = q.focal_x / tz, j02 = -(q.focal_x * tx) / (tz * tz);
    const float j11 = q.focal_y / tz, j12 = -(q.focal_y * ty) / (tz * tz);
    // rows of A = J V_rot (V_rot[r][c] = view[4 c + r]): a0 = j00 V[0][:] + j02 V[2][:], a1 = j11 V[1][:] + j12 V[2][:]
    const float a00 = sv[0] * j00 + sv[2] * j02, a01 = sv[4] * j00 + sv[6] * j02, a02 = sv[8] * j00 + sv[10] * j02;
    const float a10 = sv[1] * j11 + sv[2] * j12, a11 = sv[5] * j11 + sv[6] * j12, a12 = sv[9] * j11 + sv[10] * j12;
    // Sigma a
    const float b00 = c0 * a00 + c1 * a01 + c2 * a02, b01 = c1 * a00 + c3 * a01 + c4 * a02, b02 = c2 * a00 + c4 * a01 + c5 * a02;
    const float b10 = c0 * a10 + c1 * a11 + c2 * a12, b11 = c1 * a10 + c3 * a11 + c4 * a12, b12 = c2 * a10 + c4 * a11 + c5 * a12;
    const float covx = a00 * b00 + a01 * b01 + a02 * b02 + 0.3f;
    const float covy = a00 * b10 + a01 * b11 + a02 * b12;
    const float covz = a10 * b10 + a11 * b11 + a12 * b12 + 0.3f;

    const float det = covx * covz - covy * covy;
    if (det == 0.0f) {
        zero_rest();
        return;
    }
    const float det_inv = 1.f / det;
    const float mid = 0.5f * (covx + covz);
    const float disc = sqrtf(fmaxf(0.1f, mid * mid - det));
    const float my_radius = ceilf(3.f * sqrtf(fmaxf(mid + disc, mid - disc)));
    // ndc2Pix evaluates in double (auxiliary.h:42-45)
    const float pix_x = (float)((((double)projx + 1.0) * q.W - 1.0) * 0.5);
    const float pix_y = (float)((((double)projy + 1.0) * q.H - 1.0) * 0.5);
    const int rad = (int)my_radius;
    const unsigned int rminx = min(q.grid_x, (unsigned int)max(0, (int)((pix_x - rad) / PRE_TILE)));
    const unsigned int rminy = min(q.grid_y, (unsigned int)max(0, (int)((pix_y - rad) / PRE_TILE)));
    const unsigned int rmaxx = min(q.grid_x, (unsigned int)max(0, (int)((pix_x + rad + PRE_TILE - 1) / PRE_TILE)));
    const unsigned int rmaxy = min(q.grid_y, (unsigned int)max(0, (int)((pix_y + rad + PRE_TILE - 1) / PRE_TILE)));
    if ((rmaxx - rminx) * (rmaxy - rminy) == 0) {
        zero_rest();
        return;
    }

    // colour from the indexed SH row (forward_indexed.cu:20-77)
    {
        float dx = px - scam[0], dy = py - scam[1], dz = pz - scam[2];
        const float len = sqrtf(dx * dx + dy * dy + dz * dz);
        dx = dx / len, dy = dy / len, dz = dz / len;
        const int64_t row = q.sh_idx[idx];
        const int M = q.M;
        // 16 coefficients x 3 channels = 192 contiguous bytes: twelve 16-byte loads issued together when the
        // table has that layout (the row gather is the latency-bound part of this kernel), scalar loads otherwise
        float shv[48];
        const bool vec = q.sh_vec != 0;
        if (vec) {
            const float4 *rp = reinterpret_cast<const float4 *>(q.shs) + row * 12;
            const int n4 = q.deg > 2 ? 12 : (q.deg > 1 ? 7 : (q.deg > 0 ? 3 : 1));
#pragma unroll
            for (int i = 0; i < 12; i++) {
                const float4 t = i < n4 ? __ldg(rp + i) : make_float4(0.f, 0.f, 0.f, 0.f);
                shv[4 * i] = t.x, shv[4 * i + 1] = t.y, shv[4 * i + 2] = t.z, shv[4 * i + 3] = t.w;
            }
        }
        V3 sh = vec ? V3{shv[0], shv[1], shv[2]} : sh_row(q.shs, row, M, 0);
        float cr = kSH_C0 * sh.x, cg = kSH_C0 * sh.y, cb = kSH_C0 * sh.z;
#define SH_ACC(coef, k)                                                                            \
    {                                                                                              \
        const V3 s_ = vec ? V3{shv[3 * (k)], shv[3 * (k) + 1], shv[3 * (k) + 2]} : sh_row(q.shs, row, M, k); \
        const float w_ = (coef);                                                                   \
        cr += w_ * s_.x, cg += w_ * s_.y, cb += w_ * s_.z;                                         \
    }
        if (q.deg > 0) {
            SH_ACC(-kSH_C1 * dy, 1) SH_ACC(kSH_C1 * dz, 2) SH_ACC(-kSH_C1 * dx, 3)
            if (q.deg > 1) {
                const float xx = dx * dx, yy = dy * dy, zz = dz * dz, xy = dx * dy, yz = dy * dz, xz = dx * dz;
                SH_ACC(kSH_C2[0] * xy, 4) SH_ACC(kSH_C2[1] * yz, 5) SH_ACC(kSH_C2[2] * (2.0f * zz - xx - yy), 6)
                SH_ACC(kSH_C2[3] * xz, 7) SH_ACC(kSH_C2[4] * (xx - yy), 8)
                if (q.deg > 2) {
                    SH_ACC(kSH_C3[0] * dy * (3.0f * xx - yy), 9) SH_ACC(kSH_C3[1] * xy * dz, 10)
                    SH_ACC(kSH_C3[2] * dy * (4.0f * zz - xx - yy), 11)
                    SH_ACC(kSH_C3[3] * dz * (2.0f * zz - 3.0f * xx - 3.0f * yy), 12)
                    SH_ACC(kSH_C3[4] * dx * (4.0f * zz - xx - yy), 13) SH_ACC(kSH_C3[5] * dz * (xx - yy), 14)
                    SH_ACC(kSH_C3[6] * dx * (xx - 3.0f * yy), 15)
                }
            }
        }
#undef SH_ACC
        cr += 0.5f, cg += 0.5f, cb += 0.5f;
        uint8_t k0 = 0, k1 = 0, k2 = 0;
        if (q.clamp_color) {
            k0 = cr < 0, k1 = cg < 0, k2 = cb < 0;
            cr = fmaxf(cr, 0.0f), cg = fmaxf(cg, 0.0f), cb = fmaxf(cb, 0.0f);
        }
        q.clamped[3 * idx] = k0, q.clamped[3 * idx + 1] = k1, q.clamped[3 * idx + 2] = k2;
        q.rgb[3 * idx] = cr, q.rgb[3 * idx + 1] = cg, q.rgb[3 * idx + 2] = cb;
    }

    q.depths[idx] = pv.z;
    q.radii[idx] = rad;
    reinterpret_cast<float2 *>(q.xy)[idx] = make_float2(pix_x, pix_y);
    reinterpret_cast<float4 *>(q.conic)[idx] = make_float4(covz * det_inv, -covy * det_inv, covx * det_inv, q.opac[idx]);
    q.tiles[idx] = (rmaxy - rminy) * (rmaxx - rminx);
}

}  // namespace

int launch_preprocess_indexed(int64_t P, int deg, int M, const float *means, const float *scales, const float *factors,
                              float scale_modifier, const float *rots, const float *opac, const float *shs,
                              const float *view, const float *proj, const float *campos, int W, int H, float focal_x,
                              float focal_y, float tan_fovx, float tan_fovy, int prefiltered, const int64_t *sh_idx,
                              const int64_t *g_idx, int clamp_color, int *radii, float *xy, float *depths, float *cov3d,
                              float *rgb, float *conic, unsigned int *tiles, uint8_t *clamped, cudaStream_t s)
{
    if (P == 0) return C3VQ_OK;
    PreParams q;
    q.P = P, q.deg = deg, q.M = M;
    q.means = means, q.scales = scales, q.factors = factors, q.rots = rots, q.opac = opac, q.shs = shs;
    q.view = view, q.proj = proj, q.campos = campos, q.sh_idx = sh_idx, q.g_idx = g_idx;
    q.scale_modifier = scale_modifier, q.focal_x = focal_x, q.focal_y = focal_y, q.tan_fovx = tan_fovx, q.tan_fovy = tan_fovy;
    q.W = W, q.H = H, q.prefiltered = prefiltered, q.clamp_color = clamp_color;
    q.sh_vec = (M == 16 && ((uintptr_t)shs & 15) == 0) ? 1 : 0;
    q.grid_x = (unsigned int)((W + PRE_TILE - 1) / PRE_TILE), q.grid_y = (unsigned int)((H + PRE_TILE - 1) / PRE_TILE);
    q.radii = radii, q.xy = xy, q.depths = depths, q.cov3d = cov3d, q.rgb = rgb, q.conic = conic, q.tiles = tiles;
    q.clamped = clamped;
    const int64_t nb = (P + PRE_THREADS - 1) / PRE_THREADS;
    C3VQ_CHECK_ARG(nb < 0x7fffffffLL, "P=%lld too large", (long long)P);
    k_preprocess_indexed<<<(unsigned)nb, PRE_THREADS, 0, s>>>(q);
    C3VQ_LAUNCH_CHECK("k_preprocess_indexed");
    return C3VQ_OK;
}

}  // namespace c3vq
