// preprocess_bwd.cu — backward of the indexed rasteriser's per-Gaussian preprocess (second
// half): where finetuning writes gradients into the codebooks.  Replaces BACKWARD::preprocess_indexed
// (submodules/diff-gaussian-rasterization/cuda_rasterizer/backward_indexed.cu:469-550), i.e. the two kernels
// computeCov2D_indexedCUDA (:346-465) and preprocess_indexedCUDA (:288-341) with their helpers
// computeColorFromSH_indexed (:20-205) and computeCov3D_indexed (:210-283; auxiliary.h:58-116).
//
// ONE kernel, one thread per Gaussian (the reference launches two and hands dL_dcov3D from the first to the second
// through HBM; here it stays in registers and is written once because it is an output of the op):
//   conic gradient -> 2D covariance -> (3D covariance gradient, view-space mean gradient) -> world mean;
//   screen-position gradient -> world mean; colour gradient -> SH table row + view direction -> world mean;
//   3D covariance gradient -> shape table row (scale, rotation) + per-Gaussian size factor.
// Gradients of the TABLE rows are sums over all Gaussians that reference the row.  The reference adds every element
// with its own atomicAdd (48 + 7 per visible Gaussian); here a colour row (16 coefficients x 3 channels = 192 bytes,
// 16-byte aligned) takes twelve red.global.add.v4.f32 and a rotation row one, so a visible Gaussian costs 16
// reductions at the L2 instead of 55.  Per-Gaussian outputs are written for every Gaussian (zeros for culled ones;
// the reference leaves those rows at the zeros its wrapper allocated).
#include "common.cuh"

namespace c3vq {

namespace {

constexpr int PB_THREADS = 256;

__constant__ float bSH_C0 = 0.28209479177387814f;
__constant__ float bSH_C1 = 0.4886025119029199f;
__constant__ float bSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                -1.0925484305920792f, 0.5462742152960396f};
__constant__ float bSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                                -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

struct PbParams {
    int64_t P;
    int deg, M, sh_vec;
    const float *means, *shs, *scales, *factors, *rots, *cov3d, *view, *proj, *campos;
    const int *radii;
    const uint8_t *clamped;
    const int64_t *sh_idx, *g_idx;
    const float *dL_dmean2d, *dL_dconic, *dL_dcolor;
    float scale_modifier, focal_x, focal_y, tan_fovx, tan_fovy;
    float *dL_dmeans, *dL_dcov3d, *dL_dsh, *dL_dscale, *dL_dfactor, *dL_drot;
};

__device__ __forceinline__ void red_v4(float *addr, float a, float b, float c, float d)
{
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

__global__ void __launch_bounds__(PB_THREADS, 3) k_preprocess_indexed_bwd(const PbParams q)
{
    __shared__ float sv[16], sp[16], scam[3];
    if (threadIdx.x < 16) {
        sv[threadIdx.x] = q.view[threadIdx.x];
        sp[threadIdx.x] = q.proj[threadIdx.x];
    }
    if (threadIdx.x < 3) scam[threadIdx.x] = q.campos[threadIdx.x];
    __syncthreads();
    const int64_t idx = (int64_t)blockIdx.x * PB_THREADS + threadIdx.x;
    if (idx >= q.P) return;
    float *gm = q.dL_dmeans + 3 * idx, *gc = q.dL_dcov3d + 6 * idx;
    if (!(q.radii[idx] > 0)) {                          // culled in the forward pass: no gradient
        gm[0] = 0.f, gm[1] = 0.f, gm[2] = 0.f;
        gc[0] = 0.f, gc[1] = 0.f, gc[2] = 0.f, gc[3] = 0.f, gc[4] = 0.f, gc[5] = 0.f;
        q.dL_dfactor[idx] = 0.f;
        return;
    }
    const float mx = q.means[3 * idx], my = q.means[3 * idx + 1], mz = q.means[3 * idx + 2];

    // ---------------- conic -> 2D covariance -> 3D covariance and view-space mean (backward_indexed.cu:346-465) ------
    float dc3[6];                                       // dL/d(cov3D), an output and the input of the shape gradients
    float dmx, dmy, dmz;                                // dL/d(mean), accumulated in the reference's order
    {
        const float *cv = q.cov3d + 6 * idx;
        const float c0 = cv[0], c1 = cv[1], c2 = cv[2], c3 = cv[3], c4 = cv[4], c5 = cv[5];
        const float4 dcon = reinterpret_cast<const float4 *>(q.dL_dconic)[idx];   // [P,2,2]: entries 0, 1, 3 are used
        float tx = sv[0] * mx + sv[4] * my + sv[8] * mz + sv[12];
        float ty = sv[1] * mx + sv[5] * my + sv[9] * mz + sv[13];
        const float tz = sv[2] * mx + sv[6] * my + sv[10] * mz + sv[14];
        const float limx = 1.3f * q.tan_fovx, limy = 1.3f * q.tan_fovy;
        const float txtz = tx / tz, tytz = ty / tz;
        tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
        ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        const float hx = q.focal_x, hy = q.focal_y;
        const float j00 = hx / tz, j02 = -(hx * tx) / (tz * tz), j11 = hy / tz, j12 = -(hy * ty) / (tz * tz);
        // T = W J: the two non-zero columns, T0 = (T[0][0..2]) and T1 = (T[1][0..2]) of the reference's glm matrix
        const float T00 = sv[0] * j00 + sv[2] * j02, T01 = sv[4] * j00 + sv[6] * j02, T02 = sv[8] * j00 + sv[10] * j02;
        const float T10 = sv[1] * j11 + sv[2] * j12, T11 = sv[5] * j11 + sv[6] * j12, T12 = sv[9] * j11 + sv[10] * j12;
        // u = Vrk T0, w = Vrk T1
        const float u0 = T00 * c0 + T01 * c1 + T02 * c2, u1 = T00 * c1 + T01 * c3 + T02 * c4, u2 = T00 * c2 + T01 * c4 + T02 * c5;
        const float w0 = T10 * c0 + T11 * c1 + T12 * c2, w1 = T10 * c1 + T11 * c3 + T12 * c4, w2 = T10 * c2 + T11 * c4 + T12 * c5;
        const float a = T00 * u0 + T01 * u1 + T02 * u2 + 0.3f;
        const float b = T00 * w0 + T01 * w1 + T02 * w2;
        const float c = T10 * w0 + T11 * w1 + T12 * w2 + 0.3f;
        const float denom = a * c - b * b;
        float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0.f) {
            dL_da = denom2inv * (-c * c * dcon.x + 2.f * b * c * dcon.y + (denom - a * c) * dcon.w);
            dL_dc = denom2inv * (-a * a * dcon.w + 2.f * a * b * dcon.y + (denom - a * c) * dcon.x);
            dL_db = denom2inv * 2.f * (b * c * dcon.x - (denom + 2.f * b * b) * dcon.y + a * b * dcon.w);
            dc3[0] = T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc;
            dc3[3] = T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc;
            dc3[5] = T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc;
            dc3[1] = 2.f * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2.f * T10 * T11 * dL_dc;
            dc3[2] = 2.f * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2.f * T10 * T12 * dL_dc;
            dc3[4] = 2.f * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2.f * T11 * T12 * dL_dc;
        } else {
#pragma unroll
            for (int i = 0; i < 6; i++) dc3[i] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 6; i++) gc[i] = dc3[i];
        const float dT00 = 2.f * u0 * dL_da + w0 * dL_db, dT01 = 2.f * u1 * dL_da + w1 * dL_db, dT02 = 2.f * u2 * dL_da + w2 * dL_db;
        const float dT10 = 2.f * w0 * dL_dc + u0 * dL_db, dT11 = 2.f * w1 * dL_dc + u1 * dL_db, dT12 = 2.f * w2 * dL_dc + u2 * dL_db;
        const float dJ00 = sv[0] * dT00 + sv[4] * dT01 + sv[8] * dT02;
        const float dJ02 = sv[2] * dT00 + sv[6] * dT01 + sv[10] * dT02;
        const float dJ11 = sv[1] * dT10 + sv[5] * dT11 + sv[9] * dT12;
        const float dJ12 = sv[2] * dT10 + sv[6] * dT11 + sv[10] * dT12;
        const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
        const float dtx = x_grad_mul * -hx * itz2 * dJ02;
        const float dty = y_grad_mul * -hy * itz2 * dJ12;
        const float dtz = -hx * itz2 * dJ00 - hy * itz2 * dJ11 + (2.f * hx * tx) * itz3 * dJ02 + (2.f * hy * ty) * itz3 * dJ12;
        dmx = sv[0] * dtx + sv[1] * dty + sv[2] * dtz;       // transformVec4x3Transpose
        dmy = sv[4] * dtx + sv[5] * dty + sv[6] * dtz;
        dmz = sv[8] * dtx + sv[9] * dty + sv[10] * dtz;
    }

    // ---------------- screen position -> mean (backward_indexed.cu:316-330) ----------------
    {
        const float hw = sp[3] * mx + sp[7] * my + sp[11] * mz + sp[15];
        const float m_w = 1.0f / (hw + 0.0000001f);
        const float mul1 = (sp[0] * mx + sp[4] * my + sp[8] * mz + sp[12]) * m_w * m_w;
        const float mul2 = (sp[1] * mx + sp[5] * my + sp[9] * mz + sp[13]) * m_w * m_w;
        const float g2x = q.dL_dmean2d[3 * idx], g2y = q.dL_dmean2d[3 * idx + 1];
        dmx += (sp[0] * m_w - sp[3] * mul1) * g2x + (sp[1] * m_w - sp[3] * mul2) * g2y;
        dmy += (sp[4] * m_w - sp[7] * mul1) * g2x + (sp[5] * m_w - sp[7] * mul2) * g2y;
        dmz += (sp[8] * m_w - sp[11] * mul1) * g2x + (sp[9] * m_w - sp[11] * mul2) * g2y;
    }

    // ---------------- colour -> SH table row and view direction (backward_indexed.cu:20-205) ----------------
    if (q.shs != nullptr) {
        const float ox = mx - scam[0], oy = my - scam[1], oz = mz - scam[2];
        const float len = sqrtf(ox * ox + oy * oy + oz * oz);
        const float x = ox / len, y = oy / len, z = oz / len;
        const int64_t row = q.sh_idx[idx];
        const int M = q.M;
        float gr = q.dL_dcolor[3 * idx], gg = q.dL_dcolor[3 * idx + 1], gb = q.dL_dcolor[3 * idx + 2];
        gr *= q.clamped[3 * idx] ? 0.f : 1.f;
        gg *= q.clamped[3 * idx + 1] ? 0.f : 1.f;
        gb *= q.clamped[3 * idx + 2] ? 0.f : 1.f;
        // the row's coefficients (needed from degree 1 on, for the direction gradient)
        float shv[48];
        const bool vec = q.sh_vec != 0;
        const int ncoef = (q.deg + 1) * (q.deg + 1);
        if (vec) {
            const float4 *rp = reinterpret_cast<const float4 *>(q.shs) + row * 12;
            const int n4 = q.deg > 2 ? 12 : (q.deg > 1 ? 7 : (q.deg > 0 ? 3 : 0));
#pragma unroll
            for (int i = 0; i < 12; i++) {
                const float4 t = i < n4 ? __ldg(rp + i) : make_float4(0.f, 0.f, 0.f, 0.f);
                shv[4 * i] = t.x, shv[4 * i + 1] = t.y, shv[4 * i + 2] = t.z, shv[4 * i + 3] = t.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const bool in = k < ncoef && k > 0;
                const float *p = q.shs + (row * M + k) * 3;
                shv[3 * k] = in ? __ldg(p) : 0.f, shv[3 * k + 1] = in ? __ldg(p + 1) : 0.f, shv[3 * k + 2] = in ? __ldg(p + 2) : 0.f;
            }
        }
        // basis weights dRGB/dsh_k (zero beyond the active degree)
        float bw[16];
#pragma unroll
        for (int k = 0; k < 16; k++) bw[k] = 0.f;
        bw[0] = bSH_C0;
        float dxr = 0.f, dxg = 0.f, dxb = 0.f, dyr = 0.f, dyg = 0.f, dyb = 0.f, dzr = 0.f, dzg = 0.f, dzb = 0.f;
#define SHK(k, ch) shv[3 * (k) + (ch)]
#define ADD3(dr, dg, db, coef, k) { const float c_ = (coef); dr += c_ * SHK(k, 0), dg += c_ * SHK(k, 1), db += c_ * SHK(k, 2); }
        if (q.deg > 0) {
            bw[1] = -bSH_C1 * y, bw[2] = bSH_C1 * z, bw[3] = -bSH_C1 * x;
            ADD3(dxr, dxg, dxb, -bSH_C1, 3) ADD3(dyr, dyg, dyb, -bSH_C1, 1) ADD3(dzr, dzg, dzb, bSH_C1, 2)
            if (q.deg > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                bw[4] = bSH_C2[0] * xy, bw[5] = bSH_C2[1] * yz, bw[6] = bSH_C2[2] * (2.f * zz - xx - yy);
                bw[7] = bSH_C2[3] * xz, bw[8] = bSH_C2[4] * (xx - yy);
                ADD3(dxr, dxg, dxb, bSH_C2[0] * y, 4) ADD3(dxr, dxg, dxb, bSH_C2[2] * 2.f * -x, 6)
                ADD3(dxr, dxg, dxb, bSH_C2[3] * z, 7) ADD3(dxr, dxg, dxb, bSH_C2[4] * 2.f * x, 8)
                ADD3(dyr, dyg, dyb, bSH_C2[0] * x, 4) ADD3(dyr, dyg, dyb, bSH_C2[1] * z, 5)
                ADD3(dyr, dyg, dyb, bSH_C2[2] * 2.f * -y, 6) ADD3(dyr, dyg, dyb, bSH_C2[4] * 2.f * -y, 8)
                ADD3(dzr, dzg, dzb, bSH_C2[1] * y, 5) ADD3(dzr, dzg, dzb, bSH_C2[2] * 2.f * 2.f * z, 6)
                ADD3(dzr, dzg, dzb, bSH_C2[3] * x, 7)
                if (q.deg > 2) {
                    bw[9] = bSH_C3[0] * y * (3.f * xx - yy), bw[10] = bSH_C3[1] * xy * z;
                    bw[11] = bSH_C3[2] * y * (4.f * zz - xx - yy), bw[12] = bSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
                    bw[13] = bSH_C3[4] * x * (4.f * zz - xx - yy), bw[14] = bSH_C3[5] * z * (xx - yy);
                    bw[15] = bSH_C3[6] * x * (xx - 3.f * yy);
                    ADD3(dxr, dxg, dxb, bSH_C3[0] * 3.f * 2.f * xy, 9) ADD3(dxr, dxg, dxb, bSH_C3[1] * yz, 10)
                    ADD3(dxr, dxg, dxb, bSH_C3[2] * -2.f * xy, 11) ADD3(dxr, dxg, dxb, bSH_C3[3] * -3.f * 2.f * xz, 12)
                    ADD3(dxr, dxg, dxb, bSH_C3[4] * (-3.f * xx + 4.f * zz - yy), 13) ADD3(dxr, dxg, dxb, bSH_C3[5] * 2.f * xz, 14)
                    ADD3(dxr, dxg, dxb, bSH_C3[6] * 3.f * (xx - yy), 15)
                    ADD3(dyr, dyg, dyb, bSH_C3[0] * 3.f * (xx - yy), 9) ADD3(dyr, dyg, dyb, bSH_C3[1] * xz, 10)
                    ADD3(dyr, dyg, dyb, bSH_C3[2] * (-3.f * yy + 4.f * zz - xx), 11) ADD3(dyr, dyg, dyb, bSH_C3[3] * -3.f * 2.f * yz, 12)
                    ADD3(dyr, dyg, dyb, bSH_C3[4] * -2.f * xy, 13) ADD3(dyr, dyg, dyb, bSH_C3[5] * -2.f * yz, 14)
                    ADD3(dyr, dyg, dyb, bSH_C3[6] * -3.f * 2.f * xy, 15)
                    ADD3(dzr, dzg, dzb, bSH_C3[1] * xy, 10) ADD3(dzr, dzg, dzb, bSH_C3[2] * 4.f * 2.f * yz, 11)
                    ADD3(dzr, dzg, dzb, bSH_C3[3] * 3.f * (2.f * zz - xx - yy), 12) ADD3(dzr, dzg, dzb, bSH_C3[4] * 4.f * 2.f * xz, 13)
                    ADD3(dzr, dzg, dzb, bSH_C3[5] * (xx - yy), 14)
                }
            }
        }
#undef ADD3
#undef SHK
        // table row gradient: dL_dsh[row][k][ch] += bw[k] * dL_dRGB[ch]
        float *dst = q.dL_dsh + row * M * 3;
        if (vec) {
            float v[48];
#pragma unroll
            for (int k = 0; k < 16; k++) v[3 * k] = bw[k] * gr, v[3 * k + 1] = bw[k] * gg, v[3 * k + 2] = bw[k] * gb;
            const int n4 = q.deg > 2 ? 12 : (q.deg > 1 ? 7 : (q.deg > 0 ? 3 : 1));     // chunks that hold an active coefficient
#pragma unroll
            for (int i = 0; i < 12; i++)
                if (i < n4) red_v4(dst + 4 * i, v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        } else {
#pragma unroll
            for (int k = 0; k < 16; k++)
                if (k < ncoef) {
                    atomicAdd(dst + 3 * k, bw[k] * gr);
                    atomicAdd(dst + 3 * k + 1, bw[k] * gg);
                    atomicAdd(dst + 3 * k + 2, bw[k] * gb);
                }
        }
        // view direction -> mean (dnormvdv, auxiliary.h:100-110)
        const float ddx = dxr * gr + dxg * gg + dxb * gb, ddy = dyr * gr + dyg * gg + dyb * gb, ddz = dzr * gr + dzg * gg + dzb * gb;
        const float sum2 = ox * ox + oy * oy + oz * oz;
        const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
        dmx += ((+sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * invsum32;
        dmy += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * invsum32;
        dmz += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * invsum32;
    }
    gm[0] = dmx, gm[1] = dmy, gm[2] = dmz;

    // ---------------- 3D covariance -> shape table row and size factor (backward_indexed.cu:210-283) ----------------
    if (q.scales != nullptr) {
        const int64_t g = q.g_idx[idx];
        const float fac = q.factors[idx];
        const float sc0 = __ldg(q.scales + 3 * g), sc1 = __ldg(q.scales + 3 * g + 1), sc2 = __ldg(q.scales + 3 * g + 2);
        const float4 rq = __ldg(reinterpret_cast<const float4 *>(q.rots) + g);
        const float r = rq.x, x = rq.y, y = rq.z, z = rq.w;
        const float fm = fac * q.scale_modifier;
        const float s0 = fm * sc0, s1 = fm * sc1, s2 = fm * sc2;
        // Rs[a][b]: the standard rotation matrix of the quaternion; the reference's glm R holds its transpose
        const float Rs00 = 1.f - 2.f * (y * y + z * z), Rs01 = 2.f * (x * y - r * z), Rs02 = 2.f * (x * z + r * y);
        const float Rs10 = 2.f * (x * y + r * z), Rs11 = 1.f - 2.f * (x * x + z * z), Rs12 = 2.f * (y * z - r * x);
        const float Rs20 = 2.f * (x * z - r * y), Rs21 = 2.f * (y * z + r * x), Rs22 = 1.f - 2.f * (x * x + y * y);
        // dSigma (symmetric, off-diagonal halves)
        const float S00 = dc3[0], S01 = 0.5f * dc3[1], S02 = 0.5f * dc3[2], S11 = dc3[3], S12 = 0.5f * dc3[4], S22 = dc3[5];
        // dM[i][c] = sum_k (2 M[i][k]) dSigma[k][c], M[i][k] = s_i Rs[k][i]
        const float m00 = 2.f * (s0 * Rs00), m01 = 2.f * (s0 * Rs10), m02 = 2.f * (s0 * Rs20);
        const float m10 = 2.f * (s1 * Rs01), m11 = 2.f * (s1 * Rs11), m12 = 2.f * (s1 * Rs21);
        const float m20 = 2.f * (s2 * Rs02), m21 = 2.f * (s2 * Rs12), m22 = 2.f * (s2 * Rs22);
        float N00 = m00 * S00 + m01 * S01 + m02 * S02, N01 = m00 * S01 + m01 * S11 + m02 * S12, N02 = m00 * S02 + m01 * S12 + m02 * S22;
        float N10 = m10 * S00 + m11 * S01 + m12 * S02, N11 = m10 * S01 + m11 * S11 + m12 * S12, N12 = m10 * S02 + m11 * S12 + m12 * S22;
        float N20 = m20 * S00 + m21 * S01 + m22 * S02, N21 = m20 * S01 + m21 * S11 + m22 * S12, N22 = m20 * S02 + m21 * S12 + m22 * S22;
        const float ds0 = Rs00 * N00 + Rs10 * N01 + Rs20 * N02;
        const float ds1 = Rs01 * N10 + Rs11 * N11 + Rs21 * N12;
        const float ds2 = Rs02 * N20 + Rs12 * N21 + Rs22 * N22;
        float *gs = q.dL_dscale + 3 * g;
        atomicAdd(gs, ds0 * fac);                      // the reference scales by the size factor only (:253-255)
        atomicAdd(gs + 1, ds1 * fac);
        atomicAdd(gs + 2, ds2 * fac);
        q.dL_dfactor[idx] = ds0 * sc0 + ds1 * sc1 + ds2 * sc2;
        N00 *= s0, N01 *= s0, N02 *= s0;
        N10 *= s1, N11 *= s1, N12 *= s1;
        N20 *= s2, N21 *= s2, N22 *= s2;
        const float qr = 2.f * z * (N01 - N10) + 2.f * y * (N20 - N02) + 2.f * x * (N12 - N21);
        const float qx = 2.f * y * (N10 + N01) + 2.f * z * (N20 + N02) + 2.f * r * (N12 - N21) - 4.f * x * (N22 + N11);
        const float qy = 2.f * x * (N10 + N01) + 2.f * r * (N20 - N02) + 2.f * z * (N12 + N21) - 4.f * y * (N22 + N00);
        const float qz = 2.f * r * (N01 - N10) + 2.f * x * (N20 + N02) + 2.f * y * (N12 + N21) - 4.f * z * (N11 + N00);
        red_v4(q.dL_drot + 4 * g, qr, qx, qy, qz);
    } else {
        q.dL_dfactor[idx] = 0.f;
    }
}

}  // namespace

int launch_preprocess_indexed_bwd(int64_t P, int deg, int M, const float *means, const int *radii, const float *shs,
                                  const uint8_t *clamped, const float *scales, const float *factors, const float *rots,
                                  float scale_modifier, const float *cov3d, const float *view, const float *proj,
                                  float focal_x, float focal_y, float tan_fovx, float tan_fovy, const float *campos,
                                  const float *dL_dmean2d, const float *dL_dconic, const float *dL_dcolor,
                                  float *dL_dmeans, float *dL_dcov3d, float *dL_dsh, float *dL_dscale, float *dL_dfactor,
                                  float *dL_drot, const int64_t *sh_idx, const int64_t *g_idx, cudaStream_t s)
{
    if (P == 0) return C3VQ_OK;
    PbParams q;
    q.P = P, q.deg = deg, q.M = M;
    q.sh_vec = (M == 16 && ((uintptr_t)shs & 15) == 0 && ((uintptr_t)dL_dsh & 15) == 0) ? 1 : 0;
    q.means = means, q.shs = shs, q.scales = scales, q.factors = factors, q.rots = rots, q.cov3d = cov3d;
    q.view = view, q.proj = proj, q.campos = campos, q.radii = radii, q.clamped = clamped;
    q.sh_idx = sh_idx, q.g_idx = g_idx, q.dL_dmean2d = dL_dmean2d, q.dL_dconic = dL_dconic, q.dL_dcolor = dL_dcolor;
    q.scale_modifier = scale_modifier, q.focal_x = focal_x, q.focal_y = focal_y, q.tan_fovx = tan_fovx, q.tan_fovy = tan_fovy;
    q.dL_dmeans = dL_dmeans, q.dL_dcov3d = dL_dcov3d, q.dL_dsh = dL_dsh, q.dL_dscale = dL_dscale;
    q.dL_dfactor = dL_dfactor, q.dL_drot = dL_drot;
    const int64_t nb = (P + PB_THREADS - 1) / PB_THREADS;
    C3VQ_CHECK_ARG(nb < 0x7fffffffLL, "P=%lld too large", (long long)P);
    k_preprocess_indexed_bwd<<<(unsigned)nb, PB_THREADS, 0, s>>>(q);
    C3VQ_LAUNCH_CHECK("k_preprocess_indexed_bwd");
    return C3VQ_OK;
}

}  // namespace c3vq
