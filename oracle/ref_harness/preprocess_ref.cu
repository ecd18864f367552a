// Harness around the reference's own FORWARD::preprocess_indexed (TEST INFRASTRUCTURE ONLY).
// Compiled by `make -C oracle ref_preprocess` together with the reference's forward_indexed.cu, unmodified, from
// where it lies under /root/reference; this file only gives it a C entry point with plain device pointers so that
// tests/test_gpu_preprocess.py can run the reference kernel on the GPU box next to the product and the oracle.
#include <cstdint>
#include <cuda_runtime.h>
#include "forward.h"     // the reference's headers (cuda_rasterizer/), found through -I
#include "config.h"

extern "C" __attribute__((visibility("default")))
int ref_preprocess_indexed(int P, int D, int M, const float *means3D, const float *scales, const float *scale_factors,
                           float scale_modifier, const float *rotations, const float *opacities, const float *shs,
                           bool *clamped, const float *viewmatrix, const float *projmatrix, const float *cam_pos, int W,
                           int H, float focal_x, float focal_y, float tan_fovx, float tan_fovy, int *radii,
                           float *points_xy, float *depths, float *cov3Ds, float *rgb, float *conic_opacity,
                           uint32_t *tiles_touched, const int64_t *sh_indices, const int64_t *g_indices, int clamp_color)
{
    const dim3 grid((W + BLOCK_X - 1) / BLOCK_X, (H + BLOCK_Y - 1) / BLOCK_Y, 1);
    FORWARD::preprocess_indexed(P, D, M, means3D, (const glm::vec3 *)scales, scale_factors, scale_modifier,
                                (const glm::vec4 *)rotations, opacities, shs, clamped, nullptr, nullptr, viewmatrix,
                                projmatrix, (const glm::vec3 *)cam_pos, W, H, focal_x, focal_y, tan_fovx, tan_fovy,
                                radii, (float2 *)points_xy, depths, cov3Ds, rgb, (float4 *)conic_opacity, grid,
                                tiles_touched, false, sh_indices, g_indices, clamp_color != 0);
    return (int)cudaDeviceSynchronize();
}
