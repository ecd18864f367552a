// Harness around the reference's own BACKWARD::preprocess_indexed (TEST INFRASTRUCTURE ONLY).
// Compiled by `make -C oracle ref_preprocess_bwd` together with the reference's backward_indexed.cu, unmodified, from
// where it lies under /root/reference; this file only gives it a C entry point with plain device pointers so that
// tests/test_gpu_preprocess.py can run the reference kernels on the GPU box next to the product and the oracle.
#include <cstdint>
#include <cuda_runtime.h>
#include "backward.h"    // the reference's headers (cuda_rasterizer/), found through -I

extern "C" __attribute__((visibility("default")))
int ref_preprocess_indexed_backward(int P, int D, int M, const float *means3D, const int *radii, const float *shs,
                                    const bool *clamped, const float *scales, const float *scale_factors,
                                    const float *rotations, float scale_modifier, const float *cov3Ds,
                                    const float *viewmatrix, const float *projmatrix, float focal_x, float focal_y,
                                    float tan_fovx, float tan_fovy, const float *campos, const float *dL_dmean2D,
                                    const float *dL_dconic, float *dL_dmean3D, float *dL_dcolor, float *dL_dcov3D,
                                    float *dL_dsh, float *dL_dscale, float *dL_dscale_factor, float *dL_drot,
                                    const int64_t *sh_indices, const int64_t *g_indices)
{
    BACKWARD::preprocess_indexed(P, D, M, (const float3 *)means3D, radii, shs, clamped, (const glm::vec3 *)scales,
                                 scale_factors, (const glm::vec4 *)rotations, scale_modifier, cov3Ds, viewmatrix,
                                 projmatrix, focal_x, focal_y, tan_fovx, tan_fovy, (const glm::vec3 *)campos,
                                 (const float3 *)dL_dmean2D, dL_dconic, (glm::vec3 *)dL_dmean3D, dL_dcolor, dL_dcov3D,
                                 dL_dsh, (glm::vec3 *)dL_dscale, dL_dscale_factor, (glm::vec4 *)dL_drot, sh_indices,
                                 g_indices);
    return (int)cudaDeviceSynchronize();
}
