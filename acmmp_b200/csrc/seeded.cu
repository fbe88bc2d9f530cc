// seeded.cu - initial plane hypotheses from 16-bit prior images (the reference's `-p` mode; SURVEY.md
// 8(f) rank 4), compiled WITHOUT --use_fast_math: the reference does this on the host in float.
//
// Reference: pSampler::GetPriorPlaneEstimate with depth_normal_to_plane / getViewDirection /
// distance_to_origin / normVec3, /root/reference/src/acmmp_definitions.cpp:34-177.  One thread per
// pixel of the current level: the prior images are sampled at (row * scale, col * scale) with
// scale = prior rows / level rows (:133), decoded (depth = v * (dmax - dmin) / 65535 + dmin, normal =
// v * 2 / 65536 - 1, cv::Mat::convertTo in float), the normal is flipped to face the camera and
// multiplied by its own length (normVec3 takes the square root where it means the reciprocal, :34-41 -
// kept: with unit normals it is a no-op), w = distance of the back-projected pixel to the origin.
// The reference calls this with acmmp.GetCamera(idx) where idx is the PROBLEM index into the problem's
// own camera list (acmmp_definitions.cpp:278) - the reference view's camera only for problem 0 and past
// the end of the list for idx > number of sources; here it is always the view's own camera.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/acmmp_b200.h"
#include "kernels.h"

namespace {

struct SeedCam {
    float K0, K2, K4, K5, depth_min, depth_max;
};

__global__ void prior_maps_to_planes_kernel(const SeedCam cam, const uint16_t *__restrict__ depth16, const uint16_t *__restrict__ normal16,
                                            int prows, int pcols, int rows, int cols, float4 *__restrict__ planes) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y * blockDim.y + threadIdx.y;
    if (j >= cols || i >= rows) return;
    const int scale = prows / rows;
    const size_t at = (size_t)(i * scale) * pcols + (size_t)j * scale;
    const float range = (cam.depth_max - cam.depth_min) / 65535.0f;
    const float depth = (float)depth16[at] * range + cam.depth_min;
    const float na = 1.0f / 32768.0f;
    float n0 = (float)normal16[3 * at] * na + -1.0f;
    float n1 = (float)normal16[3 * at + 1] * na + -1.0f;
    float n2 = (float)normal16[3 * at + 2] * na + -1.0f;
    const float X0 = depth * (j - cam.K2) / cam.K0;
    const float X1 = depth * (i - cam.K5) / cam.K4;
    const float X2 = depth;
    const float norm = sqrtf(X0 * X0 + X1 * X1 + X2 * X2);
    const float v0 = X0 / norm, v1 = X1 / norm, v2 = X2 / norm;
    if (n0 * v0 + n1 * v1 + n2 * v2 > 0.0f) {
        n0 = -n0; n1 = -n1; n2 = -n2;
    }
    const float len = sqrtf(n0 * n0 + n1 * n1 + n2 * n2);
    n0 *= len; n1 *= len; n2 *= len;
    planes[(size_t)i * cols + j] = make_float4(n0, n1, n2, -(n0 * X0 + n1 * X1 + n2 * X2));
}

thread_local std::string g_seed_error;

}  // namespace

namespace acmmp {

cudaError_t launch_prior_maps_to_planes(const acmmp_camera &cam, const uint16_t *depth16, const uint16_t *normal16, int prows,
                                        int pcols, int rows, int cols, float4 *planes, cudaStream_t st) {
    SeedCam c{cam.K[0], cam.K[2], cam.K[4], cam.K[5], cam.depth_min, cam.depth_max};
    const dim3 block(32, 8), grid((cols + 31) / 32, (rows + 7) / 8);
    prior_maps_to_planes_kernel<<<grid, block, 0, st>>>(c, depth16, normal16, prows, pcols, rows, cols, planes);
    return cudaGetLastError();
}

}  // namespace acmmp

extern "C" {

const char *acmmp_seeded_last_error(void) { return g_seed_error.c_str(); }

int acmmp_prior_planes_from_maps(int device, const acmmp_camera *cam, const uint16_t *depth16, const uint16_t *normal16x3,
                                 int prior_rows, int prior_cols, int rows, int cols, float *planes4) {
    auto fail = [](int code, const char *what) {
        g_seed_error = what;
        return code;
    };
    if (!cam || !depth16 || !normal16x3 || !planes4 || rows <= 0 || cols <= 0 || prior_rows < rows || prior_cols <= 0)
        return fail(ACMMP_ERR_INVALID, "acmmp_prior_planes_from_maps: bad argument");
    const int scale = prior_rows / rows;
    if ((size_t)(rows - 1) * scale >= (size_t)prior_rows || (size_t)(cols - 1) * scale >= (size_t)prior_cols)
        return fail(ACMMP_ERR_INVALID, "acmmp_prior_planes_from_maps: prior images smaller than rows x cols x scale");
    if (cudaSetDevice(device) != cudaSuccess) return fail(ACMMP_ERR_NO_DEVICE, "acmmp_prior_planes_from_maps: no such device");
    (void)cudaGetLastError();
    const size_t PP = (size_t)prior_rows * prior_cols, P = (size_t)rows * cols;
    uint16_t *d16 = nullptr, *n16 = nullptr;
    float4 *pl = nullptr;
    cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&d16), PP * sizeof(uint16_t));
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&n16), 3 * PP * sizeof(uint16_t));
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&pl), P * sizeof(float4));
    if (e == cudaSuccess) e = cudaMemcpy(d16, depth16, PP * sizeof(uint16_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(n16, normal16x3, 3 * PP * sizeof(uint16_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = acmmp::launch_prior_maps_to_planes(*cam, d16, n16, prior_rows, prior_cols, rows, cols, pl, 0);
    if (e == cudaSuccess) e = cudaMemcpy(planes4, pl, P * sizeof(float4), cudaMemcpyDeviceToHost);
    cudaFree(pl);
    cudaFree(n16);
    cudaFree(d16);
    if (e != cudaSuccess) return fail(ACMMP_ERR_CUDA, cudaGetErrorString(e));
    return 0;
}

}  // extern "C"
