// prior.h - planar-prior construction from the current depth/cost maps (host).
// Reference: ACMMP::GetSupportPoints / DelaunayTriangulation / GetPriorPlaneParams
// / GetDepthFromPlaneParam (ACMMP.cpp:868-958) and the raster + range check in
// ProcessProblem (acmmp_definitions.cpp:332-374).
#pragma once
#include <stdint.h>

#include "../../include/acmmp_b200.h"

namespace acmmp {

// prior4: W*H*4 floats (plane per pixel, only where mask > 0); mask: W*H labels
// (triangle index + 1, 0 = none).
void build_planar_prior_host(const acmmp_camera &cam, float depth_min, float depth_max, int W, int H,
                             const float *costs, const float *depth, float *prior4, uint32_t *mask,
                             int *num_support, int *num_triangles);

}  // namespace acmmp
