// prior.h - planar-prior construction from the current depth/cost maps.
// Reference: ACMMP::GetSupportPoints / DelaunayTriangulation / GetPriorPlaneParams
// / GetDepthFromPlaneParam (ACMMP.cpp:868-958) and the raster + range check in
// ProcessProblem (acmmp_definitions.cpp:332-374).
// Split: support-point selection, triangle rasterisation, plane fit and the
// per-pixel range check are CUDA kernels (prior_kernels.cu.inc); only the 2-D
// Delaunay triangulation of the (<= P/25) support points runs on the host.
#pragma once
#include <stdint.h>

#include <vector>

#include "../../include/acmmp_b200.h"

namespace acmmp {

// Delaunay triangulation of n integer points (x0,y0,x1,y1,...).  Appends
// 6 ints per triangle (x,y of the three vertices) to `tris`; returns the count.
int delaunay_triangulate(const int32_t *xy, int n, std::vector<int32_t> &tris);
// threads one triangulation may use for the top of its recursion (1 = sequential, 0 = a third of the
// host's cores, capped at 8); the result does not depend on it
void delaunay_set_threads(int threads);

}  // namespace acmmp
