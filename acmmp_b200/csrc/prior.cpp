// prior.cpp - placeholder until the Delaunay-based construction lands.
#include "prior.h"
namespace acmmp {
void build_planar_prior_host(const acmmp_camera &, float, float, int, int, const float *, const float *, float *,
                             uint32_t *, int *num_support, int *num_triangles) {
    if (num_support) *num_support = 0;
    if (num_triangles) *num_triangles = 0;
}
}  // namespace acmmp
