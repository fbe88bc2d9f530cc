/*
 * prior.c - CPU restatement of the reference's planar-prior construction around
 * the triangulation: ACMMP::GetSupportPoints (ACMMP.cpp:868-894), the triangle
 * raster + range check of ProcessProblem (acmmp_definitions.cpp:332-374),
 * ACMMP::GetPriorPlaneParams / GetDepthFromPlaneParam (ACMMP.cpp:920-958) and the
 * per-pixel table of CudaPlanarPriorInitialization (ACMMP.cpp:811-831).
 * The triangulation itself (cv::Subdiv2D) is done by the caller through the
 * Python cv2 binding - the reference's real dependency.
 *
 * TEST INFRASTRUCTURE (see oracle.h); also the reference-side CPU code that
 * bench.py --impl reference times on the host cores (single thread, as in the
 * reference).  cv::SVD::solveZ of the 3x4 system is replaced by the equivalent
 * cross product (the null vector of [X 1] is the plane through the 3 points).
 */
#include <math.h>
#include <string.h>

#include "oracle.h"

int orc_support_points(const float *costs, int W, int H, int32_t *xy) {
    int n = 0;
    for (int col = 0; col < W; col += 5)
        for (int row = 0; row < H; row += 5) {
            float min_cost = 2.0f;
            int bx = -1, by = -1;
            const int cb = W < col + 5 ? W : col + 5, rb = H < row + 5 ? H : row + 5;
            for (int c = col; c < cb; ++c)
                for (int r = row; r < rb; ++r) {
                    const float v = costs[(size_t)r * W + c];
                    if (v < 2.0f && min_cost > v) { bx = c; by = r; min_cost = v; }
                }
            if (min_cost < 0.1f) { xy[2 * n] = bx; xy[2 * n + 1] = by; ++n; }
        }
    return n;
}

/* tris: 6 ints per triangle (already filtered to lie inside the image, in the
 * order cv::Subdiv2D::getTriangleList returned them).  prior4 / mask: W*H. */
void orc_prior_from_triangles(const orc_camera *cam, const float *depth, int W, int H, const int32_t *tris, int ntri,
                              float depth_min, float depth_max, float *tri_planes /* ntri*4 */, float *prior4,
                              uint32_t *mask) {
    float *label = (float *)prior4; /* reuse as scratch is unsafe: keep a separate pass */
    (void)label;
    memset(mask, 0, (size_t)W * H * sizeof(uint32_t));
    memset(prior4, 0, (size_t)W * H * 4 * sizeof(float));
    for (int t = 0; t < ntri; ++t) {
        const int32_t *T = tris + 6 * t;
        const int x1 = T[0], y1 = T[1], x2 = T[2], y2 = T[3], x3 = T[4], y3 = T[5];
        const float L01 = (float)sqrt(pow(x1 - x2, 2) + pow(y1 - y2, 2));
        const float L02 = (float)sqrt(pow(x1 - x3, 2) + pow(y1 - y3, 2));
        const float L12 = (float)sqrt(pow(x2 - x3, 2) + pow(y2 - y3, 2));
        float me = L01 > L02 ? L01 : L02;
        if (L12 > me) me = L12;
        const float step = (float)(1.0 / me);
        for (float p = 0; p < 1.0; p += step)
            for (float q = 0; q < 1.0 - p; q += step) {
                const int x = (int)(p * x1 + q * x2 + (1.0 - p - q) * x3);
                const int y = (int)(p * y1 + q * y2 + (1.0 - p - q) * y3);
                mask[(size_t)y * W + x] = (uint32_t)t + 1u;
            }
        /* plane through the three back-projected vertices */
        double X[3][3];
        const int vx[3] = {x1, x2, x3}, vy[3] = {y1, y2, y3};
        for (int k = 0; k < 3; ++k) {
            const float d = depth[(size_t)vy[k] * W + vx[k]];
            X[k][0] = d * (vx[k] - cam->K[2]) / cam->K[0];
            X[k][1] = d * (vy[k] - cam->K[5]) / cam->K[4];
            X[k][2] = d;
        }
        const double ax = X[1][0] - X[0][0], ay = X[1][1] - X[0][1], az = X[1][2] - X[0][2];
        const double bx = X[2][0] - X[0][0], by = X[2][1] - X[0][1], bz = X[2][2] - X[0][2];
        const double nx = ay * bz - az * by, ny = az * bx - ax * bz, nz = ax * by - ay * bx;
        const double w = -(nx * X[0][0] + ny * X[0][1] + nz * X[0][2]);
        double nrm = sqrt(nx * nx + ny * ny + nz * nz);
        if (w < 0) nrm = -nrm;
        tri_planes[4 * t] = (float)(nx / nrm);
        tri_planes[4 * t + 1] = (float)(ny / nrm);
        tri_planes[4 * t + 2] = (float)(nz / nrm);
        tri_planes[4 * t + 3] = (float)(w / nrm);
    }
    for (int i = 0; i < W; ++i)
        for (int j = 0; j < H; ++j) {
            const size_t c = (size_t)j * W + i;
            if (mask[c] > 0) {
                const float *pl = tri_planes + 4 * (mask[c] - 1);
                const float d = -pl[3] * cam->K[0] /
                                ((i - cam->K[2]) * pl[0] + (cam->K[0] / cam->K[4]) * (j - cam->K[5]) * pl[1] + cam->K[0] * pl[2]);
                if (d <= depth_max && d >= depth_min) memcpy(prior4 + 4 * c, pl, 4 * sizeof(float));
                else mask[c] = 0;
            }
        }
}
