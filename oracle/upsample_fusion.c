/*
 * upsample_fusion.c - CPU restatement of the steps either side of the
 * PatchMatch kernels: JBU depth upsample (JBU_cu, ACMMP.cu:1458-1516; RunJBU,
 * ACMMP.cpp:1022-1087), the bilinear pyramid resize the reference delegates to
 * OpenCV (cv::resize INTER_LINEAR on CV_32F, ACMMP.cpp:589), and RunFusion
 * (acmmp_definitions.cpp:828-1043) which is the reference's own CPU hot loop
 * and the timed CPU baseline of bench.py.
 *
 * TEST INFRASTRUCTURE (see oracle.h).
 *
 * Pinning: orc_fusion is checked against the reference's OWN RunFusion, compiled from
 * where it lies (oracle/ref_host_harness.cpp -> oracle/_ref/libacmmp_ref_host*.so) and run
 * on its own on-disk layout: bit-exact (points, order, normals, colours) against the IEEE
 * build, identical accepted set against the -O3 -ffast-math build
 * (tests/test_fusion_pinned.py).
 *
 * Third-party arithmetic not under /root/reference: OpenCV (unpinned; Docker
 * image installs Ubuntu 18.04 libopencv-dev 3.2).  resize_linear restates the
 * published INTER_LINEAR algorithm for float images (source coordinate
 * (d + 0.5) * scale - 0.5, clamped 2-tap rows then columns, float weights) and
 * is pinned against Python cv2 4.13 outputs in tests/golden/resize_golden.npz.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "oracle.h"

static float spatial_gauss(float x1, float y1, float x2, float y2, float sigma) { /* ACMMP.cu:151-155 */
    float dis = (float)(pow((double)(x1 - x2), 2) + pow((double)(y1 - y2), 2) - 0.0);
    return (float)exp(-1.0 * dis / (2 * sigma * sigma));
}
static float range_gauss(float x, float sigma) { /* ACMMP.cu:157-161 */
    float xp = (float)(x - 0.0);
    return (float)exp(-1.0 * (xp * xp) / (2 * sigma * sigma));
}

int orc_jbu_depth(const float *image, int w, int h, const float *depth, int sw, int sh, float *out) {
    int Imagescale = h / sh > w / sw ? h / sh : w / sw; /* ACMMP.cpp:1030 */
    if (Imagescale == 1) return 1;                      /* ACMMP.cpp:1032-1035 */
    const float scale = (float)(1.0 * sw / w);
    const float sigmad = 0.50f, sigmar = 25.5f;
    const int nb = (Imagescale * Imagescale + 1) / 2;
    for (int py = 0; py < h; ++py)
        for (int px = 0; px < w; ++px) {
            const float o_y = py * scale, o_x = px * scale;
            const float refPix = image[(size_t)py * w + px];
            float total = 0.0f, norm = 0.0f;
            for (int j = -nb; j <= nb; ++j) {
                int r_y = (int)(o_y + j);
                r_y = r_y > 0 ? (r_y < sh ? r_y : sh - 1) : 0;
                int r_ys = py + j;
                r_ys = r_ys > 0 ? (r_ys < h ? r_ys : h - 1) : 0;
                for (int i = -nb; i <= nb; ++i) {
                    int r_x = (int)(o_x + i);
                    r_x = r_x > 0 ? (r_x < sw ? r_x : sw - 1) : 0;
                    const float srcPix = depth[(size_t)r_y * sw + r_x];
                    int r_xs = px + i;
                    r_xs = r_xs > 0 ? (r_xs < w ? r_xs : w - 1) : 0;
                    const float nbp = image[(size_t)r_ys * w + r_xs];
                    const float g = spatial_gauss(o_x, o_y, (float)r_x, (float)r_y, sigmad) * range_gauss(fabsf(refPix - nbp), sigmar);
                    norm += g;
                    total += srcPix * g;
                }
            }
            out[(size_t)py * w + px] = total / norm;
        }
    return 0;
}

void orc_resize_linear(const float *src, int sw, int sh, float *dst, int dw, int dh) {
    const double scale_x = (double)sw / dw, scale_y = (double)sh / dh;
    int *xo = (int *)malloc(sizeof(int) * dw);
    float *xa = (float *)malloc(sizeof(float) * dw);
    for (int dx = 0; dx < dw; ++dx) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)floorf(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xo[dx] = sx;
        xa[dx] = fx;
    }
    for (int dy = 0; dy < dh; ++dy) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)floorf(fy);
        fy -= sy;
        if (sy < 0) { fy = 0; sy = 0; }
        if (sy >= sh - 1) { fy = 0; sy = sh - 1; }
        const int sy1 = sy + 1 < sh ? sy + 1 : sh - 1;
        for (int dx = 0; dx < dw; ++dx) {
            const int sx = xo[dx], sx1 = sx + 1 < sw ? sx + 1 : sw - 1;
            const float a1 = xa[dx], a0 = 1.f - a1, b1 = fy, b0 = 1.f - fy;
            const float r0 = src[(size_t)sy * sw + sx] * a0 + src[(size_t)sy * sw + sx1] * a1;
            const float r1 = src[(size_t)sy1 * sw + sx] * a0 + src[(size_t)sy1 * sw + sx1] * a1;
            dst[(size_t)dy * dw + dx] = r0 * b0 + r1 * b1;
        }
    }
    free(xo);
    free(xa);
}

/* Get3DPointonWorld / ProjectonCamera / GetAngle, ACMMP.cpp:204-262 */
static void world_point(const orc_camera *c, int x, int y, float depth, float W[3]) {
    const float X0 = depth * (x - c->K[2]) / c->K[0];
    const float X1 = depth * (y - c->K[5]) / c->K[4];
    const float X2 = depth;
    for (int k = 0; k < 3; ++k) {
        const float t = c->R[k] * X0 + c->R[3 + k] * X1 + c->R[6 + k] * X2;
        const float C = -(c->R[k] * c->t[0] + c->R[3 + k] * c->t[1] + c->R[6 + k] * c->t[2]);
        W[k] = t + C;
    }
}
static void project(const orc_camera *c, const float W[3], float *x, float *y, float *d) {
    float t[3];
    for (int k = 0; k < 3; ++k) t[k] = c->R[3 * k] * W[0] + c->R[3 * k + 1] * W[1] + c->R[3 * k + 2] * W[2] + c->t[k];
    *d = c->K[6] * t[0] + c->K[7] * t[1] + c->K[8] * t[2];
    *x = (c->K[0] * t[0] + c->K[1] * t[1] + c->K[2] * t[2]) / *d;
    *y = (c->K[3] * t[0] + c->K[4] * t[1] + c->K[5] * t[2]) / *d;
}
static float get_angle(const float *a, const float *b) {
    const float dot = a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
    const float ang = acosf(dot);
    return ang != ang ? 0.0f : ang;
}

/* RunFusion, acmmp_definitions.cpp:921-1040.  Views are addressed by index;
 * pair_ids holds indices (the reference maps image ids to indices with
 * image_id_2_index, :850,970).  The per-VIEW used_list (H9: declared once per
 * view at :936, so stale source coordinates of earlier pixels are re-masked by
 * later acceptances) is replicated.  colors: 3 bytes / pixel in the stored
 * channel order, copied through. */
long orc_fusion(int n, const orc_camera *cams, const float *const *depths, const float *const *normals,
                const uint8_t *const *colors, const int *pair_offsets, const int *pair_ids, float consistency_scalar,
                int con_num_thresh, float *points, long capacity) {
    uint8_t **masks = (uint8_t **)malloc(sizeof(uint8_t *) * n);
    for (int i = 0; i < n; ++i) masks[i] = (uint8_t *)calloc((size_t)cams[i].width * cams[i].height, 1);
    long count = 0;
    for (int i = 0; i < n; ++i) {
        const int cols = cams[i].width, rows = cams[i].height;
        const int num_ngb = pair_offsets[i + 1] - pair_offsets[i];
        const int *ngb = pair_ids + pair_offsets[i];
        const float depth_max = cams[i].depth_max;
        int *used = (int *)malloc(sizeof(int) * 2 * (num_ngb > 0 ? num_ngb : 1));
        for (int j = 0; j < num_ngb; ++j) used[2 * j] = used[2 * j + 1] = -1;
        for (int r = 0; r < rows; ++r)
            for (int c = 0; c < cols; ++c) {
                const size_t idx = (size_t)r * cols + c;
                if (masks[i][idx] == 1) continue;
                const float ref_depth = depths[i][idx];
                const float *ref_normal = normals[i] + idx * 3;
                if (ref_depth <= 0.0 || ref_depth >= depth_max) continue;
                float X[3];
                world_point(&cams[i], c, r, ref_depth, X);
                int num_consistent = 0;
                float dyn = 0;
                for (int j = 0; j < num_ngb; ++j) {
                    const int s = ngb[j];
                    const int scols = cams[s].width, srows = cams[s].height;
                    float qx, qy, pd;
                    project(&cams[s], X, &qx, &qy, &pd);
                    const int src_r = (int)(qy + 0.5f), src_c = (int)(qx + 0.5f);
                    if (src_c >= 0 && src_c < scols && src_r >= 0 && src_r < srows) {
                        const size_t sidx = (size_t)src_r * scols + src_c;
                        if (masks[s][sidx] == 1) continue;
                        const float src_depth = depths[s][sidx];
                        if (src_depth <= 0.0) continue;
                        float T[3], tx, ty;
                        world_point(&cams[s], src_c, src_r, src_depth, T);
                        project(&cams[i], T, &tx, &ty, &pd);
                        const float reproj = (float)sqrt(pow(c - tx, 2) + pow(r - ty, 2));
                        const float rel = fabsf(pd - ref_depth) / ref_depth;
                        const float ang = get_angle(ref_normal, normals[s] + sidx * 3);
                        if (reproj < 2.0f && rel < 0.01f && ang < 0.174533f) {
                            used[2 * j] = src_c;
                            used[2 * j + 1] = src_r;
                            const float ti = reproj + 200 * rel + ang * 10;
                            dyn += (float)exp(-ti);
                            num_consistent++;
                        }
                    }
                }
                if (num_consistent >= con_num_thresh && dyn > consistency_scalar * num_consistent) {
                    if (points && count < capacity) {
                        float *o = points + count * 9;
                        o[0] = X[0]; o[1] = X[1]; o[2] = X[2];
                        o[3] = ref_normal[0]; o[4] = ref_normal[1]; o[5] = ref_normal[2];
                        if (colors && colors[i]) { o[6] = colors[i][idx * 3]; o[7] = colors[i][idx * 3 + 1]; o[8] = colors[i][idx * 3 + 2]; }
                        else { o[6] = o[7] = o[8] = 0; }
                    }
                    count++;
                    for (int j = 0; j < num_ngb; ++j) {
                        if (used[2 * j] == -1) continue;
                        const int s = ngb[j];
                        /* coordinates may come from an earlier pixel (stale); they are valid for view s */
                        masks[s][(size_t)used[2 * j + 1] * cams[s].width + used[2 * j]] = 1;
                    }
                }
            }
        free(used);
    }
    for (int i = 0; i < n; ++i) free(masks[i]);
    free(masks);
    return count;
}
