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

/* ---- RunPriorAwareFusion, acmmp_definitions.cpp:443-826 ------------------------------------------
 * Two (depth, normal) maps per view: set 0 from the fusion folder, set 1 from the run's own output
 * folder (:607-643).  A pixel tries both of its hypotheses against both maps of every source view;
 * the hypothesis with more consistent sources wins, a hypothesis that passes alone needs
 * single_match_penalty more consistent sources (:746-776).  Only the winner's consistent source
 * pixels are masked (:816-820) - there is no stale used_list here.
 * Pinned against the reference's own function through oracle/ref_host_harness.cpp
 * (tests/test_fusion_pinned.py). */
typedef struct {
    int below;     /* any of the two source maps passes the three thresholds (:487-516) */
    float dyn;
    size_t sidx;
} paf_cand;

static void paf_metric(const orc_camera *ref, const orc_camera *src, int r, int c, float ref_depth, const float *ref_normal,
                       int src_r, int src_c, float src_depth, const float *src_normal, int *ok, float *dc) {
    float reproj = 1e6f, rel = 1e6f, ang = 1e6f;                      /* cmetric defaults (:450-452) */
    if (src_depth > 0) {
        float T[3], tx, ty, pd;
        world_point(src, src_c, src_r, src_depth, T);
        project(ref, T, &tx, &ty, &pd);
        reproj = (float)sqrt(pow(c - tx, 2) + pow(r - ty, 2));
        rel = fabsf(pd - ref_depth) / ref_depth;
        ang = get_angle(ref_normal, src_normal);
    }
    *ok = reproj < 2.0f && rel < 0.01f && ang < 0.174533f;
    *dc = (float)exp(-(reproj + 200 * rel + ang * 10));
}

static int paf_candidates(int n_unused, const orc_camera *cams, const float *const *depths, const float *const *normals,
                          const float *const *depths2, const float *const *normals2, uint8_t *const *masks, const int *ngb,
                          int num_ngb, int i, int r, int c, float ref_depth, const float *ref_normal, paf_cand *out) {
    (void)n_unused;
    int m = 0;
    float X[3];
    for (int j = 0; j < num_ngb; ++j) {                               /* getCandidates (:519-573) */
        const int s = ngb[j];
        float qx, qy, pd;
        world_point(&cams[i], c, r, ref_depth, X);
        project(&cams[s], X, &qx, &qy, &pd);
        const int src_r = (int)(qy + 0.5f), src_c = (int)(qx + 0.5f);
        if (!(src_c >= 0 && src_c < cams[s].width && src_r >= 0 && src_r < cams[s].height)) continue;
        const size_t sidx = (size_t)src_r * cams[s].width + src_c;
        if (masks[s][sidx] == 1) continue;
        int ok0, ok1;
        float dc0, dc1;
        paf_metric(&cams[i], &cams[s], r, c, ref_depth, ref_normal, src_r, src_c, depths[s][sidx], normals[s] + 3 * sidx, &ok0, &dc0);
        paf_metric(&cams[i], &cams[s], r, c, ref_depth, ref_normal, src_r, src_c, depths2[s][sidx], normals2[s] + 3 * sidx, &ok1, &dc1);
        paf_cand cd;
        cd.below = ok0 || ok1;
        cd.dyn = ok0 && ok1 ? (float)fmax(dc0, dc1) : (ok0 ? dc0 : (ok1 ? dc1 : 0.0f));
        cd.sidx = sidx;
        cd.below = cd.below ? 1 + s : 0;                              /* keeps the view index for the marking */
        out[m++] = cd;
    }
    return m;
}

long orc_prior_fusion(int n, const orc_camera *cams, const float *const *depths, const float *const *normals,
                      const float *const *depths2, const float *const *normals2, const uint8_t *const *colors,
                      const uint8_t *const *masks0, const int *pair_offsets, const int *pair_ids, float consistency_scalar,
                      int num_consistent_thresh, int single_match_penalty, float *points, long capacity) {
    uint8_t **masks = (uint8_t **)malloc(sizeof(uint8_t *) * n);
    for (int i = 0; i < n; ++i) {
        const size_t P = (size_t)cams[i].width * cams[i].height;
        masks[i] = (uint8_t *)calloc(P, 1);
        if (masks0 && masks0[i]) memcpy(masks[i], masks0[i], P);
    }
    long count = 0;
    for (int i = 0; i < n; ++i) {
        const int cols = cams[i].width, rows = cams[i].height;
        const int num_ngb = pair_offsets[i + 1] - pair_offsets[i];
        const int *ngb = pair_ids + pair_offsets[i];
        paf_cand *c0 = (paf_cand *)malloc(sizeof(paf_cand) * (num_ngb > 0 ? num_ngb : 1));
        paf_cand *c1 = (paf_cand *)malloc(sizeof(paf_cand) * (num_ngb > 0 ? num_ngb : 1));
        for (int r = 0; r < rows; ++r)
            for (int c = 0; c < cols; ++c) {
                const size_t idx = (size_t)r * cols + c;
                if (masks[i][idx] == 1) continue;
                const float d0 = depths[i][idx], d1 = depths2[i][idx];
                const float *n0 = normals[i] + 3 * idx, *n1 = normals2[i] + 3 * idx;
                if (d0 <= 0.0 && d1 <= 0.0) continue;
                float cons0 = 0, cons1 = 0;
                int k0 = 0, k1 = 0, m0 = 0, m1 = 0, t0 = 0, t1 = 0;
                if (d0 > 0.0) {
                    m0 = paf_candidates(n, cams, depths, normals, depths2, normals2, masks, ngb, num_ngb, i, r, c, d0, n0, c0);
                    for (int k = 0; k < m0; ++k)
                        if (c0[k].below) { k0++; cons0 += c0[k].dyn; }
                    t0 = k0 >= num_consistent_thresh && cons0 > consistency_scalar * k0;
                }
                if (d1 > 0.0) {
                    m1 = paf_candidates(n, cams, depths, normals, depths2, normals2, masks, ngb, num_ngb, i, r, c, d1, n1, c1);
                    for (int k = 0; k < m1; ++k)
                        if (c1[k].below) { k1++; cons1 += c1[k].dyn; }
                    t1 = k1 >= num_consistent_thresh && cons1 > consistency_scalar * k1;
                }
                int passing, sel;                                      /* (:746-776) */
                if (t0 && t1) { passing = 1; sel = k1 >= k0; }
                else if (t1) { passing = k1 >= num_consistent_thresh + single_match_penalty; sel = 1; }
                else { passing = t0 && k0 >= num_consistent_thresh + single_match_penalty; sel = 0; }
                if (!passing) continue;
                const float gd = sel ? d1 : d0;
                const float *gn = sel ? n1 : n0;
                if (points && count < capacity) {
                    float *o = points + count * 9;
                    world_point(&cams[i], c, r, gd, o);
                    o[3] = gn[0]; o[4] = gn[1]; o[5] = gn[2];
                    if (colors && colors[i]) { o[6] = colors[i][idx * 3]; o[7] = colors[i][idx * 3 + 1]; o[8] = colors[i][idx * 3 + 2]; }
                    else { o[6] = o[7] = o[8] = 0; }
                }
                count++;
                const paf_cand *it = sel ? c1 : c0;
                const int m = sel ? m1 : m0;
                for (int k = 0; k < m; ++k)
                    if (it[k].below) masks[it[k].below - 1][it[k].sidx] = 1;
            }
        free(c0);
        free(c1);
    }
    for (int i = 0; i < n; ++i) free(masks[i]);
    free(masks);
    return count;
}

/* ---- pSampler::GetPriorPlaneEstimate, acmmp_definitions.cpp:98-177 (seeded mode, `-p`) ------------
 * depth16 / normal16: the 16-bit prior images as cv::imread(path, -1) returns them (prows x pcols; the
 * normal image's three channels in stored order: channel 0 -> x).  Depth = v * (dmax - dmin) / 65535 +
 * dmin, normal = v * 2 / 65536 - 1 (cv::Mat::convertTo, float arithmetic for 16-bit sources); sampled
 * at (i * scale, j * scale), scale = prows / rows (:133).  The plane is the normal flipped to face the
 * camera, "normalised" by normVec3 - which MULTIPLIES by the length (:34-41), kept - and the distance
 * of the back-projected pixel to the origin.  Pinned against the reference's own function
 * (oracle/ref_host_harness.cpp: ref_host_prior_planes, IEEE build) in tests/test_seeded.py. */
void orc_prior_planes_from_maps(const orc_camera *cam, const uint16_t *depth16, const uint16_t *normal16, int prows, int pcols,
                                int rows, int cols, float *planes4) {
    (void)pcols;
    const float dist = cam->depth_max - cam->depth_min;
    const float range = dist / 65535.0f;
    const float na = (float)(2.0 / 65536.0), nb = -1.0f;
    const int scale = prows / rows;
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) {
            const size_t at = (size_t)(i * scale) * pcols + (size_t)j * scale;
            const float dprod = (float)depth16[at] * range;
            const float depth = dprod + cam->depth_min;
            float n[3];
            for (int k = 0; k < 3; ++k) {
                const float prod = (float)normal16[3 * at + k] * na;
                n[k] = prod + nb;
            }
            float X[3];
            X[0] = depth * (j - cam->K[2]) / cam->K[0];
            X[1] = depth * (i - cam->K[5]) / cam->K[4];
            X[2] = depth;
            const float norm = (float)sqrt(X[0] * X[0] + X[1] * X[1] + X[2] * X[2]);
            const float v0 = X[0] / norm, v1 = X[1] / norm, v2 = X[2] / norm;
            const float dot = n[0] * v0 + n[1] * v1 + n[2] * v2;
            if (dot > 0.0f) { n[0] = -n[0]; n[1] = -n[1]; n[2] = -n[2]; }
            const float len = sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
            n[0] *= len; n[1] *= len; n[2] *= len;
            float *o = planes4 + 4 * ((size_t)i * cols + j);
            o[0] = n[0]; o[1] = n[1]; o[2] = n[2];
            o[3] = -(n[0] * X[0] + n[1] * X[1] + n[2] * X[2]);
        }
}
