/*
 * cost.c - CPU restatement of the per-(pixel, plane, view) arithmetic of the
 * reference's PatchMatch path: texture sampling, plane-induced homography,
 * bilateral NCC, geometric-consistency cost, initial multi-view cost + view mask.
 *
 * TEST INFRASTRUCTURE (see oracle.h).  Plain IEEE float, reference expression
 * order, no FMA contraction; the reference build uses --use_fast_math, so the
 * agreement with the GPU is ~1e-5 on costs, not bit-exact.
 */
#include <math.h>
#include <string.h>

#include "oracle.h"

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* Hardware bilinear filter of a float texture, unnormalised coordinates,
 * clamp addressing (cudaAddressModeWrap is not honoured for unnormalised
 * coordinates; texture set-up at ACMMP.cpp:646-663).  CUDA programming guide,
 * "Linear filtering": xB = x - 0.5, i = floor(xB), alpha = frac(xB) held in
 * 9-bit fixed point with 8 fractional bits. */
float orc_tex_bilinear(const float *img, int w, int h, float x, float y) {
    const float xb = x - 0.5f, yb = y - 0.5f;
    float fx = floorf(xb), fy = floorf(yb);
    /* 8 fractional bits, round to nearest; a fraction that rounds to 1 carries into the index */
    int ia = (int)floorf((xb - fx) * 256.0f + 0.5f);
    int ib = (int)floorf((yb - fy) * 256.0f + 0.5f);
    if (ia >= 256) { ia = 0; fx += 1.0f; }
    if (ib >= 256) { ib = 0; fy += 1.0f; }
    /* out-of-range coordinates saturate before the int conversion */
    const float fxc = fx < -2.0f ? -2.0f : (fx > (float)w ? (float)w : fx);
    const float fyc = fy < -2.0f ? -2.0f : (fy > (float)h ? (float)h : fy);
    const int i0 = clampi((int)fxc, 0, w - 1), i1 = clampi((int)fxc + 1, 0, w - 1);
    const int j0 = clampi((int)fyc, 0, h - 1), j1 = clampi((int)fyc + 1, 0, h - 1);
    const float t00 = img[(size_t)j0 * w + i0], t10 = img[(size_t)j0 * w + i1];
    const float t01 = img[(size_t)j1 * w + i0], t11 = img[(size_t)j1 * w + i1];
    /* The four 2-D weights are themselves multiples of 1/256 that always sum to 1 (measured on
       B200 with tools/tex_weight_probe.cu over all 65536 (a, b)):
         w11 = (a b + 128) >> 8,  w10 = a - w11,  w01 = b - w11,  w00 = 256 - a - b + w11.
       With these weights the blend is exact for integer texels (8-bit grey images) and equals the
       hardware's result for 99.66 % of samples on arbitrary fp32 texels, the rest being 1 ulp off
       (tools/tex_emul_probe.cu): the filter keeps more than 24 bits internally. */
    const int w11 = (ia * ib + 128) >> 8, w10 = ia - w11, w01 = ib - w11, w00 = 256 - ia - ib + w11;
    return (float)(((double)w00 * t00 + (double)w10 * t10 + (double)w01 * t01 + (double)w11 * t11) * (1.0 / 256.0));
}

/* texel-centre fetch (x + 0.5, y + 0.5): returns the texel itself, clamped */
static inline float texel(const float *img, int w, int h, int x, int y) {
    return img[(size_t)clampi(y, 0, h - 1) * w + clampi(x, 0, w - 1)];
}

/* ComputeHomography, ACMMP.cu:262-322 */
void orc_homography(const orc_camera *r, const orc_camera *s, const float pl[4], float H[9]) {
    float rC[3], sC[3];
    for (int k = 0; k < 3; ++k) {
        rC[k] = -(r->R[k] * r->t[0] + r->R[3 + k] * r->t[1] + r->R[6 + k] * r->t[2]);
        sC[k] = -(s->R[k] * s->t[0] + s->R[3 + k] * s->t[1] + s->R[6 + k] * s->t[2]);
    }
    float Rr[9], Cr[3], tr[3];
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b)
            Rr[3 * a + b] = s->R[3 * a] * r->R[3 * b] + s->R[3 * a + 1] * r->R[3 * b + 1] + s->R[3 * a + 2] * r->R[3 * b + 2];
    for (int k = 0; k < 3; ++k) Cr[k] = rC[k] - sC[k];
    for (int a = 0; a < 3; ++a) tr[a] = s->R[3 * a] * Cr[0] + s->R[3 * a + 1] * Cr[1] + s->R[3 * a + 2] * Cr[2];
    float G[9];
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) G[3 * a + b] = Rr[3 * a + b] - tr[a] * pl[b] / pl[3];
    float T[9];
    for (int a = 0; a < 3; ++a) {
        T[3 * a] = G[3 * a] / r->K[0];
        T[3 * a + 1] = G[3 * a + 1] / r->K[4];
        T[3 * a + 2] = -G[3 * a] * r->K[2] / r->K[0] - G[3 * a + 1] * r->K[5] / r->K[4] + G[3 * a + 2];
    }
    H[0] = s->K[0] * T[0] + s->K[2] * T[6];
    H[1] = s->K[0] * T[1] + s->K[2] * T[7];
    H[2] = s->K[0] * T[2] + s->K[2] * T[8];
    H[3] = s->K[4] * T[3] + s->K[5] * T[6];
    H[4] = s->K[4] * T[4] + s->K[5] * T[7];
    H[5] = s->K[4] * T[5] + s->K[5] * T[8];
    H[6] = s->K[8] * T[6];
    H[7] = s->K[8] * T[7];
    H[8] = s->K[8] * T[8];
}

/* ComputeCorrespondingPoint, ACMMP.cu:324-331 */
static inline void warp(const float H[9], int x, int y, float *ox, float *oy) {
    const float px = H[0] * x + H[1] * y + H[2];
    const float py = H[3] * x + H[4] * y + H[5];
    const float pz = H[6] * x + H[7] * y + H[8];
    *ox = px / pz;
    *oy = py / pz;
}

/* ComputeBilateralNCC + ComputeBilateralWeight, ACMMP.cu:353-432
 * (patch 11, stride 2: offsets -5..5, outer loop over x offsets). */
float orc_ncc(const orc_problem *p, int px, int py, const float plane[4], int view) {
    const float cost_max = 2.0f;
    const orc_camera *rc = &p->cams[0], *sc = &p->cams[view];
    const float *rimg = p->images[0], *simg = p->images[view];
    float H[9];
    orc_homography(rc, sc, plane, H);
    float cx, cy;
    warp(H, px, py, &cx, &cy);
    if (cx >= sc->width || cx < 0.0f || cy >= sc->height || cy < 0.0f) return cost_max;
    const float ss = p->sigma_spatial, scol = p->sigma_color;
    float sum_r = 0, sum_rr = 0, sum_s = 0, sum_ss = 0, sum_rs = 0, sum_w = 0;
    const float rcen = texel(rimg, rc->width, rc->height, px, py);
    for (int i = -5; i <= 5; i += 2) {
        float r_r = 0, r_rr = 0, r_s = 0, r_ss = 0, r_rs = 0, r_w = 0;
        for (int j = -5; j <= 5; j += 2) {
            const int qx = px + i, qy = py + j;
            const float rp = texel(rimg, rc->width, rc->height, qx, qy);
            float sx, sy;
            warp(H, qx, qy, &sx, &sy);
            const float sp = orc_tex_bilinear(simg, sc->width, sc->height, sx + 0.5f, sy + 0.5f);
            const float sd = sqrtf((float)i * (float)i + (float)j * (float)j);
            const float cd = fabsf(rp - rcen);
            const float w = expf(-sd / (2.0f * ss * ss) - cd / (2.0f * scol * scol));
            r_r += w * rp;
            r_rr += w * rp * rp;
            r_s += w * sp;
            r_ss += w * sp * sp;
            r_rs += w * rp * sp;
            r_w += w;
        }
        sum_r += r_r; sum_rr += r_rr; sum_s += r_s; sum_ss += r_ss; sum_rs += r_rs; sum_w += r_w;
    }
    const float inv = 1.0f / sum_w;
    sum_r *= inv; sum_rr *= inv; sum_s *= inv; sum_ss *= inv; sum_rs *= inv;
    const float var_r = sum_rr - sum_r * sum_r;
    const float var_s = sum_ss - sum_s * sum_s;
    if (var_r < 1e-5f || var_s < 1e-5f) return cost_max;
    const float cov = sum_rs - sum_r * sum_s;
    const float c = 1.0f - cov / sqrtf(var_r * var_s);
    return fmaxf(0.0f, fminf(cost_max, c));
}

/* ComputeDepthfromPlaneHypothesis, ACMMP.cu:163-168 */
float orc_plane_depth(const orc_camera *c, const float pl[4], int px, int py) {
    return -pl[3] * c->K[0] / ((px - c->K[2]) * pl[0] + (c->K[0] / c->K[4]) * (py - c->K[5]) * pl[1] + c->K[0] * pl[2]);
}

/* GetDistance2Origin, ACMMP.cu:144-149 */
float orc_plane_offset(const orc_camera *c, int px, int py, float depth, const float n[3]) {
    const float X0 = depth * (px - c->K[2]) / c->K[0];
    const float X1 = depth * (py - c->K[5]) / c->K[4];
    return -(n[0] * X0 + n[1] * X1 + n[2] * depth);
}

/* Get3DPointonWorld_cu / ProjectonCamera_cu, ACMMP.cu:480-516 */
static void to_world(const orc_camera *c, float x, float y, float depth, float W[3]) {
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

/* ComputeGeomConsistencyCost, ACMMP.cu:518-543 */
float orc_geom_cost(const orc_problem *p, int px, int py, const float plane[4], int view) {
    const float max_cost = 3.0f;
    const orc_camera *rc = &p->cams[0], *sc = &p->cams[view];
    const float depth = orc_plane_depth(rc, plane, px, py);
    float Wp[3], sx, sy, sd;
    to_world(rc, (float)px, (float)py, depth, Wp);
    project(sc, Wp, &sx, &sy, &sd);
    const float src_depth = texel(p->depths[view], sc->width, sc->height, (int)sx, (int)sy);
    if (src_depth == 0.0f) return max_cost;
    float Ws[3], bx, by, bd;
    to_world(sc, sx, sy, src_depth, Ws);
    project(rc, Ws, &bx, &by, &bd);
    const float dc = px - bx, dr = py - by;
    return fminf(max_cost, sqrtf(dc * dc + dr * dr));
}

/* ComputeMultiViewInitialCostandSelectedViews, ACMMP.cu:434-471 */
float orc_initial_cost(const orc_problem *p, int px, int py, const float plane[4], uint32_t *sel, float *view_costs) {
    const float cost_max = 2.0f;
    float cv[32], sorted[32];
    const int nsrc = p->n - 1;
    int valid = 0;
    for (int i = 0; i < nsrc; ++i) {
        cv[i] = orc_ncc(p, px, py, plane, i + 1);
        sorted[i] = cv[i];
        if (cv[i] < cost_max) valid++;
        if (view_costs) view_costs[i] = cv[i];
    }
    for (int i = 1; i < nsrc; ++i) { /* sort_small, ACMMP.cu:24-34 */
        const float t = sorted[i];
        int j = i;
        for (; j >= 1 && t < sorted[j - 1]; --j) sorted[j] = sorted[j - 1];
        sorted[j] = t;
    }
    *sel = 0;
    const int top_k = valid < p->top_k ? valid : p->top_k;
    if (top_k <= 0) return cost_max;
    float cost = 0.0f;
    for (int i = 0; i < top_k; ++i) cost += sorted[i];
    const float thr = sorted[top_k - 1];
    for (int i = 0; i < nsrc; ++i)
        if (cv[i] <= thr) *sel |= 1u << i;
    return cost / top_k;
}
