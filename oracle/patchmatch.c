/*
 * patchmatch.c - CPU restatement of the reference's PatchMatch kernels:
 * RandomInitialization (ACMMP.cu:609-705), CheckerboardPropagation +
 * PlaneHypothesisRefinement (ACMMP.cu:707-1173), GetDepthandNormal
 * (ACMMP.cu:1199-1212), CheckerboardFilter (ACMMP.cu:1214-1328).
 *
 * TEST INFRASTRUCTURE (see oracle.h).
 *
 * Two places of the reference have no defined meaning and are pinned here the
 * same way as in the product (DESIGN.md "UB pinning"):
 *  - same-colour candidate reads race with the writes of the same launch
 *    (near-arm V candidates): they are served from the state before the
 *    half-sweep (snapshot), which is what the reference sees whenever the
 *    neighbour has not finished yet;
 *  - `float4 plane_hypotheses_now` (ACMMP.cu:1149) is uninitialised unless a
 *    neighbour is adopted in a non-prior run: compat 0 = what the sm_100 build
 *    does (takes the best arm's plane whenever that arm exists - the arg-min arm,
 *    or under a planar prior that covers the pixel the arm with the largest
 *    restricted cost - and the live plane otherwise; measured against the
 *    reference's kernels with tools/prior_ub_probe.py), compat 1 = upstream
 *    semantics (refine the live plane).
 * A defined quirk kept as it is: under a prior the adoption assigns a block-local
 * `depth_now` that shadows the one handed to the refinement (ACMMP.cu:1120,1129).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "oracle.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

static void normalize3(float v[3]) { /* NormalizeVec3, ACMMP.cu:98-105 (rsqrtf) */
    const float inv = 1.0f / sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    v[0] *= inv; v[1] *= inv; v[2] *= inv;
}

static void view_direction(const orc_camera *c, int px, int py, float depth, float v[3]) { /* ACMMP.cu:130-142 */
    const float X0 = depth * (px - c->K[2]) / c->K[0];
    const float X1 = depth * (py - c->K[5]) / c->K[4];
    const float X2 = depth;
    const float n = sqrtf(X0 * X0 + X1 * X1 + X2 * X2);
    v[0] = X0 / n; v[1] = X1 / n; v[2] = X2 / n;
}

static void random_normal(const orc_camera *c, int px, int py, uint32_t *rng, float depth, float n[3]) { /* ACMMP.cu:170-196 */
    float q1 = 1.0f, q2 = 1.0f, s = 2.0f;
    while (s >= 1.0f) {
        q1 = 2.0f * orc_rng_uniform(rng) - 1.0f;
        q2 = 2.0f * orc_rng_uniform(rng) - 1.0f;
        s = q1 * q1 + q2 * q2;
    }
    const float sq = sqrtf(1.0f - s);
    n[0] = 2.0f * q1 * sq;
    n[1] = 2.0f * q2 * sq;
    n[2] = 1.0f - 2.0f * s;
    float v[3];
    view_direction(c, px, py, depth, v);
    if (n[0] * v[0] + n[1] * v[1] + n[2] * v[2] > 0.0f) { n[0] = -n[0]; n[1] = -n[1]; n[2] = -n[2]; }
    normalize3(n);
}

static void perturbed_normal(const orc_camera *c, int px, int py, const float normal[3], uint32_t *rng, float pert, float out[3]) { /* ACMMP.cu:198-233 */
    float v[3];
    view_direction(c, px, py, 1.0f, v);
    const float a1 = (orc_rng_uniform(rng) - 0.5f) * pert;
    const float a2 = (orc_rng_uniform(rng) - 0.5f) * pert;
    const float a3 = (orc_rng_uniform(rng) - 0.5f) * pert;
    const float s1 = sinf(a1), s2 = sinf(a2), s3 = sinf(a3), c1 = cosf(a1), c2 = cosf(a2), c3 = cosf(a3);
    float R[9];
    R[0] = c2 * c3;
    R[1] = c3 * s1 * s2 - c1 * s3;
    R[2] = s1 * s3 + c1 * c3 * s2;
    R[3] = c2 * s3;
    R[4] = c1 * c3 + s1 * s2 * s3;
    R[5] = c1 * s2 * s3 - c3 * s1;
    R[6] = -s2;
    R[7] = c2 * s1;
    R[8] = c1 * c2;
    float np[3];
    for (int k = 0; k < 3; ++k) np[k] = R[3 * k] * normal[0] + R[3 * k + 1] * normal[1] + R[3 * k + 2] * normal[2];
    if (np[0] * v[0] + np[1] * v[1] + np[2] * v[2] >= 0.0f) memcpy(np, normal, sizeof np);
    normalize3(np);
    memcpy(out, np, sizeof np);
}

static void to_cam(const orc_camera *c, const float n[3], float o[3]) { /* TransformNormal2RefCam, ACMMP.cu:343-351 */
    for (int k = 0; k < 3; ++k) o[k] = c->R[3 * k] * n[0] + c->R[3 * k + 1] * n[1] + c->R[3 * k + 2] * n[2];
}
static void to_world_n(const orc_camera *c, const float n[3], float o[3]) { /* TransformNormal, ACMMP.cu:333-341 */
    for (int k = 0; k < 3; ++k) o[k] = c->R[k] * n[0] + c->R[3 + k] * n[1] + c->R[6 + k] * n[2];
}

/* SpatialGauss / RangeGauss, ACMMP.cu:151-161: float arguments, double math */
static float spatial_gauss(float x1, float y1, float x2, float y2, float sigma) {
    float dis = (float)(pow((double)(x1 - x2), 2) + pow((double)(y1 - y2), 2) - 0.0);
    return (float)exp(-1.0 * dis / (2 * sigma * sigma));
}
static float range_gauss(float x, float sigma) {
    float xp = (float)(x - 0.0);
    return (float)exp(-1.0 * (xp * xp) / (2 * sigma * sigma));
}

/* upscale_normal, ACMMP.cu:548-607 (normal part; the upsampled cost is
 * overwritten right after, ACMMP.cu:680) */
static void upscale_normal(const orc_problem *p, int px, int py, float out[3]) {
    const orc_camera *rc = &p->cams[0];
    const int W = rc->width, H = rc->height;
    const float scale = (float)(1.0 * p->scaled_cols / W);
    const float sigmad = 0.50f, sigmar = 25.5f;
    const float fs = fmaxf(W / p->scaled_cols, H / p->scaled_rows);
    const int Imagescale = (int)fs;
    const int nb = (Imagescale * Imagescale + 1) / 2;
    const float o_y = py * scale, o_x = px * scale;
    const float refPix = p->images[0][(size_t)py * W + px];
    float norm = 0.0f, nx = 0, ny = 0, nz = 0;
    for (int j = -nb; j <= nb; ++j) {
        int r_y = (int)(o_y + j);
        r_y = (int)(r_y > 0 ? (r_y < p->scaled_rows ? r_y : p->scaled_rows - 1) : 0);
        const int r_ys = py + j;
        for (int i = -nb; i <= nb; ++i) {
            int r_x = (int)(o_x + i);
            r_x = (int)(r_x > 0 ? (r_x < p->scaled_cols ? r_x : p->scaled_cols - 1) : 0);
            const int s_center = (int)(r_y * p->scaled_cols + r_x);
            const float *sn = p->scaled_planes + (size_t)s_center * 4;
            const int r_xs = px + i;
            const float nbp = p->images[0][(size_t)clampi(r_ys, 0, H - 1) * W + clampi(r_xs, 0, W - 1)];
            const float g = spatial_gauss(o_x, o_y, (float)r_x, (float)r_y, sigmad) * range_gauss(fabsf(refPix - nbp), sigmar);
            norm += g;
            nx = nx + sn[0] * g;
            ny = ny + sn[1] * g;
            nz = nz + sn[2] * g;
        }
    }
    out[0] = nx / norm; out[1] = ny / norm; out[2] = nz / norm;
    normalize3(out);
}

void orc_init(orc_problem *p, int mode, int x0, int y0, int x1, int y1) {
    const orc_camera *rc = &p->cams[0];
    const int W = rc->width;
    for (int py = y0; py < y1; ++py)
        for (int px = x0; px < x1; ++px) {
            const size_t c = (size_t)py * W + px;
            uint32_t *rng = p->rng6 + c * 6;
            float *pl = p->planes + c * 4;
            if (mode == ORC_INIT_RANDOM) { /* ACMMP.cu:626-633, 235-241 */
                const float depth = orc_rng_uniform(rng) * (p->depth_max - p->depth_min) + p->depth_min;
                random_normal(rc, px, py, rng, depth, pl);
                pl[3] = orc_plane_offset(rc, px, py, depth, pl);
            } else if (mode == ORC_INIT_SEEDED) { /* ACMMP.cu:634-639 */
                memcpy(pl, p->seeded + c * 4, 4 * sizeof(float));
            } else if (mode == ORC_INIT_PRIOR) { /* ACMMP.cu:640-661 */
                if (p->masks[c] > 0 && p->costs[c] >= 0.1f) {
                    const float pert = 0.02f;
                    const float *pr = p->prior_planes + c * 4;
                    float d = pr[3];
                    const float dmin = (1 - 3 * pert) * d, dmax = (1 + 3 * pert) * d;
                    d = orc_rng_uniform(rng) * (dmax - dmin) + dmin;
                    float n[3];
                    perturbed_normal(rc, px, py, pr, rng, (float)(3 * pert * M_PI), n);
                    pl[0] = n[0]; pl[1] = n[1]; pl[2] = n[2]; pl[3] = d;
                } else {
                    const float depth = pl[3];
                    pl[3] = orc_plane_offset(rc, px, py, depth, pl);
                }
            } else if (mode == ORC_INIT_UPSAMPLE) { /* ACMMP.cu:663-689 */
                float nw[3], nc[3];
                upscale_normal(p, px, py, nw);
                uint32_t s0;
                p->pre_costs[c] = orc_initial_cost(p, px, py, pl, &s0, NULL);
                to_cam(rc, nw, nc);
                const float depth = pl[3];
                pl[0] = nc[0]; pl[1] = nc[1]; pl[2] = nc[2];
                pl[3] = orc_plane_offset(rc, px, py, depth, pl);
            } else { /* warm start, ACMMP.cu:690-703 */
                const float *src = p->hierarchy ? p->scaled_planes + c * 4 : pl;
                float nc[3];
                to_cam(rc, src, nc);
                const float depth = src[3];
                pl[0] = nc[0]; pl[1] = nc[1]; pl[2] = nc[2];
                pl[3] = orc_plane_offset(rc, px, py, depth, pl);
            }
            p->costs[c] = orc_initial_cost(p, px, py, pl, &p->sel[c], NULL);
        }
}

static float prior_factor(float depth_diff, float angle_cos, float tds2, float tas2) { /* ACMMP.cu:767,1112 */
    const float ad = acosf(angle_cos);
    return 0.5f + expf(-depth_diff * depth_diff / tds2) * expf(-ad * ad / tas2);
}

/* One pixel of CheckerboardPropagation.  P/C: snapshot planes / costs. */
static void sweep_pixel(orc_problem *p, const float *P, const float *C, int iter, int px, int py) {
    const orc_camera *rc = &p->cams[0];
    const int W = rc->width, H = rc->height, nsrc = p->n - 1;
    const int center = py * W + px;
    float cost_array[8][32];
    memset(cost_array, 0, sizeof cost_array);
    cost_array[0][0] = 2.0f; /* `= {2.0f}`, ACMMP.cu:805 */
    int flag[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int pos[8];
    for (int k = 0; k < 8; ++k) pos[k] = center;
    /* arms: 0 up_near 1 up_far 2 down_near 3 down_far 4 left_near 5 left_far 6 right_near 7 right_far */
    for (int arm = 0; arm < 8; ++arm) {
        const int dirx = arm >= 6 ? 1 : (arm >= 4 ? -1 : 0);
        const int diry = arm < 2 ? -1 : (arm < 4 ? 1 : 0);
        const int step = dirx + diry * W;
        const int room = arm < 2 ? py : arm < 4 ? (H - 1 - py) : arm < 6 ? px : (W - 1 - px);
        if (arm & 1) { /* far arms, ACMMP.cu:813-888 */
            if (room > 2) {
                flag[arm] = 1;
                int best = center + 3 * step;
                float cm = C[best];
                for (int i = 1; i < 11; ++i)
                    if (room > 2 + 2 * i) {
                        const int t = center + (3 + 2 * i) * step;
                        const int take = arm == 7 ? (cm < C[t]) : (C[t] < cm); /* right_far keeps the max, :879 */
                        if (take) { cm = C[t]; best = t; }
                    }
                pos[arm] = best;
            }
        } else { /* near arms, ACMMP.cu:890-993 */
            if (room > 0) {
                flag[arm] = 1;
                const int vertical = arm < 4;
                const int lat = vertical ? 1 : W;
                const int room_neg = vertical ? px : py;
                const int room_pos = vertical ? (W - 1 - px) : (H - 1 - py);
                int best = center + step;
                float cm = C[best];
                for (int i = 0; i < 3; ++i) {
                    if (room > 1 + i && room_neg > i) {
                        const int t = center + (2 + i) * step - i * lat;
                        if (C[t] < cm) { cm = C[t]; best = t; }
                    }
                    if (room > 1 + i && room_pos > i) {
                        const int t = center + (2 + i) * step + i * lat;
                        if (C[t] < cm) { cm = C[t]; best = t; }
                    }
                }
                pos[arm] = best;
            }
        }
        if (flag[arm])
            for (int v = 0; v < nsrc; ++v) cost_array[arm][v] = orc_ncc(p, px, py, P + (size_t)pos[arm] * 4, v + 1);
    }
    /* view selection priors, ACMMP.cu:994-1008 */
    float view_weights[32], priors[32], probs[32];
    memset(view_weights, 0, sizeof view_weights);
    memset(priors, 0, sizeof priors);
    memset(probs, 0, sizeof probs);
    const int nbpos[4] = {center - W, center + W, center - 1, center + 1};
    for (int i = 0; i < 4; ++i)
        if (flag[2 * i])
            for (int j = 0; j < nsrc; ++j) priors[j] += ((p->sel[nbpos[i]] >> j) & 1u) ? 0.9f : 0.1f;
    const float thr = (float)(0.8 * expf((iter) * (iter) / (-90.0f)));
    for (int i = 0; i < nsrc; ++i) {
        float count = 0, tmpw = 0;
        int count_false = 0;
        for (int j = 0; j < 8; ++j) {
            if (cost_array[j][i] < thr) { tmpw += expf(cost_array[j][i] * cost_array[j][i] / (-0.18f)); count++; }
            if (cost_array[j][i] > 1.2f) count_false++;
        }
        if (count > 2 && count_false < 3) probs[i] = tmpw / count;
        else if (count_false < 3) probs[i] = expf(thr * thr / (-0.32f));
        probs[i] = probs[i] * priors[i];
    }
    { /* TransformPDFToCDF, ACMMP.cu:107-121 */
        float s = 0.0f;
        for (int i = 0; i < nsrc; ++i) s += probs[i];
        const float inv = 1.0f / s;
        float cum = 0.0f;
        for (int i = 0; i < nsrc; ++i) { cum += probs[i] * inv; probs[i] = cum; }
    }
    uint32_t *rng = p->rng6 + (size_t)center * 6;
    for (int s = 0; s < 15; ++s) {
        const float r = orc_rng_uniform(rng) - 1.1920929e-07f;
        for (int v = 0; v < nsrc; ++v)
            if (probs[v] > r) { view_weights[v] += 1.0f; break; }
    }
    uint32_t tsel = 0;
    float wn = 0;
    for (int i = 0; i < nsrc; ++i)
        if (view_weights[i] > 0) { tsel |= 1u << i; wn += view_weights[i]; }
    float final_costs[8];
    for (int i = 0; i < 8; ++i) {
        float fc = 0.0f;
        for (int j = 0; j < nsrc; ++j)
            if (view_weights[j] > 0) {
                if (p->geom) {
                    if (flag[i]) fc += view_weights[j] * (cost_array[i][j] + 0.2f * orc_geom_cost(p, px, py, P + (size_t)pos[i] * 4, j + 1));
                    else fc += view_weights[j] * (cost_array[i][j] + 0.1f * 3.0f);
                } else fc += view_weights[j] * cost_array[i][j];
            }
        final_costs[i] = fc / wn;
    }
    int min_idx = 0; /* FindMinCostIndex: `<=`, last minimum wins (ACMMP.cu:50-61) */
    for (int i = 1; i < 8; ++i)
        if (final_costs[i] <= final_costs[min_idx]) min_idx = i;
    float plane_c[4];
    memcpy(plane_c, P + (size_t)center * 4, sizeof plane_c);
    float cost_now = 0.0f;
    for (int i = 0; i < nsrc; ++i) {
        if (!(view_weights[i] > 0)) continue; /* weight 0 adds +0 in the reference */
        float c = orc_ncc(p, px, py, plane_c, i + 1);
        if (p->geom) c = c + 0.2f * orc_geom_cost(p, px, py, plane_c, i + 1);
        cost_now += view_weights[i] * c;
    }
    cost_now /= wn;
    float live_plane[4], live_cost = cost_now;
    memcpy(live_plane, plane_c, sizeof live_plane);
    int write_sel = 0;
    float depth_now = orc_plane_depth(rc, plane_c, px, py);
    float restricted_cost = 0.0f;
    const float dmin = p->depth_min, dmax = p->depth_max;
    const float depth_sigma = (dmax - dmin) / 64.0f;
    const float tds2 = 2 * depth_sigma * depth_sigma;
    const float angle_sigma = (float)(M_PI * (5.0f / 180.0f));
    const float tas2 = 2 * angle_sigma * angle_sigma;
    const float beta = 0.18f;
    const uint32_t mask_c = p->planar_prior ? p->masks[center] : 0u;
    const float *prior_c = p->planar_prior ? p->prior_planes + (size_t)center * 4 : NULL;
    float plane_now[4];
    memcpy(plane_now, plane_c, sizeof plane_now);
    if (p->planar_prior) { /* ACMMP.cu:1095-1147 */
        const float depth_prior = orc_plane_depth(rc, prior_c, px, py);
        if (mask_c > 0) {
            float rf[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int i = 0; i < 8; ++i)
                if (flag[i]) {
                    const float *q = P + (size_t)pos[i] * 4;
                    const float dd = orc_plane_depth(rc, q, px, py) - depth_prior;
                    const float ac = prior_c[0] * q[0] + prior_c[1] * q[1] + prior_c[2] * q[2];
                    rf[i] = expf(-final_costs[i] * final_costs[i] / beta) * prior_factor(dd, ac, tds2, tas2);
                }
            int max_idx = 0; /* FindMaxCostIndex: `>=`, last maximum wins */
            for (int i = 1; i < 8; ++i)
                if (rf[i] >= rf[max_idx]) max_idx = i;
            const float dd = orc_plane_depth(rc, plane_c, px, py) - depth_prior;
            const float ac = prior_c[0] * plane_c[0] + prior_c[1] * plane_c[1] + prior_c[2] * plane_c[2];
            const float rnow = expf(-cost_now * cost_now / beta) * prior_factor(dd, ac, tds2, tas2);
            memcpy(plane_now, live_plane, sizeof plane_now);
            if (flag[max_idx]) {
                const float *q = P + (size_t)pos[max_idx] * 4;
                const float db = orc_plane_depth(rc, q, px, py);
                if (db >= dmin && db <= dmax && rf[max_idx] > rnow) {
                    /* not depth_now: the reference's assignment hits a shadowing local (ACMMP.cu:1120,1129) */
                    memcpy(live_plane, q, sizeof live_plane);
                    live_cost = final_costs[max_idx];
                    restricted_cost = rf[max_idx];
                    write_sel = 1;
                }
                memcpy(plane_now, p->compat == 0 ? q : live_plane, sizeof plane_now);
            }
        } else {
            memcpy(plane_now, live_plane, sizeof plane_now);
            if (flag[min_idx]) {
                const float *q = P + (size_t)pos[min_idx] * 4;
                const float db = orc_plane_depth(rc, q, px, py);
                if (db >= dmin && db <= dmax && final_costs[min_idx] < cost_now) {
                    depth_now = db;
                    memcpy(live_plane, q, sizeof live_plane);
                    live_cost = final_costs[min_idx];
                }
                memcpy(plane_now, p->compat == 0 ? q : live_plane, sizeof plane_now);
            }
        }
        if (p->compat == 1) cost_now = live_cost;
    } else if (flag[min_idx]) { /* ACMMP.cu:1149-1159 */
        const float *q = P + (size_t)pos[min_idx] * 4;
        const float db = orc_plane_depth(rc, q, px, py);
        if (p->compat == 0) memcpy(plane_now, q, sizeof plane_now);
        if (db >= dmin && db <= dmax && final_costs[min_idx] < cost_now) {
            depth_now = db;
            memcpy(plane_now, q, sizeof plane_now);
            cost_now = final_costs[min_idx];
            write_sel = 1;
        }
    }
    /* PlaneHypothesisRefinement, ACMMP.cu:707-784 */
    {
        const float pert = 0.02f;
        const int use_prior = p->planar_prior && mask_c > 0;
        float depth_prior = 0.0f, depth_rand, n_rand[3];
        if (use_prior) {
            depth_prior = orc_plane_depth(rc, prior_c, px, py);
            depth_rand = orc_rng_uniform(rng) * 6 * depth_sigma + (depth_prior - 3 * depth_sigma);
            perturbed_normal(rc, px, py, prior_c, rng, angle_sigma, n_rand);
        } else {
            depth_rand = orc_rng_uniform(rng) * (dmax - dmin) + dmin;
            random_normal(rc, px, py, rng, depth_now, n_rand);
        }
        float depth_pert = depth_now;
        const float pmin = (1 - pert) * depth_pert, pmax = (1 + pert) * depth_pert;
        depth_pert = orc_rng_uniform(rng) * (pmax - pmin) + pmin; /* the do/while never repeats, :734-736 */
        float n_pert[3];
        perturbed_normal(rc, px, py, plane_now, rng, (float)(pert * M_PI), n_pert);
        const float cd[5] = {depth_rand, depth_now, depth_rand, depth_now, depth_pert};
        float cn[5][4];
        memcpy(cn[0], plane_now, 3 * sizeof(float));
        memcpy(cn[1], n_rand, 3 * sizeof(float));
        memcpy(cn[2], n_rand, 3 * sizeof(float));
        memcpy(cn[3], n_pert, 3 * sizeof(float));
        memcpy(cn[4], plane_now, 3 * sizeof(float));
        for (int i = 0; i < 5; ++i) cn[i][3] = orc_plane_offset(rc, px, py, cd[i], cn[i]);
        for (int i = 0; i < 5; ++i) {
            float tc = 0.0f;
            for (int j = 0; j < nsrc; ++j)
                if (view_weights[j] > 0) {
                    float c = orc_ncc(p, px, py, cn[i], j + 1);
                    if (p->geom) c = c + 0.2f * orc_geom_cost(p, px, py, cn[i], j + 1);
                    tc += view_weights[j] * c;
                }
            tc /= wn;
            const float db = orc_plane_depth(rc, cn[i], px, py);
            if (use_prior) {
                const float dd = cd[i] - depth_prior;
                const float ac = prior_c[0] * cn[i][0] + prior_c[1] * cn[i][1] + prior_c[2] * cn[i][2];
                const float rt = expf(-tc * tc / beta) * prior_factor(dd, ac, tds2, tas2);
                if (db >= dmin && db <= dmax && rt > restricted_cost) {
                    depth_now = db; memcpy(plane_now, cn[i], sizeof plane_now); cost_now = tc; restricted_cost = rt;
                }
            } else if (db >= dmin && db <= dmax && tc < cost_now) {
                depth_now = db; memcpy(plane_now, cn[i], sizeof plane_now); cost_now = tc;
            }
        }
    }
    /* write back, ACMMP.cu:1163-1172 */
    float out_cost = cost_now;
    const float *out_plane = plane_now;
    if (p->hierarchy && !(cost_now < p->pre_costs[center] - 0.1f)) { out_cost = live_cost; out_plane = live_plane; }
    p->costs[center] = out_cost;
    memcpy(p->planes + (size_t)center * 4, out_plane, 4 * sizeof(float));
    if (write_sel) p->sel[center] = tsel;
}

void orc_sweep(orc_problem *p, int iter, int color, int x0, int y0, int x1, int y1) {
    const int W = p->cams[0].width, H = p->cams[0].height;
    const size_t n = (size_t)W * H;
    float *P = (float *)malloc(n * 4 * sizeof(float));
    float *C = (float *)malloc(n * sizeof(float));
    memcpy(P, p->planes, n * 4 * sizeof(float));
    memcpy(C, p->costs, n * sizeof(float));
    for (int py = y0; py < y1; ++py)
        for (int px = x0; px < x1; ++px)
            if (((px + py) & 1) == color) sweep_pixel(p, P, C, iter, px, py);
    free(P);
    free(C);
}

void orc_depth_normal(orc_problem *p) { /* GetDepthandNormal, ACMMP.cu:1199-1212 */
    const orc_camera *rc = &p->cams[0];
    const int W = rc->width, H = rc->height;
    for (int py = 0; py < H; ++py)
        for (int px = 0; px < W; ++px) {
            float *pl = p->planes + ((size_t)py * W + px) * 4;
            const float d = orc_plane_depth(rc, pl, px, py);
            float nw[3];
            to_world_n(rc, pl, nw);
            pl[0] = nw[0]; pl[1] = nw[1]; pl[2] = nw[2]; pl[3] = d;
        }
}

void orc_median_filter(orc_problem *p, int color) { /* CheckerboardFilter, ACMMP.cu:1214-1328 */
    const int W = p->cams[0].width, H = p->cams[0].height;
    static const int off[20][2] = {{0, -1}, {0, -3}, {0, -5}, {0, 1}, {0, 3}, {0, 5}, {-1, 0}, {-3, 0}, {-5, 0}, {1, 0},
                                   {3, 0}, {5, 0}, {2, -1}, {2, 1}, {-2, -1}, {-2, 1}, {-1, -2}, {1, -2}, {-1, 2}, {1, 2}};
    for (int py = 0; py < H; ++py)
        for (int px = 0; px < W; ++px) {
            if (((px + py) & 1) != color) continue;
            const size_t c = (size_t)py * W + px;
            if (p->costs[c] < 0.001f) continue;
            float f[21];
            int n = 0;
            f[n++] = p->planes[c * 4 + 3];
            for (int k = 0; k < 20; ++k) {
                const int dx = off[k][0], dy = off[k][1];
                /* the reference's guards (ACMMP.cu:1249-1319) */
                int ok = 1;
                if (dy == -2) ok = py > 2; /* (+-1,-2) is guarded by p.y > 2, not p.y > 1 */
                else if (dy < 0) ok = py > -dy - 1;
                else if (dy > 0) ok = py < H - dy;
                if (dx < 0 && !(px > -dx - 1)) ok = 0;
                if (dx > 0 && !(px < W - dx)) ok = 0;
                if (ok) f[n++] = p->planes[(size_t)((long)c + (long)dy * W + dx) * 4 + 3];
            }
            for (int i = 1; i < n; ++i) {
                const float t = f[i];
                int j = i;
                for (; j >= 1 && t < f[j - 1]; --j) f[j] = f[j - 1];
                f[j] = t;
            }
            const int m = n / 2;
            p->planes[c * 4 + 3] = (n % 2 == 0) ? (f[m - 1] + f[m]) / 2 : f[m];
        }
}

void orc_run(orc_problem *p, int mode, int iterations) { /* ACMMP::RunPatchMatch, ACMMP.cu:1415-1448 */
    const int W = p->cams[0].width, H = p->cams[0].height;
    orc_init(p, mode, 0, 0, W, H);
    for (int it = 0; it < iterations; ++it) {
        orc_sweep(p, it, 0, 0, 0, W, H);
        orc_sweep(p, it, 1, 0, 0, W, H);
    }
    orc_depth_normal(p);
    orc_median_filter(p, 0);
    orc_median_filter(p, 1);
}
