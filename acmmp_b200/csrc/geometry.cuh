// geometry.cuh - plane / depth / normal helpers of the PatchMatch path.
//
// These are the small closed-form pieces of the reference's per-pixel math
// (ACMMP.cu:123-260, 333-351, 480-543).  They sit outside the 36-tap loop, so
// they are written as ordinary float expressions and compiled with the
// reference's own arithmetic flags (--use_fast_math); the association order of
// every sum/product below is the one the reference's expressions have, because
// discrete decisions (arg-min, adoption, range gates) hang on the last bit.
#pragma once
#include "common.cuh"

namespace acmmp {

struct Plane {
    float x, y, z, w;
};

ACMMP_DEV Plane make_plane(float4 v) { return Plane{v.x, v.y, v.z, v.w}; }
ACMMP_DEV float4 to_float4(const Plane &p) { return make_float4(p.x, p.y, p.z, p.w); }

// camera-frame point of pixel (px,py) at `depth`            [ACMMP.cu:123-128]
ACMMP_DEV void backproject(const RefConst &rc, int px, int py, float depth, float X[3]) {
    X[0] = depth * (px - rc.K2) / rc.K0;
    X[1] = depth * (py - rc.K5) / rc.K4;
    X[2] = depth;
}

// unit viewing ray through the pixel                          [ACMMP.cu:130-142]
ACMMP_DEV void view_direction(const RefConst &rc, int px, int py, float depth, float v[3]) {
    float X[3];
    backproject(rc, px, py, depth, X);
    float norm = sqrtf(X[0] * X[0] + X[1] * X[1] + X[2] * X[2]);
    v[0] = X[0] / norm;
    v[1] = X[1] / norm;
    v[2] = X[2] / norm;
}

// plane offset d = -n.X                                        [ACMMP.cu:144-149]
ACMMP_DEV float plane_offset(const RefConst &rc, int px, int py, float depth, const Plane &n) {
    float X[3];
    backproject(rc, px, py, depth, X);
    return -(n.x * X[0] + n.y * X[1] + n.z * X[2]);
}

// depth of plane (n,d) along the pixel's ray                   [ACMMP.cu:163-168]
ACMMP_DEV float plane_depth(const RefConst &rc, const Plane &pl, int px, int py) {
    return -pl.w * rc.K0 /
           ((px - rc.K2) * pl.x + (rc.K0 / rc.K4) * (py - rc.K5) * pl.y + rc.K0 * pl.z);
}

ACMMP_DEV void normalize3(Plane &v) {
    const float n2 = v.x * v.x + v.y * v.y + v.z * v.z;
    const float inv = rsqrtf(n2);
    v.x *= inv;
    v.y *= inv;
    v.z *= inv;
}

// uniform random unit normal facing the camera (Marsaglia)     [ACMMP.cu:170-196]
ACMMP_DEV Plane random_normal(const RefConst &rc, int px, int py, Rng &rng, float depth) {
    float q1 = 1.0f, q2 = 1.0f, s = 2.0f;
    while (s >= 1.0f) {
        q1 = 2.0f * rng_uniform(rng) - 1.0f;
        q2 = 2.0f * rng_uniform(rng) - 1.0f;
        s = q1 * q1 + q2 * q2;
    }
    const float sq = sqrtf(1.0f - s);
    Plane n;
    n.x = 2.0f * q1 * sq;
    n.y = 2.0f * q2 * sq;
    n.z = 1.0f - 2.0f * s;
    n.w = 0.0f;
    float v[3];
    view_direction(rc, px, py, depth, v);
    const float dot = n.x * v[0] + n.y * v[1] + n.z * v[2];
    if (dot > 0.0f) {
        n.x = -n.x;
        n.y = -n.y;
        n.z = -n.z;
    }
    normalize3(n);
    return n;
}

// small random rotation of a normal; reverted if it turns away  [ACMMP.cu:198-233]
ACMMP_DEV Plane perturbed_normal(const RefConst &rc, int px, int py, const Plane &normal,
                                 Rng &rng, float perturbation) {
    float v[3];
    view_direction(rc, px, py, 1.0f, v);
    const float a1 = (rng_uniform(rng) - 0.5f) * perturbation;
    const float a2 = (rng_uniform(rng) - 0.5f) * perturbation;
    const float a3 = (rng_uniform(rng) - 0.5f) * perturbation;
    const float s1 = sinf(a1), s2 = sinf(a2), s3 = sinf(a3);
    const float c1 = cosf(a1), c2 = cosf(a2), c3 = cosf(a3);
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
    Plane np;
    np.x = R[0] * normal.x + R[1] * normal.y + R[2] * normal.z;
    np.y = R[3] * normal.x + R[4] * normal.y + R[5] * normal.z;
    np.z = R[6] * normal.x + R[7] * normal.y + R[8] * normal.z;
    np.w = normal.w;
    if (np.x * v[0] + np.y * v[1] + np.z * v[2] >= 0.0f) {
        np = normal;
    }
    normalize3(np);
    return np;
}

// camera -> world (R^T n) and world -> camera (R n)             [ACMMP.cu:333-351]
ACMMP_DEV Plane normal_to_world(const RefConst &rc, const Plane &p) {
    Plane o;
    o.x = rc.R[0] * p.x + rc.R[3] * p.y + rc.R[6] * p.z;
    o.y = rc.R[1] * p.x + rc.R[4] * p.y + rc.R[7] * p.z;
    o.z = rc.R[2] * p.x + rc.R[5] * p.y + rc.R[8] * p.z;
    o.w = p.w;
    return o;
}
ACMMP_DEV Plane normal_to_cam(const RefConst &rc, const Plane &p) {
    Plane o;
    o.x = rc.R[0] * p.x + rc.R[1] * p.y + rc.R[2] * p.z;
    o.y = rc.R[3] * p.x + rc.R[4] * p.y + rc.R[5] * p.z;
    o.z = rc.R[6] * p.x + rc.R[7] * p.y + rc.R[8] * p.z;
    o.w = p.w;
    return o;
}

// forward-backward reprojection error against a source depth map, clamped to
// 3 px                                                          [ACMMP.cu:480-543]
ACMMP_DEV float geom_consistency_cost(const RefConst &rc, const PairConst &pc, const Plane &pl,
                                      int px, int py) {
    const float max_cost = 3.0f;
    const float depth = plane_depth(rc, pl, px, py);
    // reference camera -> world
    float X0 = depth * (px - rc.K[2]) / rc.K[0];
    float X1 = depth * (py - rc.K[5]) / rc.K[4];
    float X2 = depth;
    float T0 = rc.R[0] * X0 + rc.R[3] * X1 + rc.R[6] * X2;
    float T1 = rc.R[1] * X0 + rc.R[4] * X1 + rc.R[7] * X2;
    float T2 = rc.R[2] * X0 + rc.R[5] * X1 + rc.R[8] * X2;
    float C0 = -(rc.R[0] * rc.t[0] + rc.R[3] * rc.t[1] + rc.R[6] * rc.t[2]);
    float C1 = -(rc.R[1] * rc.t[0] + rc.R[4] * rc.t[1] + rc.R[7] * rc.t[2]);
    float C2 = -(rc.R[2] * rc.t[0] + rc.R[5] * rc.t[1] + rc.R[8] * rc.t[2]);
    const float W0 = T0 + C0, W1 = T1 + C1, W2 = T2 + C2;
    // world -> source pixel
    float s0 = pc.Rs[0] * W0 + pc.Rs[1] * W1 + pc.Rs[2] * W2 + pc.ts[0];
    float s1 = pc.Rs[3] * W0 + pc.Rs[4] * W1 + pc.Rs[5] * W2 + pc.ts[1];
    float s2 = pc.Rs[6] * W0 + pc.Rs[7] * W1 + pc.Rs[8] * W2 + pc.ts[2];
    const float sd = pc.Ks[6] * s0 + pc.Ks[7] * s1 + pc.Ks[8] * s2;
    const float sx = (pc.Ks[0] * s0 + pc.Ks[1] * s1 + pc.Ks[2] * s2) / sd;
    const float sy = (pc.Ks[3] * s0 + pc.Ks[4] * s1 + pc.Ks[5] * s2) / sd;
    // point sample of the source depth at the truncated pixel, clamp-to-edge
    // (the reference fetches a linear-filter texture at the texel centre
    // (int)x+0.5, which returns that texel exactly; ACMMP.cu:528)
    int ix = (int)sx, iy = (int)sy;
    ix = min(max(ix, 0), pc.dw - 1);
    iy = min(max(iy, 0), pc.dh - 1);
    const float src_depth = __ldg(pc.depth + (size_t)iy * pc.dw + ix);
    if (src_depth == 0.0f) return max_cost;
    // source pixel at its own depth -> world
    float Y0 = src_depth * (sx - pc.Ks[2]) / pc.Ks[0];
    float Y1 = src_depth * (sy - pc.Ks[5]) / pc.Ks[4];
    float Y2 = src_depth;
    float U0 = pc.Rs[0] * Y0 + pc.Rs[3] * Y1 + pc.Rs[6] * Y2;
    float U1 = pc.Rs[1] * Y0 + pc.Rs[4] * Y1 + pc.Rs[7] * Y2;
    float U2 = pc.Rs[2] * Y0 + pc.Rs[5] * Y1 + pc.Rs[8] * Y2;
    float D0 = -(pc.Rs[0] * pc.ts[0] + pc.Rs[3] * pc.ts[1] + pc.Rs[6] * pc.ts[2]);
    float D1 = -(pc.Rs[1] * pc.ts[0] + pc.Rs[4] * pc.ts[1] + pc.Rs[7] * pc.ts[2]);
    float D2 = -(pc.Rs[2] * pc.ts[0] + pc.Rs[5] * pc.ts[1] + pc.Rs[8] * pc.ts[2]);
    const float V0 = U0 + D0, V1 = U1 + D1, V2 = U2 + D2;
    // world -> reference pixel
    float r0 = rc.R[0] * V0 + rc.R[1] * V1 + rc.R[2] * V2 + rc.t[0];
    float r1 = rc.R[3] * V0 + rc.R[4] * V1 + rc.R[5] * V2 + rc.t[1];
    float r2 = rc.R[6] * V0 + rc.R[7] * V1 + rc.R[8] * V2 + rc.t[2];
    const float rd = rc.K[6] * r0 + rc.K[7] * r1 + rc.K[8] * r2;
    const float bx = (rc.K[0] * r0 + rc.K[1] * r1 + rc.K[2] * r2) / rd;
    const float by = (rc.K[3] * r0 + rc.K[4] * r1 + rc.K[5] * r2) / rd;
    const float dc = px - bx;
    const float dr = py - by;
    return fminf(max_cost, sqrtf(dc * dc + dr * dr));
}

}  // namespace acmmp
