// fusion.cu - depth-map fusion, compiled WITHOUT --use_fast_math:
// the reference runs this on the host in IEEE float / double arithmetic.
//
// Reference: RunFusion, /root/reference/src/acmmp_definitions.cpp:828-1043 with
// Get3DPointonWorld / ProjectonCamera / GetAngle (/root/reference/src/ACMMP.cpp:204-262).
// The reference is one sequential loop over views and, inside a view, over pixels in row
// order: every accepted pixel masks the source pixels it was consistent with, later
// pixels (and later views) skip masked source pixels.  `used_list` is declared once per
// VIEW (:936), so an accepted pixel also masks the source coordinates left behind by
// EARLIER pixels for the sources it is not consistent with itself (H9) - kept.
//
// Here, per view:
//   1. fusion_score_kernel (one thread per pixel): everything that does not depend on
//      the masks - for each source j the source pixel hit and exp(-(reproj + 200 rel +
//      10 angle)) if the geometric test passes.
//   2. the greedy pass as a fixed-point iteration: a pixel's state (consistent sources,
//      accepted) depends only on states of EARLIER pixels through "source pixel x was
//      masked at time q < p".  Each round recomputes every pixel in parallel from the
//      mask times of the previous round (decide kernel), then rebuilds the mask times:
//      the source pixel of a consistent (p, j) is masked at the first accepted pixel q >= p
//      that comes before the next pixel consistent with j - two suffix-min scans fused
//      with the atomicMin scatter (scan kernels A, B, C).  The fixed point is unique and
//      equals the sequential result (induction over the pixel order); rounds needed = the
//      longest dependency chain, a handful in practice.
//   3. masks |= marks; accepted pixels are compacted in pixel order (same point order as
//      the reference's PLY) into (x, y, z, nx, ny, nz, c0, c1, c2) floats.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/acmmp_b200.h"

namespace {

constexpr int kInf = 0x7fffffff;
constexpr int kChunk = 2048;          // pixels per scan block
constexpr int kScanThreads = 256;     // 8 pixels per thread
constexpr int kPerThread = kChunk / kScanThreads;

struct FusionCam {
    float K[9], R[9], t[3];
    int width, height;
    float depth_max;
};

struct FusionView {
    float *depth = nullptr;      // P
    float *normal = nullptr;     // 3P
    uint8_t *color = nullptr;    // 3P or null
    float *depth2 = nullptr;     // P: second map set (prior-aware fusion), else null
    float *normal2 = nullptr;    // 3P
    bool set2 = false;
    uint8_t *mask0 = nullptr;    // P: caller's mask (1 = ignore the pixel), kept so that a run can be repeated
    uint8_t *mask = nullptr;     // P: working mask: 1 = ignore / already used
    int *mark = nullptr;         // P: time (pixel index of the view being fused) at which the pixel gets masked
    bool set = false;
};

__device__ __forceinline__ void world_point(const FusionCam &c, int x, int y, float depth, float W[3]) {
    const float X0 = depth * (x - c.K[2]) / c.K[0];
    const float X1 = depth * (y - c.K[5]) / c.K[4];
    const float X2 = depth;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float t = c.R[k] * X0 + c.R[3 + k] * X1 + c.R[6 + k] * X2;
        const float C = -(c.R[k] * c.t[0] + c.R[3 + k] * c.t[1] + c.R[6 + k] * c.t[2]);
        W[k] = t + C;
    }
}

__device__ __forceinline__ void project(const FusionCam &c, const float W[3], float &x, float &y, float &d) {
    float t[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) t[k] = c.R[3 * k] * W[0] + c.R[3 * k + 1] * W[1] + c.R[3 * k + 2] * W[2] + c.t[k];
    d = c.K[6] * t[0] + c.K[7] * t[1] + c.K[8] * t[2];
    x = (c.K[0] * t[0] + c.K[1] * t[1] + c.K[2] * t[2]) / d;
    y = (c.K[3] * t[0] + c.K[4] * t[1] + c.K[5] * t[2]) / d;
}

struct ScoreArgs {
    FusionCam ref;
    FusionCam src[ACMMP_MAX_SRC_VIEWS];
    const float *src_depth[ACMMP_MAX_SRC_VIEWS];
    const float *src_normal[ACMMP_MAX_SRC_VIEWS];
    int nsrc;
};

// loc[j * P + p] = source pixel index hit by pixel p in source j if the geometric test of
// acmmp_definitions.cpp:973-990 passes (ignoring masks), else -1;  w = exp(-index).
// valid[p] = 0 where the reference skips the pixel altogether (:946-952).
__global__ void fusion_score_kernel(const __grid_constant__ ScoreArgs a, const float *depth, const float *normal,
                                    const uint8_t *mask, int *loc, float *w, uint8_t *valid) {
    const int P = a.ref.width * a.ref.height;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const int r = p / a.ref.width, c = p - r * a.ref.width;
    const float ref_depth = depth[p];
    const bool ok = mask[p] != 1 && !(ref_depth <= 0.0f || ref_depth >= a.ref.depth_max);
    valid[p] = ok ? 1 : 0;
    float X[3];
    if (ok) world_point(a.ref, c, r, ref_depth, X);
    const float n0 = normal[3 * (size_t)p], n1 = normal[3 * (size_t)p + 1], n2 = normal[3 * (size_t)p + 2];
    for (int j = 0; j < a.nsrc; ++j) {
        int l = -1;
        float wt = 0.0f;
        if (ok) {
            const FusionCam &s = a.src[j];
            float qx, qy, pd;
            project(s, X, qx, qy, pd);
            const int src_r = (int)(qy + 0.5f), src_c = (int)(qx + 0.5f);
            if (src_c >= 0 && src_c < s.width && src_r >= 0 && src_r < s.height) {
                const int sidx = src_r * s.width + src_c;
                const float src_depth = a.src_depth[j][sidx];
                if (src_depth > 0.0f) {
                    float T[3], tx, ty;
                    world_point(s, src_c, src_r, src_depth, T);
                    project(a.ref, T, tx, ty, pd);
                    const double dx = (double)(c - tx), dy = (double)(r - ty);
                    const float reproj = (float)sqrt(dx * dx + dy * dy);
                    const float rel = fabsf(pd - ref_depth) / ref_depth;
                    const float *sn = a.src_normal[j] + 3 * (size_t)sidx;
                    const float dot = n0 * sn[0] + n1 * sn[1] + n2 * sn[2];
                    float ang = acosf(dot);
                    if (ang != ang) ang = 0.0f;
                    if (reproj < 2.0f && rel < 0.01f && ang < 0.174533f) {
                        l = sidx;
                        const float ti = reproj + 200 * rel + ang * 10;
                        wt = expf(-ti);
                    }
                }
            }
        }
        loc[(size_t)j * P + p] = l;
        w[(size_t)j * P + p] = wt;
    }
}

struct IterArgs {
    int *mark[ACMMP_MAX_SRC_VIEWS];   // per source: mask time of every source pixel
    int nsrc;
    int P;
    float consistency_scalar;
    int con_num_thresh;
};

// state of pixel p given the mask times: bit j of cons = consistent with source j and the
// source pixel not masked before p; appr = accepted (:1002-1004)
__global__ void fusion_decide_kernel(const __grid_constant__ IterArgs a, const int *loc, const float *w, const uint8_t *valid,
                                     uint32_t *cons, uint8_t *appr, int *changed) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.P) return;
    uint32_t bits = 0u;
    uint8_t ap = 0;
    if (valid[p]) {
        int num = 0;
        float dyn = 0.0f;
        for (int j = 0; j < a.nsrc; ++j) {
            const int l = loc[(size_t)j * a.P + p];
            if (l >= 0 && !(a.mark[j][l] < p)) {
                bits |= 1u << j;
                dyn += w[(size_t)j * a.P + p];
                num++;
            }
        }
        ap = (num >= a.con_num_thresh && dyn > a.consistency_scalar * num) ? 1 : 0;
    }
    if (bits != cons[p] || ap != appr[p]) {
        cons[p] = bits;
        appr[p] = ap;
        *changed = 1;
    }
}

// ---- suffix-min scans over the pixel order, fused with the mask-time scatter -----------------
// array 0: accepted flags, array 1 + j: "consistent with source j" flags.  value = pixel
// index where the flag is set, else +inf; suffix-min = next flagged pixel at or after p.

__device__ __forceinline__ int warp_suffix_min_inclusive(int v) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_down_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) + d < 32) v = min(v, o);
    }
    return v;
}

// A: per chunk, the smallest flagged pixel of every array
__global__ void __launch_bounds__(kScanThreads) fusion_chunk_min_kernel(const uint32_t *cons, const uint8_t *appr, int P, int narr,
                                                                        int nchunks, int *agg) {
    __shared__ int s_warp[ACMMP_MAX_SRC_VIEWS + 1][kScanThreads / 32];
    const int base = blockIdx.x * kChunk + threadIdx.x * kPerThread;
    uint32_t bits[kPerThread];
    uint32_t ap = 0u;
#pragma unroll
    for (int k = 0; k < kPerThread; ++k) {
        const int p = base + k;
        bits[k] = p < P ? cons[p] : 0u;
        if (p < P && appr[p]) ap |= 1u << k;
    }
    for (int a = 0; a < narr; ++a) {
        int m = kInf;
#pragma unroll
        for (int k = kPerThread - 1; k >= 0; --k) {
            const bool f = a == 0 ? ((ap >> k) & 1u) : ((bits[k] >> (a - 1)) & 1u);
            if (f) m = base + k;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) m = min(m, __shfl_xor_sync(0xffffffffu, m, d));
        if ((threadIdx.x & 31) == 0) s_warp[a][threadIdx.x >> 5] = m;
    }
    __syncthreads();
    if (threadIdx.x < narr) {
        int m = kInf;
#pragma unroll
        for (int k = 0; k < kScanThreads / 32; ++k) m = min(m, s_warp[threadIdx.x][k]);
        agg[threadIdx.x * nchunks + blockIdx.x] = m;
    }
}

// B: exclusive suffix-min over the chunk aggregates; one warp per array
__global__ void fusion_agg_scan_kernel(const int *agg, int nchunks, int *carry) {
    const int a = blockIdx.x, lane = threadIdx.x;
    int run = kInf;
    for (int tile = (nchunks + 31) / 32 - 1; tile >= 0; --tile) {
        const int b = tile * 32 + lane;
        const int v = b < nchunks ? agg[a * nchunks + b] : kInf;
        const int incl = warp_suffix_min_inclusive(v);
        int excl = __shfl_down_sync(0xffffffffu, incl, 1);
        if (lane == 31) excl = kInf;
        if (b < nchunks) carry[a * nchunks + b] = min(excl, run);
        run = min(run, __shfl_sync(0xffffffffu, incl, 0));
    }
}

// block-wide exclusive suffix-min of one value per thread (+ carry of the following chunks)
__device__ __forceinline__ int block_suffix_min_exclusive(int v, int carry, int *s_warp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int incl = warp_suffix_min_inclusive(v);
    int excl = __shfl_down_sync(0xffffffffu, incl, 1);
    if (lane == 31) excl = kInf;
    __syncthreads();   // s_warp free again
    if (lane == 0) s_warp[warp] = incl;
    __syncthreads();
    int rest = carry;
    for (int k = warp + 1; k < kScanThreads / 32; ++k) rest = min(rest, s_warp[k]);
    return min(excl, rest);
}

// C: rescan every chunk with its carry and scatter the mask times:
// the source pixel of a consistent (p, j) is masked at time q = next accepted pixel >= p
// provided no other pixel consistent with j lies in (p, q] (that one overwrites
// used_list[j] before q's marks are written, acmmp_definitions.cpp:992-993,1024-1032).
__global__ void __launch_bounds__(kScanThreads) fusion_mark_kernel(const __grid_constant__ IterArgs a, const uint32_t *cons,
                                                                   const uint8_t *appr, const int *loc, const int *carry, int nchunks) {
    __shared__ int s_warp[kScanThreads / 32];
    const int P = a.P;
    const int base = blockIdx.x * kChunk + threadIdx.x * kPerThread;
    uint32_t bits[kPerThread];
    uint32_t ap = 0u;
#pragma unroll
    for (int k = 0; k < kPerThread; ++k) {
        const int p = base + k;
        bits[k] = p < P ? cons[p] : 0u;
        if (p < P && appr[p]) ap |= 1u << k;
    }
    // next accepted pixel at or after each of my pixels
    int next_ap[kPerThread];
    {
        int m = kInf;
#pragma unroll
        for (int k = kPerThread - 1; k >= 0; --k) {
            if ((ap >> k) & 1u) m = base + k;
            next_ap[k] = m;
        }
        const int after = block_suffix_min_exclusive(m, carry[blockIdx.x], s_warp);
#pragma unroll
        for (int k = 0; k < kPerThread; ++k) next_ap[k] = min(next_ap[k], after);
    }
    for (int j = 0; j < a.nsrc; ++j) {
        int m = kInf;
#pragma unroll
        for (int k = kPerThread - 1; k >= 0; --k)
            if ((bits[k] >> j) & 1u) m = base + k;
        const int after = block_suffix_min_exclusive(m, carry[(1 + j) * nchunks + blockIdx.x], s_warp);
        int nxt = after;   // next pixel consistent with j strictly after pixel k
#pragma unroll
        for (int k = kPerThread - 1; k >= 0; --k) {
            if ((bits[k] >> j) & 1u) {
                const int q = next_ap[k];
                if (q != kInf && q < nxt) atomicMin(a.mark[j] + loc[(size_t)j * P + base + k], q);
                nxt = base + k;
            }
        }
    }
}

__global__ void fusion_reset_marks_kernel(const uint8_t *mask, int *mark, int P) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x < P) mark[x] = mask[x] == 1 ? -1 : kInf;
}

__global__ void fusion_commit_marks_kernel(uint8_t *mask, const int *mark, int P) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x < P && mark[x] != kInf) mask[x] = 1;
}

// ---- order-preserving compaction of the accepted pixels -------------------------------------
__global__ void __launch_bounds__(kScanThreads) fusion_count_kernel(const uint8_t *appr, int P, int *cnt) {
    __shared__ int s_warp[kScanThreads / 32];
    const int base = blockIdx.x * kChunk + threadIdx.x * kPerThread;
    int n = 0;
#pragma unroll
    for (int k = 0; k < kPerThread; ++k)
        if (base + k < P && appr[base + k]) n++;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) n += __shfl_xor_sync(0xffffffffu, n, d);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int k = 0; k < kScanThreads / 32; ++k) t += s_warp[k];
        cnt[blockIdx.x] = t;
    }
}

// exclusive prefix sum over the chunk counts (one warp); offs[nchunks] = total
__global__ void fusion_offsets_kernel(const int *cnt, int nchunks, int *offs) {
    const int lane = threadIdx.x;
    int run = 0;
    for (int tile = 0; tile * 32 < nchunks; ++tile) {
        const int b = tile * 32 + lane;
        const int v = b < nchunks ? cnt[b] : 0;
        int incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (b < nchunks) offs[b] = run + incl - v;
        run += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) offs[nchunks] = run;
}

// points: 9 floats each = world point, normal of the reference pixel, colour bytes as stored
// (acmmp_definitions.cpp:954-958,1013-1017)
__global__ void __launch_bounds__(kScanThreads) fusion_emit_kernel(const __grid_constant__ FusionCam cam, const float *depth,
                                                                   const float *normal, const float *depth2, const float *normal2,
                                                                   const uint8_t *color, const uint8_t *appr, const int *offs,
                                                                   float *points) {
    __shared__ int s_warp[kScanThreads / 32];
    const int P = cam.width * cam.height;
    const int base = blockIdx.x * kChunk + threadIdx.x * kPerThread;
    int n = 0;
    uint32_t ap = 0u;
#pragma unroll
    for (int k = 0; k < kPerThread; ++k)
        if (base + k < P && appr[base + k]) {
            ap |= 1u << k;
            n++;
        }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = n;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int before = offs[blockIdx.x] + incl - n;
    for (int k = 0; k < warp; ++k) before += s_warp[k];
#pragma unroll
    for (int k = 0; k < kPerThread; ++k)
        if ((ap >> k) & 1u) {
            const int p = base + k;
            const int r = p / cam.width, c = p - r * cam.width;
            const bool second = appr[p] == 2;   // prior-aware fusion: the second hypothesis won
            const float *nm = (second ? normal2 : normal) + 3 * (size_t)p;
            float X[3];
            world_point(cam, c, r, second ? depth2[p] : depth[p], X);
            float *o = points + (size_t)before * 9;
            o[0] = X[0]; o[1] = X[1]; o[2] = X[2];
            o[3] = nm[0]; o[4] = nm[1]; o[5] = nm[2];
            if (color) {
                o[6] = (float)color[3 * (size_t)p]; o[7] = (float)color[3 * (size_t)p + 1]; o[8] = (float)color[3 * (size_t)p + 2];
            } else {
                o[6] = o[7] = o[8] = 0.0f;
            }
            before++;
        }
}

// ---- prior-aware fusion (RunPriorAwareFusion, acmmp_definitions.cpp:443-826) ------------------
// Two (depth, normal) sets per view.  A pixel tries both of its hypotheses (set 0, set 1) against
// both maps of every source; per hypothesis h and source j the mask-independent part is
//   loc[(h ns + j) P + p] = source pixel if either source map passes the three thresholds, else -1
//   w  [(h ns + j) P + p] = its dynamic consistency (max over the passing source maps)   (:487-516)
// The greedy pass is again a fixed point over mask times, simpler than RunFusion's: an accepted pixel
// masks exactly the consistent source pixels of its winning hypothesis (:816-820), so
// mark[x] = min over accepted pixels p that include x - one atomicMin scatter per round.
struct PriorScoreArgs {
    FusionCam ref;
    FusionCam src[ACMMP_MAX_SRC_VIEWS];
    const float *src_depth[2][ACMMP_MAX_SRC_VIEWS];
    const float *src_normal[2][ACMMP_MAX_SRC_VIEWS];
    int nsrc;
};

// get_consistency_metrics for one source map (:458-484, :498-503): passes = the three thresholds
__device__ __forceinline__ void prior_metric(const FusionCam &ref, const FusionCam &s, int r, int c, float ref_depth, float n0,
                                             float n1, float n2, int src_r, int src_c, float src_depth, const float *sn,
                                             bool &passes, float &dc) {
    float reproj = 1e6f, rel = 1e6f, ang = 1e6f;
    if (src_depth > 0.0f) {
        float T[3], tx, ty, pd;
        world_point(s, src_c, src_r, src_depth, T);
        project(ref, T, tx, ty, pd);
        const double dx = (double)(c - tx), dy = (double)(r - ty);
        reproj = (float)sqrt(dx * dx + dy * dy);
        rel = fabsf(pd - ref_depth) / ref_depth;
        const float dot = n0 * sn[0] + n1 * sn[1] + n2 * sn[2];
        ang = acosf(dot);
        if (ang != ang) ang = 0.0f;
    }
    passes = reproj < 2.0f && rel < 0.01f && ang < 0.174533f;
    const float ti = reproj + 200 * rel + ang * 10;
    dc = (float)exp((double)-ti);
}

__global__ void prior_score_kernel(const __grid_constant__ PriorScoreArgs a, const float *depth0, const float *normal0,
                                   const float *depth1, const float *normal1, const uint8_t *mask, int *loc, float *w,
                                   uint8_t *valid) {
    const int P = a.ref.width * a.ref.height;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const int r = p / a.ref.width, c = p - r * a.ref.width;
    const float d[2] = {depth0[p], depth1[p]};
    const bool ok = mask[p] != 1 && !(d[0] <= 0.0f && d[1] <= 0.0f);     // (:686-697)
    valid[p] = ok ? 1 : 0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const float *nm = (h ? normal1 : normal0) + 3 * (size_t)p;
        const float n0 = nm[0], n1 = nm[1], n2 = nm[2];
        const bool live = ok && d[h] > 0.0f;
        float X[3];
        if (live) world_point(a.ref, c, r, d[h], X);
        for (int j = 0; j < a.nsrc; ++j) {
            int l = -1;
            float wt = 0.0f;
            if (live) {
                const FusionCam &s = a.src[j];
                float qx, qy, pd;
                project(s, X, qx, qy, pd);
                const int src_r = (int)(qy + 0.5f), src_c = (int)(qx + 0.5f);
                if (src_c >= 0 && src_c < s.width && src_r >= 0 && src_r < s.height) {
                    const int sidx = src_r * s.width + src_c;
                    bool p0, p1;
                    float dc0, dc1;
                    prior_metric(a.ref, s, r, c, d[h], n0, n1, n2, src_r, src_c, a.src_depth[0][j][sidx],
                                 a.src_normal[0][j] + 3 * (size_t)sidx, p0, dc0);
                    prior_metric(a.ref, s, r, c, d[h], n0, n1, n2, src_r, src_c, a.src_depth[1][j][sidx],
                                 a.src_normal[1][j] + 3 * (size_t)sidx, p1, dc1);
                    if (p0 || p1) {
                        l = sidx;
                        wt = p0 && p1 ? fmaxf(dc0, dc1) : (p0 ? dc0 : dc1);
                    }
                }
            }
            loc[((size_t)h * a.nsrc + j) * P + p] = l;
            w[((size_t)h * a.nsrc + j) * P + p] = wt;
        }
    }
}

// state of pixel p given the mask times: cons[h P + p] = sources consistent with hypothesis h and not
// masked before p; appr = 0 rejected, 1 + h accepted with hypothesis h (:699-776)
__global__ void prior_decide_kernel(const __grid_constant__ IterArgs a, int single_match_penalty, const int *loc, const float *w,
                                    const uint8_t *valid, uint32_t *cons, uint8_t *appr, int *changed) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.P) return;
    uint32_t bits[2] = {0u, 0u};
    uint8_t ap = 0;
    if (valid[p]) {
        int num[2] = {0, 0};
        bool thr[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float dyn = 0.0f;
            for (int j = 0; j < a.nsrc; ++j) {
                const size_t at = ((size_t)h * a.nsrc + j) * a.P + p;
                const int l = loc[at];
                if (l >= 0 && !(a.mark[j][l] < p)) {
                    bits[h] |= 1u << j;
                    dyn += w[at];
                    num[h]++;
                }
            }
            thr[h] = num[h] >= a.con_num_thresh && dyn > a.consistency_scalar * num[h];
        }
        if (thr[0] && thr[1]) ap = num[1] >= num[0] ? 2 : 1;
        else if (thr[1]) ap = num[1] >= a.con_num_thresh + single_match_penalty ? 2 : 0;
        else ap = thr[0] && num[0] >= a.con_num_thresh + single_match_penalty ? 1 : 0;
    }
    if (bits[0] != cons[p] || bits[1] != cons[(size_t)a.P + p] || ap != appr[p]) {
        cons[p] = bits[0];
        cons[(size_t)a.P + p] = bits[1];
        appr[p] = ap;
        *changed = 1;
    }
}

__global__ void prior_mark_kernel(const __grid_constant__ IterArgs a, const int *loc, const uint32_t *cons, const uint8_t *appr) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.P) return;
    const int ap = appr[p];
    if (!ap) return;
    const int h = ap - 1;
    uint32_t bits = cons[(size_t)h * a.P + p];
    while (bits) {
        const int j = __ffs(bits) - 1;
        bits &= bits - 1;
        atomicMin(a.mark[j] + loc[((size_t)h * a.nsrc + j) * a.P + p], p);
    }
}

}  // namespace

// ---- C ABI ------------------------------------------------------------------------------------

struct acmmp_fusion {
    int device = 0;
    int n = 0;
    std::string error;
    cudaStream_t stream = nullptr;
    std::vector<acmmp_camera> cams;
    std::vector<int> pair_off, pair_ids;
    std::vector<FusionView> views;
    // per-view scratch, sized for the largest view
    size_t cap_pixels = 0;
    int max_src = 0;
    int cap_hyps = 0;
    int *loc = nullptr;
    float *w = nullptr;
    uint8_t *valid = nullptr, *appr = nullptr;
    uint32_t *cons = nullptr;
    int *agg = nullptr, *carry = nullptr, *cnt = nullptr, *offs = nullptr, *changed = nullptr;
    float *points = nullptr;          // device, 9 floats per point
    size_t points_cap = 0, points_n = 0;
    std::vector<int64_t> per_view;    // accepted pixels per view
    std::vector<int> rounds;          // fixed-point rounds per view
    double score_ms = 0.0, greedy_ms = 0.0, emit_ms = 0.0;
};

namespace {

thread_local std::string g_fusion_error;

int ffail(acmmp_fusion *f, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (f) f->error = buf;
    else g_fusion_error = buf;
    return code;
}

#define FCU(f, call)                                                                                              \
    do {                                                                                                          \
        cudaError_t e__ = (call);                                                                                 \
        if (e__ != cudaSuccess)                                                                                   \
            return ffail(f, ACMMP_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

FusionCam make_cam(const acmmp_camera &c) {
    FusionCam o;
    memcpy(o.K, c.K, sizeof o.K);
    memcpy(o.R, c.R, sizeof o.R);
    memcpy(o.t, c.t, sizeof o.t);
    o.width = c.width;
    o.height = c.height;
    o.depth_max = c.depth_max;
    return o;
}

template <class T> cudaError_t dalloc(T **p, size_t count) { return cudaMalloc(reinterpret_cast<void **>(p), count * sizeof(T)); }

void free_scratch(acmmp_fusion *f) {
    cudaFree(f->loc); cudaFree(f->w); cudaFree(f->valid); cudaFree(f->appr); cudaFree(f->cons);
    cudaFree(f->agg); cudaFree(f->carry); cudaFree(f->cnt); cudaFree(f->offs);
    f->loc = nullptr; f->w = nullptr; f->valid = f->appr = nullptr; f->cons = nullptr;
    f->agg = f->carry = f->cnt = f->offs = nullptr;
}

}  // namespace

extern "C" {

int acmmp_fusion_create(acmmp_fusion **out, int device, int n_views, const acmmp_camera *cams, const int32_t *pair_offsets,
                        const int32_t *pair_ids) {
    if (!out || !cams || !pair_offsets || !pair_ids || n_views < 1)
        return ffail(nullptr, ACMMP_ERR_INVALID, "acmmp_fusion_create: bad arguments");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return ffail(nullptr, ACMMP_ERR_NO_DEVICE, "no CUDA device");
    if (device < 0 || device >= count) return ffail(nullptr, ACMMP_ERR_INVALID, "device %d out of range", device);
    for (int v = 0; v < n_views; ++v) {
        const int ns = pair_offsets[v + 1] - pair_offsets[v];
        if (ns < 0 || ns > ACMMP_MAX_SRC_VIEWS) return ffail(nullptr, ACMMP_ERR_INVALID, "view %d has %d source views (max %d)", v, ns, ACMMP_MAX_SRC_VIEWS);
        for (int k = pair_offsets[v]; k < pair_offsets[v + 1]; ++k)
            if (pair_ids[k] < 0 || pair_ids[k] >= n_views || pair_ids[k] == v)
                return ffail(nullptr, ACMMP_ERR_INVALID, "view %d: bad source id %d", v, pair_ids[k]);
        if (cams[v].width <= 0 || cams[v].height <= 0) return ffail(nullptr, ACMMP_ERR_INVALID, "view %d: bad size", v);
    }
    acmmp_fusion *f = new acmmp_fusion();
    f->device = device;
    f->n = n_views;
    f->cams.assign(cams, cams + n_views);
    f->pair_off.assign(pair_offsets, pair_offsets + n_views + 1);
    f->pair_ids.assign(pair_ids, pair_ids + pair_offsets[n_views]);
    f->views.resize(n_views);
    f->per_view.assign(n_views, 0);
    f->rounds.assign(n_views, 0);
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMalloc(reinterpret_cast<void **>(&f->changed), sizeof(int)) != cudaSuccess) {
        delete f;
        return ffail(nullptr, ACMMP_ERR_CUDA, "acmmp_fusion_create: cannot set up device %d", device);
    }
    *out = f;
    return 0;
}

void acmmp_fusion_destroy(acmmp_fusion *f) {
    if (!f) return;
    cudaSetDevice(f->device);
    if (f->stream) cudaStreamSynchronize(f->stream);
    for (auto &v : f->views) {
        cudaFree(v.depth); cudaFree(v.normal); cudaFree(v.color); cudaFree(v.mask0); cudaFree(v.mask); cudaFree(v.mark);
        cudaFree(v.depth2); cudaFree(v.normal2);
    }
    free_scratch(f);
    cudaFree(f->changed);
    cudaFree(f->points);
    if (f->stream) cudaStreamDestroy(f->stream);
    delete f;
}

const char *acmmp_fusion_last_error(const acmmp_fusion *f) { return f ? f->error.c_str() : g_fusion_error.c_str(); }

int acmmp_fusion_set_view(acmmp_fusion *f, int view, const float *depth, const float *normal3, const uint8_t *color3,
                          const uint8_t *mask, int on_device) {
    if (!f || !depth || !normal3) return ACMMP_ERR_INVALID;
    if (view < 0 || view >= f->n) return ffail(f, ACMMP_ERR_INVALID, "view %d out of range", view);
    FCU(f, cudaSetDevice(f->device));
    (void)cudaGetLastError();   // drop a stale non-sticky error left by other code in the process
    const size_t P = (size_t)f->cams[view].width * f->cams[view].height;
    FusionView &v = f->views[view];
    const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    if (!v.depth) {
        FCU(f, dalloc(&v.depth, P));
        FCU(f, dalloc(&v.normal, 3 * P));
        FCU(f, dalloc(&v.mask0, P));
        FCU(f, dalloc(&v.mask, P));
        FCU(f, dalloc(&v.mark, P));
    }
    FCU(f, cudaMemcpyAsync(v.depth, depth, P * sizeof(float), kind, f->stream));
    FCU(f, cudaMemcpyAsync(v.normal, normal3, 3 * P * sizeof(float), kind, f->stream));
    if (color3) {
        if (!v.color) FCU(f, dalloc(&v.color, 3 * P));
        FCU(f, cudaMemcpyAsync(v.color, color3, 3 * P, kind, f->stream));
    } else if (v.color) {
        cudaFree(v.color);
        v.color = nullptr;
    }
    if (mask) FCU(f, cudaMemcpyAsync(v.mask0, mask, P, kind, f->stream));
    else FCU(f, cudaMemsetAsync(v.mask0, 0, P, f->stream));
    if (!on_device) FCU(f, cudaStreamSynchronize(f->stream));   // the host buffers may go away
    v.set = true;
    return 0;
}

// both fusions: per view score -> fixed-point greedy pass -> ordered compaction
static int fusion_run_impl(acmmp_fusion *f, float consistency_scalar, int con_num_thresh, bool prior, int single_match_penalty,
                           int64_t *num_points) {
    if (!f) return ACMMP_ERR_INVALID;
    const int hyps = prior ? 2 : 1;
    FCU(f, cudaSetDevice(f->device));
    (void)cudaGetLastError();   // drop a stale non-sticky error left by other code in the process
    size_t maxP = 0;
    int max_src = 1;
    for (int v = 0; v < f->n; ++v) {
        if (!f->views[v].set) return ffail(f, ACMMP_ERR_STATE, "maps of view %d not set", v);
        if (prior && !f->views[v].set2) return ffail(f, ACMMP_ERR_STATE, "second maps of view %d not set", v);
        const size_t P = (size_t)f->cams[v].width * f->cams[v].height;
        if (P > maxP) maxP = P;
        const int ns = f->pair_off[v + 1] - f->pair_off[v];
        if (ns > max_src) max_src = ns;
    }
    const int max_chunks = (int)((maxP + kChunk - 1) / kChunk);
    if (maxP > f->cap_pixels || max_src > f->max_src || hyps > f->cap_hyps) {
        FCU(f, cudaStreamSynchronize(f->stream));
        free_scratch(f);
        FCU(f, dalloc(&f->loc, maxP * max_src * hyps));
        FCU(f, dalloc(&f->w, maxP * max_src * hyps));
        FCU(f, dalloc(&f->valid, maxP));
        FCU(f, dalloc(&f->appr, maxP));
        FCU(f, dalloc(&f->cons, maxP * hyps));
        FCU(f, dalloc(&f->agg, (size_t)(max_src + 1) * max_chunks));
        FCU(f, dalloc(&f->carry, (size_t)(max_src + 1) * max_chunks));
        FCU(f, dalloc(&f->cnt, (size_t)max_chunks));
        FCU(f, dalloc(&f->offs, (size_t)max_chunks + 1));
        f->cap_pixels = maxP;
        f->max_src = max_src;
        f->cap_hyps = hyps;
    }
    cudaStream_t st = f->stream;
    for (int v = 0; v < f->n; ++v) {
        const size_t P = (size_t)f->cams[v].width * f->cams[v].height;
        FCU(f, cudaMemcpyAsync(f->views[v].mask, f->views[v].mask0, P, cudaMemcpyDeviceToDevice, st));
    }
    f->points_n = 0;
    f->score_ms = f->greedy_ms = f->emit_ms = 0.0;
    cudaEvent_t ev[4];
    for (auto &e : ev) FCU(f, cudaEventCreate(&e));
    int rc = 0;
    for (int i = 0; i < f->n && rc == 0; ++i) {
        const int ns = f->pair_off[i + 1] - f->pair_off[i];
        const int *ngb = f->pair_ids.data() + f->pair_off[i];
        const int P = f->cams[i].width * f->cams[i].height;
        const int nchunks = (P + kChunk - 1) / kChunk;
        const int pblocks = (P + 255) / 256;
        FusionView &V = f->views[i];
        ScoreArgs sa;
        memset(&sa, 0, sizeof sa);
        IterArgs ia;
        memset(&ia, 0, sizeof ia);
        sa.ref = make_cam(f->cams[i]);
        sa.nsrc = ns;
        ia.nsrc = ns;
        ia.P = P;
        ia.consistency_scalar = consistency_scalar;
        ia.con_num_thresh = con_num_thresh;
        for (int j = 0; j < ns; ++j) {
            sa.src[j] = make_cam(f->cams[ngb[j]]);
            sa.src_depth[j] = f->views[ngb[j]].depth;
            sa.src_normal[j] = f->views[ngb[j]].normal;
            ia.mark[j] = f->views[ngb[j]].mark;
        }
        auto reset_marks = [&]() -> cudaError_t {
            for (int j = 0; j < ns; ++j) {
                const int SP = f->cams[ngb[j]].width * f->cams[ngb[j]].height;
                fusion_reset_marks_kernel<<<(SP + 255) / 256, 256, 0, st>>>(f->views[ngb[j]].mask, f->views[ngb[j]].mark, SP);
            }
            return cudaGetLastError();
        };
        FCU(f, cudaEventRecord(ev[0], st));
        if (prior) {
            PriorScoreArgs pa;
            memset(&pa, 0, sizeof pa);
            pa.ref = sa.ref;
            pa.nsrc = ns;
            for (int j = 0; j < ns; ++j) {
                pa.src[j] = sa.src[j];
                pa.src_depth[0][j] = f->views[ngb[j]].depth;
                pa.src_normal[0][j] = f->views[ngb[j]].normal;
                pa.src_depth[1][j] = f->views[ngb[j]].depth2;
                pa.src_normal[1][j] = f->views[ngb[j]].normal2;
            }
            prior_score_kernel<<<pblocks, 256, 0, st>>>(pa, V.depth, V.normal, V.depth2, V.normal2, V.mask, f->loc, f->w, f->valid);
        } else {
            fusion_score_kernel<<<pblocks, 256, 0, st>>>(sa, V.depth, V.normal, V.mask, f->loc, f->w, f->valid);
        }
        FCU(f, cudaGetLastError());
        FCU(f, cudaEventRecord(ev[1], st));
        FCU(f, cudaMemsetAsync(f->cons, 0, (size_t)P * hyps * sizeof(uint32_t), st));
        FCU(f, cudaMemsetAsync(f->appr, 0, (size_t)P, st));
        FCU(f, reset_marks());
        int rounds = 0;
        for (;;) {
            int changed = 0;
            FCU(f, cudaMemsetAsync(f->changed, 0, sizeof(int), st));
            if (prior)
                prior_decide_kernel<<<pblocks, 256, 0, st>>>(ia, single_match_penalty, f->loc, f->w, f->valid, f->cons, f->appr, f->changed);
            else
                fusion_decide_kernel<<<pblocks, 256, 0, st>>>(ia, f->loc, f->w, f->valid, f->cons, f->appr, f->changed);
            FCU(f, cudaGetLastError());
            FCU(f, cudaMemcpyAsync(&changed, f->changed, sizeof(int), cudaMemcpyDeviceToHost, st));
            FCU(f, cudaStreamSynchronize(st));
            ++rounds;
            if (!changed) break;
            if (rounds > P) {   // cannot happen: pixel p is final after round p + 1
                rc = ffail(f, ACMMP_ERR_STATE, "fusion of view %d did not converge", i);
                break;
            }
            FCU(f, reset_marks());
            if (prior) {
                prior_mark_kernel<<<pblocks, 256, 0, st>>>(ia, f->loc, f->cons, f->appr);
            } else {
                fusion_chunk_min_kernel<<<nchunks, kScanThreads, 0, st>>>(f->cons, f->appr, P, ns + 1, nchunks, f->agg);
                fusion_agg_scan_kernel<<<ns + 1, 32, 0, st>>>(f->agg, nchunks, f->carry);
                fusion_mark_kernel<<<nchunks, kScanThreads, 0, st>>>(ia, f->cons, f->appr, f->loc, f->carry, nchunks);
            }
            FCU(f, cudaGetLastError());
        }
        if (rc) break;
        f->rounds[i] = rounds;
        for (int j = 0; j < ns; ++j) {
            const int SP = f->cams[ngb[j]].width * f->cams[ngb[j]].height;
            fusion_commit_marks_kernel<<<(SP + 255) / 256, 256, 0, st>>>(f->views[ngb[j]].mask, f->views[ngb[j]].mark, SP);
        }
        FCU(f, cudaGetLastError());
        FCU(f, cudaEventRecord(ev[2], st));
        fusion_count_kernel<<<nchunks, kScanThreads, 0, st>>>(f->appr, P, f->cnt);
        fusion_offsets_kernel<<<1, 32, 0, st>>>(f->cnt, nchunks, f->offs);
        FCU(f, cudaGetLastError());
        int total = 0;
        FCU(f, cudaMemcpyAsync(&total, f->offs + nchunks, sizeof(int), cudaMemcpyDeviceToHost, st));
        FCU(f, cudaStreamSynchronize(st));
        f->per_view[i] = total;
        if (f->points_n + (size_t)total > f->points_cap) {
            size_t cap = f->points_cap ? f->points_cap : (size_t)1 << 20;
            while (cap < f->points_n + (size_t)total) cap *= 2;
            float *np_ = nullptr;
            FCU(f, dalloc(&np_, cap * 9));
            if (f->points_n) FCU(f, cudaMemcpyAsync(np_, f->points, f->points_n * 9 * sizeof(float), cudaMemcpyDeviceToDevice, st));
            FCU(f, cudaStreamSynchronize(st));
            cudaFree(f->points);
            f->points = np_;
            f->points_cap = cap;
        }
        if (total > 0) {
            fusion_emit_kernel<<<nchunks, kScanThreads, 0, st>>>(sa.ref, V.depth, V.normal, prior ? V.depth2 : nullptr,
                                                                 prior ? V.normal2 : nullptr, V.color, f->appr,
                                                                 f->offs, f->points + f->points_n * 9);
            FCU(f, cudaGetLastError());
        }
        f->points_n += (size_t)total;
        FCU(f, cudaEventRecord(ev[3], st));
        FCU(f, cudaEventSynchronize(ev[3]));
        float a = 0, b = 0, c = 0;
        cudaEventElapsedTime(&a, ev[0], ev[1]);
        cudaEventElapsedTime(&b, ev[1], ev[2]);
        cudaEventElapsedTime(&c, ev[2], ev[3]);
        f->score_ms += a;
        f->greedy_ms += b;
        f->emit_ms += c;
    }
    for (auto &e : ev) cudaEventDestroy(e);
    if (rc) return rc;
    if (num_points) *num_points = (int64_t)f->points_n;
    return 0;
}

int acmmp_fusion_run(acmmp_fusion *f, float consistency_scalar, int con_num_thresh, int64_t *num_points) {
    return fusion_run_impl(f, consistency_scalar, con_num_thresh, false, 0, num_points);
}

int acmmp_fusion_run_prior_aware(acmmp_fusion *f, float consistency_scalar, int num_consistent_thresh, int single_match_penalty,
                                 int64_t *num_points) {
    return fusion_run_impl(f, consistency_scalar, num_consistent_thresh, true, single_match_penalty, num_points);
}

int acmmp_fusion_set_view_second(acmmp_fusion *f, int view, const float *depth2, const float *normal3_2, int on_device) {
    if (!f || !depth2 || !normal3_2) return ACMMP_ERR_INVALID;
    if (view < 0 || view >= f->n) return ffail(f, ACMMP_ERR_INVALID, "view %d out of range", view);
    FCU(f, cudaSetDevice(f->device));
    (void)cudaGetLastError();   // drop a stale non-sticky error left by other code in the process
    const size_t P = (size_t)f->cams[view].width * f->cams[view].height;
    FusionView &v = f->views[view];
    const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    if (!v.depth2) {
        FCU(f, dalloc(&v.depth2, P));
        FCU(f, dalloc(&v.normal2, 3 * P));
    }
    FCU(f, cudaMemcpyAsync(v.depth2, depth2, P * sizeof(float), kind, f->stream));
    FCU(f, cudaMemcpyAsync(v.normal2, normal3_2, 3 * P * sizeof(float), kind, f->stream));
    if (!on_device) FCU(f, cudaStreamSynchronize(f->stream));
    v.set2 = true;
    return 0;
}

int acmmp_fusion_get_points(acmmp_fusion *f, float *points9, int64_t capacity, int64_t *per_view_counts) {
    if (!f) return ACMMP_ERR_INVALID;
    FCU(f, cudaSetDevice(f->device));
    (void)cudaGetLastError();   // drop a stale non-sticky error left by other code in the process
    if (points9) {
        if (capacity < (int64_t)f->points_n) return ffail(f, ACMMP_ERR_INVALID, "capacity %lld < %zu points", (long long)capacity, f->points_n);
        if (f->points_n) FCU(f, cudaMemcpyAsync(points9, f->points, f->points_n * 9 * sizeof(float), cudaMemcpyDeviceToHost, f->stream));
        FCU(f, cudaStreamSynchronize(f->stream));
    }
    if (per_view_counts)
        for (int v = 0; v < f->n; ++v) per_view_counts[v] = f->per_view[v];
    return 0;
}

int acmmp_fusion_get_mask(acmmp_fusion *f, int view, uint8_t *mask) {
    if (!f || !mask || view < 0 || view >= f->n || !f->views[view].set) return ACMMP_ERR_INVALID;
    FCU(f, cudaSetDevice(f->device));
    (void)cudaGetLastError();   // drop a stale non-sticky error left by other code in the process
    const size_t P = (size_t)f->cams[view].width * f->cams[view].height;
    FCU(f, cudaMemcpyAsync(mask, f->views[view].mask, P, cudaMemcpyDeviceToHost, f->stream));
    FCU(f, cudaStreamSynchronize(f->stream));
    return 0;
}

int acmmp_fusion_stats(acmmp_fusion *f, double ms[3], int32_t *rounds_per_view) {
    if (!f) return ACMMP_ERR_INVALID;
    if (ms) { ms[0] = f->score_ms; ms[1] = f->greedy_ms; ms[2] = f->emit_ms; }
    if (rounds_per_view)
        for (int v = 0; v < f->n; ++v) rounds_per_view[v] = f->rounds[v];
    return 0;
}

}  // extern "C"
