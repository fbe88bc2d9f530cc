// sweep_kernel.cuh - one red or black half-sweep of adaptive-checkerboard
// propagation + multi-hypothesis joint view selection + plane refinement.
//
// What it computes is the reference's CheckerboardPropagation +
// PlaneHypothesisRefinement (ACMMP.cu:707-1173) for every pixel of one colour.
// How it is organised is new:
//
//  * 8 lanes per pixel ("group"), 4 pixels per warp, 32 pixels per 256-thread
//    CTA.  Lane j owns neighbour arm j (0 up_near, 1 up_far, 2 down_near,
//    3 down_far, 4 left_near, 5 left_far, 6 right_near, 7 right_far): it scans
//    its candidate strip, fetches the winning plane and scores it against all
//    sources.  The 8 x (N-1) cost matrix lives in shared memory (the reference
//    keeps it, plus six more 32-float arrays, in 1,888 B of per-thread local
//    memory).  Votes and the arg-min / arg-max run on shuffles inside the
//    group, the multinomial view sampling on ballots (every lane replays the
//    pixel's random stream; lane l owns the CDF entries of views l, l + 8, ...).
//  * ONE inlined copy of the single-lane NCC loop serves the cost matrix
//    (pass 0) and the refinement candidates (pass 1); the arms' planes, the
//    pixel's plane, the candidates and the decision state are parked in shared
//    memory around it (SweepGroupSmem), which is what lets four CTAs of 64
//    registers share an SM.  Cold launches (right after a random
//    initialisation) score the cost matrix by lane pairs instead
//    (sweep_kernel<..., kPairMatrix>): a texture quad then holds two arm
//    hypotheses instead of four.
//  * the hypothesis-independent half of every NCC (36 bilateral weights and
//    the reference-patch moments) is computed once per pixel per launch
//    (RefPatch, ncc.cuh); camera-pair constants once per problem (PairConst).
//  * the current plane and the 5 refinement candidates are only scored against
//    the views that were actually sampled (view weight > 0) - the reference
//    scores all N-1 and multiplies the rest by zero - and those (candidate,
//    view) tasks are spread over the 8 lanes.
//  * the XORWOW state stays in registers from the first to the last draw of the
//    pixel (every lane of a group carries an identical copy, so all lanes make
//    the same draws).
//  * same-colour reads (the V-shaped near-arm candidates sit on the colour
//    being rewritten, a data race in the reference) see the state before the
//    launch: planes / costs are ping-pong buffers (SweepArgs: read one, write
//    the other, no copies), which is deterministic and what the reference does
//    whenever the neighbour has not finished yet.
#pragma once
#include <cuda.h>   // CUtensorMap (type only)

#include "common.cuh"
#include "geometry.cuh"
#include "ncc.cuh"

namespace acmmp {

constexpr int kGroup = 8;                 // lanes per pixel
#ifndef ACMMP_SWEEP_GROUPS
#define ACMMP_SWEEP_GROUPS 32
#endif
constexpr int kGroupsPerCta = ACMMP_SWEEP_GROUPS;   // pixels per CTA (multiple of 8)
constexpr int kSweepThreads = kGroup * kGroupsPerCta;
constexpr int kTileW = 16, kTileH = kGroupsPerCta / 8;   // 16 x kTileH pixels, half of each colour
constexpr int kInitGroup = 4;             // init_kernel: lanes per pixel
constexpr int kInitTileW = 16, kInitTileH = 4;
constexpr int kInitGroups = kInitTileW * kInitTileH;   // pixels per CTA
constexpr int kInitThreads = kInitGroup * kInitGroups;

struct SweepArgs {
    // State before the half-sweep (read) and after it (write): a ping-pong pair X / Y without copies.  The black and
    // the red half-sweep of an iteration form a pair: the first reads everything from X and writes its colour to Y;
    // the second reads the OTHER colour (odd Manhattan distance: direct neighbours, far arms) from Y - the values the
    // first just wrote - and its own colour (the V candidates, its own plane) from X, and writes to Y, which then
    // holds the whole state.  Same-colour reads therefore always see the state before the launch.
    // Diagnostics (acmmp_run_sweep_snapshot): all neighbour reads from an injected snapshot, own plane from the live
    // state, written in place.
    const float4 *planes_other;  // planes / costs of the other colour
    const float *costs_other;
    const float4 *planes_in;     // planes / costs of the same colour before this half-sweep
    const float *costs_in;
    const float4 *planes_self;   // the pixel's own plane
    const float *costs_self;
    float4 *planes;
    float *costs;
    uint32_t *selected_views;
    const float *pre_costs;
    const float4 *prior_planes;
    const uint32_t *plane_masks;
    RngPlanes rng;
    const float *ref_image;   // row-major reference image
    cudaTextureObject_t tex;  // layered texture of all views (uniform handle)
    const PairConst *pairs;   // nsrc
    const RefConst *rc;
    int nsrc;
    int iter;
    int color;                // 0 black ((x+y) even), 1 red
    float sigma_spatial, sigma_color;
    int geom_consistency, planar_prior, hierarchy;
    int compat;               // acmmp_compat_mode
    // Emulation of the reference launch's block waves (acmmp_set_wave_emulation; 0 = off).  The reference
    // covers the image with 32 x 32-pixel blocks (32 x 16 threads, one colour) of which `resident` run at a
    // time (148 on a B200: one 512-thread block of 128 registers per SM); blocks of a later wave start when
    // those of the earlier wave have finished and stored their results, so their same-colour reads of earlier
    // blocks see NEW planes / costs.  This sub-launch handles the reference blocks [wave_lo, wave_hi) and reads
    // same-colour neighbours in blocks < wave_lo from the output buffers.
    int wave_lo, wave_hi, ref_grid_x;
    unsigned long long *ncc_counter;   // optional instrumentation: kNccCounterSlots slots, NCC evaluations issued
    // sweep_kernel<true, true, ...> only: the R8 layers as a pitch-linear (x, y, layer) byte tensor for TMA (box staging, ncc.cuh)
    CUtensorMap views_map;
    int layer_w, layer_h;
};
constexpr int kNccCounterSlots = 4096;

// mbarrier / TMA primitives of the box staging
ACMMP_DEV uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
ACMMP_DEV void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
ACMMP_DEV void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
ACMMP_DEV void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_addr(bar)),
        "r"(parity)
        : "memory");
}
ACMMP_DEV void tma_load_box(void *dst, const CUtensorMap *map, int x, int y, int layer, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_addr(dst)),
                 "l"(map), "r"(x), "r"(y), "r"(layer), "r"(smem_addr(bar))
                 : "memory");
}

// Per-pixel shared-memory block.  nsrc-sized arrays follow the fixed part.
struct GroupSmemFixed {
    RefPatch rp;
};
// sweep_kernel: the five refinement candidates of the pixel live here between their generation and the decision, so
// that they do not occupy 25 registers across the NCC loop (which both passes share)
struct SweepGroupSmem {
    RefPatch rp;
    // operands of the shared NCC loop: planes[lane] = the candidate of checkerboard arm `lane` (pass 0),
    // planes[8 + i] = refinement candidate i (pass 1)
    float4 planes[13];
    float4 plane_c;                          // the pixel's own plane before the half-sweep
    float4 plane_now, live_plane, prior;     // parked across the second pass (lane 0 writes, every lane reads back)
    float cand_depth[8];
    float scal[8];                           // parked scalars: cost_now, live_cost, depth_now, restricted_cost, depth_prior
};

ACMMP_DEV size_t group_smem_floats(int nsrc) {
    // cm[nsrc][8] + task costs [6][nsrc] + probs[nsrc]
    return (size_t)nsrc * 8 + (size_t)6 * nsrc + nsrc;
}

__host__ __device__ inline size_t sweep_smem_bytes(int nsrc, bool boxes = false) {
    size_t per_group = sizeof(SweepGroupSmem) + ((size_t)nsrc * 8 + (size_t)6 * nsrc + nsrc) * sizeof(float);
    per_group = (per_group + 15) & ~(size_t)15;
    size_t pairs = ((size_t)nsrc * sizeof(PairConst) + 15) & ~(size_t)15;
    size_t rc = (sizeof(RefConst) + 15) & ~(size_t)15;
    size_t total = pairs + rc + per_group * kGroupsPerCta;
    if (boxes) {   // [mbarrier 16 B][box origins nsrc x 8 B] -> 128-byte boundary -> nsrc boxes
        total += 16 + (size_t)nsrc * 8;
        total = (total + 127) & ~(size_t)127;
        total += (size_t)nsrc * kBoxW * kBoxH;
    }
    return total;
}

// group-scoped shuffle helpers
struct Grp {
    unsigned mask;   // lanes of this group inside the warp
    int base;        // first lane of the group
    int lane;        // 0..7
    ACMMP_DEV float shfl(float v, int src) const { return __shfl_sync(mask, v, base + src); }
    ACMMP_DEV int shfl(int v, int src) const { return __shfl_sync(mask, v, base + src); }
    ACMMP_DEV unsigned shfl(unsigned v, int src) const { return __shfl_sync(mask, v, base + src); }
    ACMMP_DEV void sync() const { __syncwarp(mask); }
    ACMMP_DEV unsigned ballot(bool p) const { return (__ballot_sync(mask, p) >> base) & 0xffu; }
};

}  // namespace acmmp
