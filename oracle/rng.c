/*
 * rng.c - XORWOW generator, stream-compatible with the cuRAND device API calls
 * the reference makes: curand_init(seed, subsequence = p.y, offset = p.x, &st)
 * (ACMMP.cu:624) and curand_uniform(&st) (ACMMP.cu:177,202,237,648,724,1036).
 *
 * cuRAND 12.9 (toolkit header curand_kernel.h, not part of /root/reference)
 * defines the algorithm restated here: seed scrambling, one generator step,
 * "skip ahead n subsequences" = 2^67 * n steps and "skip ahead n" steps applied
 * to the five xorshift words (a linear map over GF(2)), the Weyl word d advanced
 * by 362437 per step.  cuRAND ships pre-computed skip matrices; this file
 * derives the same maps by repeated squaring of the one-step matrix instead.
 * Pinned by tests/test_oracle_rng.py against tests/golden/xorwow_golden.npz,
 * generated from cuRAND's own host tables by tests/golden/make_xorwow_golden.cu.
 *
 * Test infrastructure (see oracle.h).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "oracle.h"

#define NW 5            /* xorshift words */
#define NB (NW * 32)    /* state bits */

/* A linear map on the 160-bit state: row i = image of basis vector e_i. */
typedef struct { uint32_t row[NB][NW]; } gf2_map;

static void xorshift_step(uint32_t v[NW]) {
    uint32_t t = v[0] ^ (v[0] >> 2);
    v[0] = v[1]; v[1] = v[2]; v[2] = v[3]; v[3] = v[4];
    v[4] = (v[4] ^ (v[4] << 4)) ^ (t ^ (t << 1));
}

static void map_apply(const gf2_map *m, const uint32_t in[NW], uint32_t out[NW]) {
    uint32_t acc[NW] = {0, 0, 0, 0, 0};
    for (int i = 0; i < NB; ++i)
        if ((in[i >> 5] >> (i & 31)) & 1u)
            for (int k = 0; k < NW; ++k) acc[k] ^= m->row[i][k];
    memcpy(out, acc, sizeof acc);
}

/* out = b after a (apply a first) */
static void map_compose(const gf2_map *a, const gf2_map *b, gf2_map *out) {
    gf2_map *tmp = (gf2_map *)malloc(sizeof(gf2_map));
    for (int i = 0; i < NB; ++i) map_apply(b, a->row[i], tmp->row[i]);
    memcpy(out, tmp, sizeof(gf2_map));
    free(tmp);
}

static gf2_map *g_step_pow2 = NULL;   /* [k] = one step, applied 2^k times, k < 64+67+... */
#define NPOW 140

static void build_tables(void) {
    if (g_step_pow2) return;
    gf2_map *t = (gf2_map *)malloc(sizeof(gf2_map) * NPOW);
    for (int i = 0; i < NB; ++i) {
        uint32_t v[NW] = {0, 0, 0, 0, 0};
        v[i >> 5] = 1u << (i & 31);
        xorshift_step(v);
        memcpy(t[0].row[i], v, sizeof v);
    }
    for (int k = 1; k < NPOW; ++k) map_compose(&t[k - 1], &t[k - 1], &t[k]);
    g_step_pow2 = t;
}

/* advance v by n * 2^shift steps */
static void skip(uint32_t v[NW], uint64_t n, int shift) {
    build_tables();
    for (int b = 0; n; ++b, n >>= 1)
        if (n & 1u) map_apply(&g_step_pow2[b + shift], v, v);
}

void orc_rng_seed(uint64_t seed, uint64_t subsequence, uint64_t offset, uint32_t st[6]) {
    /* seed scrambling of curand_init for XORWOW */
    uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
    uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
    uint32_t t0 = 1099087573u * s0;
    uint32_t t1 = 2591861531u * s1;
    uint32_t d = 6615241u + t1 + t0;
    uint32_t v[NW];
    v[0] = 123456789u + t0;
    v[1] = 362436069u ^ t0;
    v[2] = 521288629u + t1;
    v[3] = 88675123u ^ t1;
    v[4] = 5783321u + t0;
    skip(v, subsequence, 67);   /* one subsequence = 2^67 draws; d unchanged (multiple of 2^32) */
    skip(v, offset, 0);
    d += 362437u * (uint32_t)offset;
    st[0] = d;
    memcpy(st + 1, v, sizeof v);
}

uint32_t orc_rng_next(uint32_t st[6]) {
    xorshift_step(st + 1);
    st[0] += 362437u;
    return st[5] + st[0];
}

/* curand_uniform: x * 2^-32 + 2^-33, evaluated by the GPU as a single FMA */
float orc_rng_uniform(uint32_t st[6]) {
    return fmaf((float)orc_rng_next(st), 2.3283064e-10f, 2.3283064e-10f / 2.0f);
}

void orc_rng_init_image(uint64_t seed, int w, int h, uint32_t *rng6) {
    for (int y = 0; y < h; ++y) {
        uint32_t st[6];
        orc_rng_seed(seed, (uint64_t)y, 0, st);
        for (int x = 0; x < w; ++x) {
            memcpy(rng6 + ((size_t)y * w + x) * 6, st, sizeof st);
            orc_rng_next(st);
        }
    }
}
