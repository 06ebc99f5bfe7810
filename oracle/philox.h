/* oracle/philox.h — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
 *
 * Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy
 * as 1, 2, 3", SC'11), restated from the paper.  Known-answer vectors of the
 * Random123 distribution are checked in tests/test_oracle_primitives.py.
 *
 * Draw contract: for burst-sample (global burst g, repeat r),
 * dwell step s and slot q the four words are
 *     philox4x32_10(ctr = (g, r, s, q), key = (seed, epoch)).
 */
#ifndef ORACLE_PHILOX_H
#define ORACLE_PHILOX_H
#include <stdint.h>

typedef struct { uint32_t v[4]; } philox4_t;

static inline philox4_t philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                      uint32_t k0, uint32_t k1)
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    const uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int round = 0; round < 10; ++round) {
        uint64_t p0 = (uint64_t)M0 * c0;
        uint64_t p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    philox4_t r = {{c0, c1, c2, c3}};
    return r;
}

/* slot numbers (counter word c3) of the draw contract (v2: one block per dwell step, plus one
 * per step when the background path is on; no per-sample block) */
enum {
    PDA_SLOT_STEP = 0u, /* x0: next-state uniform, x1: dwell uniform, x2: normal word;
                           step 0 only: x3 seeds the xor128 binomial stream (rk_seed(x3)), and the
                           initial-state uniform is built from the low bytes of x0, x1, x2, which
                           the uniforms above (top 24 bits of a word) do not use */
    PDA_SLOT_BG = 1u    /* x0: high word of the acceptor-background Poisson draw (low word 0),
                           x1: its fallback uniform, x2 / x3: the same for the donor background */
};

/* the initial-state word of step 0 */
static inline uint32_t pda_init_word(const philox4_t *w)
{
    return ((w->v[0] & 0xffu) << 24) | ((w->v[1] & 0xffu) << 16) | ((w->v[2] & 0xffu) << 8) | 0x80u;
}

#endif
