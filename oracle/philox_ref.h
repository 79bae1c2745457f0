/* TEST INFRASTRUCTURE ONLY — part of the CPU oracle, never linked into the product.
 *
 * Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11),
 * restated from the published algorithm.  The reference (OpenSpiel) has no counter-based
 * RNG: its evaluators draw from std::mt19937 / numpy RandomState
 * (open_spiel/algorithms/mcts.h:105, open_spiel/python/algorithms/mcts.py:56) and no
 * reference test pins those streams.  The replay contract of this repo therefore is
 *     r(seed, playout, t) = philox4x32_10(ctr={playout_lo, playout_hi, t>>2, 0},
 *                                         key={seed_lo, seed_hi})[t & 3]
 *     action_t = LegalActions()[ r mod n ]
 * and both the oracle and the CUDA engine compute r from this definition independently.
 */
#ifndef BSG_ORACLE_PHILOX_REF_H_
#define BSG_ORACLE_PHILOX_REF_H_
#include <stdint.h>

static inline void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2],
                                        uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int round = 0; round < 10; ++round) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* The t-th draw of playout `playout` under `seed`. */
static inline uint32_t oracle_draw(uint64_t seed, uint64_t playout, uint32_t t) {
  uint32_t ctr[4] = {(uint32_t)playout, (uint32_t)(playout >> 32), t >> 2, 0u};
  uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  uint32_t out[4];
  oracle_philox4x32_10(ctr, key, out);
  return out[t & 3];
}
#endif
