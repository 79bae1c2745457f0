// Shared host/device primitives of the batched playout engine (sm_100a).
// Everything here is integer work on registers; no tensor cores by design.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define BSG_HD __host__ __device__ __forceinline__
#define BSG_D __device__ __forceinline__
#define BSG_HDC __host__ __device__ constexpr
#else
#define BSG_HD inline
#define BSG_D inline
#define BSG_HDC constexpr
#endif

namespace bsg {

typedef uint32_t u32;
typedef uint64_t u64;

BSG_HD int popc32(u32 x) {
#if defined(__CUDA_ARCH__)
  return __popc(x);
#else
  return __builtin_popcount(x);
#endif
}
BSG_HD int popc64(u64 x) {
#if defined(__CUDA_ARCH__)
  return __popcll(x);
#else
  return __builtin_popcountll(x);
#endif
}
BSG_HD int ctz64(u64 x) {  // index of the lowest set bit; 64 for x = 0
#if defined(__CUDA_ARCH__)
  return x ? __ffsll((long long)x) - 1 : 64;
#else
  return x ? __builtin_ctzll(x) : 64;
#endif
}
BSG_HD u32 umulhi32(u32 a, u32 b) {
#if defined(__CUDA_ARCH__)
  return __umulhi(a, b);
#else
  return (u32)(((u64)a * b) >> 32);
#endif
}
BSG_HD u32 umin32(u32 a, u32 b) { return a < b ? a : b; }

// Funnel shifts over a (hi:lo) pair, shift in [0,31].
BSG_HD u32 fshl(u32 lo, u32 hi, int s) {  // high word of (hi:lo) << s
#if defined(__CUDA_ARCH__)
  return __funnelshift_l(lo, hi, s);
#else
  return s ? (hi << s) | (lo >> (32 - s)) : hi;
#endif
}
BSG_HD u32 fshr(u32 lo, u32 hi, int s) {  // low word of (hi:lo) >> s
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(lo, hi, s);
#else
  return s ? (lo >> s) | (hi << (32 - s)) : lo;
#endif
}

// ---------------------------------------------------------------------------------------
// Philox4x32-10, counter based.  Replay contract of this engine (see include/bsg.h):
//   r(seed, playout, t) = philox4x32_10(ctr = {playout_lo, playout_hi, t >> 2, 0},
//                                       key = {seed_lo, seed_hi})[t & 3]
// The key schedule depends only on the seed, so a kernel computes it once per launch.
// ---------------------------------------------------------------------------------------
struct PhiloxKeys {
  u32 k0[10];
  u32 k1[10];
};
BSG_HD void philox_schedule(u64 seed, PhiloxKeys& ks) {
  u32 a = (u32)seed, b = (u32)(seed >> 32);
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    ks.k0[i] = a;
    ks.k1[i] = b;
    a += 0x9E3779B9u;
    b += 0xBB67AE85u;
  }
}
BSG_HD void philox4x32_10(u32 c0, u32 c1, u32 c2, u32 c3, const PhiloxKeys& ks, u32 out[4]) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    u32 h0 = umulhi32(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
    u32 h1 = umulhi32(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
    u32 n0 = h1 ^ c1 ^ ks.k0[i];
    u32 n2 = h0 ^ c3 ^ ks.k1[i];
    c0 = n0; c1 = l1; c2 = n2; c3 = l0;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// ---------------------------------------------------------------------------------------
// r mod n for 1 <= n <= 127 with one table word: M[n] = floor(2^32 / n) (M[1] = 2^32-1).
// q' = mulhi(r, M) is floor(r/n) or one less, so r - q'n is in [0, 2n) and one unsigned
// min folds it back.  n = 0 (a finished lane running dead steps) returns garbage, never traps.
// ---------------------------------------------------------------------------------------
BSG_HD u32 modM_entry(u32 n) {
  if (n == 0) return 0u;
  if (n == 1) return 0xFFFFFFFFu;
  return (u32)(0x100000000ull / n);
}
BSG_HD u32 fastmod(u32 r, u32 n, u32 M) {
  u32 rem = r - umulhi32(r, M) * n;
  return umin32(rem, rem - n);
}

// Index of the k-th (0-based) set bit of x; k must be < popc(x) for a meaningful answer.
BSG_HD int select32(u32 x, int k) {
  int pos = 0, c;
  c = popc32(x & 0xFFFFu); if (k >= c) { k -= c; pos += 16; x >>= 16; }
  c = popc32(x & 0xFFu);   if (k >= c) { k -= c; pos += 8;  x >>= 8; }
  c = popc32(x & 0xFu);    if (k >= c) { k -= c; pos += 4;  x >>= 4; }
  c = popc32(x & 0x3u);    if (k >= c) { k -= c; pos += 2;  x >>= 2; }
  c = (int)(x & 1u);       if (k >= c) { pos += 1; }
  return pos;
}

// Outcome codes shared by all games (packed-state meta and kernel results).
enum : int { kOutcomeP0 = 0, kOutcomeP1 = 1, kOutcomeDraw = 2, kOutcomeNone = 3 };
BSG_HD int outcome_ret0(int outcome) { return outcome == kOutcomeP0 ? 1 : (outcome == kOutcomeP1 ? -1 : 0); }

enum : int { kGameTicTacToe = 0, kGameConnectFour = 1, kGameBreakthrough = 2, kGameHex = 3, kGamePentago = 5, kGameOthello = 6, kGameY = 7 };

}  // namespace bsg
