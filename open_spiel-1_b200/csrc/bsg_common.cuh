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
BSG_HD int ctz32(u32 x) {  // index of the lowest set bit of a non-zero word
#if defined(__CUDA_ARCH__)
  return __ffs((int)x) - 1;
#else
  return __builtin_ctz(x);
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
// Multi-word shifts on the MULTIPLIER pipe (device only; an A/B switch, OFF by default because it
// measured slower — see BSG_MULSHIFT below).  The flood kernels (hex, y) are bound
// by the ALU pipe — LOP3 / SHF at 88-93 % with the multiplier pipe at 4-14 % (profiles/r1f, r1s)
// — and a third of a dilation ring is funnel shifts.  A shift is a multiplication by a power of
// two: (f[w] << k) | (f[w-1] >> (32-k)) = f[w] * 2^k + hi32(f[w-1] * 2^k), one IMAD.HI and one
// IMAD with the addend folded in, no ALU instruction.  The multipliers are read from a
// __constant__ table (host-writable, hence opaque to ptxas, which otherwise turns multiplications
// by visible powers of two back into SHF / LEA) and travel in uniform registers.
// ---------------------------------------------------------------------------------------
#ifndef BSG_MULSHIFT
#define BSG_MULSHIFT 0   // measured on B200: hex 71.8 -> 68.1, y 23.9 -> 23.5 G steps/s with it (IMAD.HI is quarter rate)
#endif
#if defined(__CUDACC__)
static __constant__ u32 g_pow2[33] = {
    1u << 0, 1u << 1, 1u << 2, 1u << 3, 1u << 4, 1u << 5, 1u << 6, 1u << 7, 1u << 8, 1u << 9, 1u << 10,
    1u << 11, 1u << 12, 1u << 13, 1u << 14, 1u << 15, 1u << 16, 1u << 17, 1u << 18, 1u << 19, 1u << 20,
    1u << 21, 1u << 22, 1u << 23, 1u << 24, 1u << 25, 1u << 26, 1u << 27, 1u << 28, 1u << 29, 1u << 30,
    1u << 31, 0u};
#endif
// o = f shifted towards HIGHER bit indices by K (0 < K < 32) over W words (bits leaving the top
// word are dropped).
template <int K, int W> BSG_HD void shl_words(const u32* f, u32* o) {
#if defined(__CUDA_ARCH__) && BSG_MULSHIFT
  const u32 m = g_pow2[K];
  u32 carry = 0;
#pragma unroll
  for (int w = 0; w < W; ++w) {
    u32 r, h = 0;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(f[w]), "r"(m), "r"(carry));
    if (w < W - 1) asm("mul.hi.u32 %0, %1, %2;" : "=r"(h) : "r"(f[w]), "r"(m));
    o[w] = r;
    carry = h;
  }
#else
#pragma unroll
  for (int w = W - 1; w >= 0; --w) o[w] = w == 0 ? f[0] << K : fshl(f[w - 1], f[w], K);
#endif
}
// o = f shifted towards LOWER bit indices by K (0 < K < 32) over W words.
template <int K, int W> BSG_HD void shr_words(const u32* f, u32* o) {
#if defined(__CUDA_ARCH__) && BSG_MULSHIFT
  const u32 m = g_pow2[32 - K];
#pragma unroll
  for (int w = 0; w < W; ++w) {
    u32 lo = 0, r;
    if (w < W - 1) asm("mul.lo.u32 %0, %1, %2;" : "=r"(lo) : "r"(f[w + 1]), "r"(m));
    asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(f[w]), "r"(m), "r"(lo));
    o[w] = r;
  }
#else
#pragma unroll
  for (int w = 0; w < W; ++w) o[w] = w == W - 1 ? f[w] >> K : fshr(f[w], f[w + 1], K);
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

enum : int { kGameTicTacToe = 0, kGameConnectFour = 1, kGameBreakthrough = 2, kGameHex = 3, kGamePentago = 5, kGameOthello = 6, kGameY = 7, kGameHavannah = 9, kGameGo = 11 };

}  // namespace bsg
