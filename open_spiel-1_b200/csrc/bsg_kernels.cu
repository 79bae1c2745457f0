// CUDA kernels + C ABI of the batched playout engine.  sm_100a only; integer pipes only.
//
// Playout kernel design:
//  * one playout per lane, whole game state in registers (per-game bitboards, bsg_games.cuh);
//  * persistent warps: every warp draws chunks of kChunkPlayouts consecutive playouts from one
//    global work counter (one atomicAdd per chunk) and its lanes claim the next playout of the
//    chunk from a warp-uniform cursor with one ballot when they finish (lane refill; no
//    per-playout atomics), so a warp keeps ~all lanes busy although game lengths differ;
//  * steps run in groups of four: one Philox4x32-10 block (4 draws) per group, and lanes are
//    refilled only at group boundaries, which keeps every lane's step index congruent to the
//    group phase — the RNG and the refill code stay warp-convergent;
//  * the refill code runs whenever a lane waits (measured and rejected: waiting until >= 4..16
//    lanes are idle; it was slower for all four games);
//  * a step is unconditional straight-line code (finished lanes run dead steps on a dead
//    state); only the bookkeeping is predicated;
//  * per-game launch traits (members of the game struct, trait templates below): group length,
//    resident blocks, rolled step group (long steps: instruction cache), L1 carve-out (games
//    whose state is local memory), lockstep replay of a root's move history (go).
#include <type_traits>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>   // types and enums only: the library is loaded with dlopen (bsg_comm_*)
#include <stdio.h>

#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "bsg_host.cuh"
#include "bsg_mcts.cuh"
#include "bsg_forest.cuh"
#include "bsg_actor.cuh"

namespace bsg {

#ifndef BSG_SMEM_DRAWS
#define BSG_SMEM_DRAWS 1    // measured on B200: hex 71.8 -> 73.9, y 23.9 -> 25.5 G steps/s
#endif
#ifndef BSG_STAGE_ROOTS
#define BSG_STAGE_ROOTS 0   // A/B switch: chunk roots staged in shared memory (lean, single-plane games); measured slower on B200: connect_four 264.1 -> 260.8 G steps/s (46 -> 56 registers)
#endif
#ifndef BSG_EXTRA_RINGS
#define BSG_EXTRA_RINGS 0   // A/B switch: up to this many warp-voted extra flood rings per iteration
#endif
#ifndef BSG_EXTRA_RINGS_T
#define BSG_EXTRA_RINGS_T 8
#endif
constexpr int kThreads = 256;
constexpr int kChunkPlayouts = 128;   // playouts per work-counter draw of a warp
constexpr int kWorkSlots = 1024;      // launches that may be in flight at once, per engine
constexpr int kMaxDevices = 64;       // per-device caches of function attributes
constexpr size_t kSmallCallBytes = 64 << 10;   // bsg_playouts_host: roots up to this size take the lean single-stream path

struct PlayoutArgs {
  const void* roots;
  int64_t n_roots;   // capacity of the packed SoA root buffer
  int64_t root0;     // first root of this launch (chunked host path); playout p uses root0 + p / rollouts
  uint32_t n_playouts;
  uint32_t rollouts;
  uint64_t first_playout;
  PhiloxKeys keys;
  bsg_playout_io io;
  uint32_t* work_counter;  // zeroed before the launch; warps draw chunks of playouts from it
  uint32_t out_flags;  // kOut* bits: which optional per-playout outputs of io are present
  uint32_t compact;    // != 0: `roots` is in the game's compact format (compact_bits<G> per root)
  uint32_t* fill_boards;  // value-only launches: kFillWords words per playout for fill_resolve_kernel
  uint64_t fill_base;     // host path: this launch parks its boards from playout slot fill_base on
};
enum : uint32_t { kOutRet = 1, kOutLength = 2, kOutTerminal = 4, kOutRootSums = 8 };
inline uint32_t out_flags_of(const bsg_playout_io& io) {
  return (io.ret0 ? kOutRet : 0u) | (io.length ? kOutLength : 0u) | (io.terminal ? kOutTerminal : 0u) |
         ((io.root_sum_ret0 || io.root_sum_steps) ? kOutRootSums : 0u);
}

// Resident CTAs per SM the lean kernel is compiled for (register cap); 1 = no cap.
template <class G, class = void> struct min_blocks : std::integral_constant<int, 1> {};
template <class G> struct min_blocks<G, std::enable_if_t<(G::kMinBlocks > 1)>> : std::integral_constant<int, G::kMinBlocks> {};

// Steps per group of the lean kernel (one Philox block per four steps; lanes refill only at group
// boundaries).  4 unless the game says otherwise (G::kGroupSteps, a multiple of 4).
template <class G, class = void> struct group_steps : std::integral_constant<int, 4> {};
template <class G> struct group_steps<G, std::enable_if_t<(G::kGroupSteps > 4)>> : std::integral_constant<int, G::kGroupSteps> {};

// Games whose per-lane state is mostly local memory (G::kLocalHeavy), see PlayoutOp.
template <class G, class = void> struct local_heavy : std::false_type {};
template <class G> struct local_heavy<G, std::enable_if_t<G::kLocalHeavy>> : std::true_type {};
#ifndef BSG_LOCAL_CARVEOUT
#define BSG_LOCAL_CARVEOUT 15
#endif

// Games whose step is too long to unroll the group (G::kRolledGroup), see the step loop.
template <class G, class = void> struct rolled_group : std::false_type {};
template <class G> struct rolled_group<G, std::enable_if_t<G::kRolledGroup>> : std::true_type {};

// Games whose full board is inert (G::kInertWhenFull): a step on a full board places nothing and
// cannot report a win, so the lean kernel need not test for the draw in every step — a playout
// that has made as many moves as its root had empty cells is over, which one comparison per
// GROUP finds out (the steps in between are dead steps on an unchanged board).
template <class G, class = void> struct inert_when_full : std::false_type {};
template <class G> struct inert_when_full<G, std::enable_if_t<G::kInertWhenFull>> : std::true_type {};
template <class G, bool kOn> struct moves_left_of { BSG_HD static u32 get(const typename G::State&) { return 0xFFFFFFFFu; } };
template <class G> struct moves_left_of<G, true> { BSG_HD static u32 get(const typename G::State& s) { return G::moves_left(s); } };

// Games whose packed state must be replayed to be loaded (G::kReplayLoad, G::load_lazy,
// G::group_begin): the lean kernel spreads the replay over the step loop.
template <class G, class = void> struct replay_load : std::false_type {};
template <class G> struct replay_load<G, std::enable_if_t<G::kReplayLoad>> : std::true_type {};
template <class G, bool kOn> struct replay_ops {
  BSG_HD static void load(typename G::State&, const void*, int64_t, int64_t, u32&, u32&) {}
  BSG_HD static u32 group_begin(typename G::State&) { return 0u; }
};
template <class G> struct replay_ops<G, true> {
  BSG_HD static void load(typename G::State& s, const void* b, int64_t n, int64_t i, u32& mv, u32& oc) { G::load_lazy(s, b, n, i, mv, oc); }
  BSG_HD static u32 group_begin(typename G::State& s) { return G::group_begin(s); }
};

// Games with a compact root format for host -> device transfers (G::kCompactBits per root, bit
// strings back to back; G::load_compact / store_compact).  compact_bytes<G>::value > 0 says "has one".
template <class G, class = void> struct compact_bytes : std::integral_constant<int, 0> {};
template <class G> struct compact_bytes<G, std::enable_if_t<(G::kCompactBits > 0)>> : std::integral_constant<int, (G::kCompactBits + 7) / 8> {};
template <class G, class = void> struct compact_bits : std::integral_constant<int, 8 * G::kStateBytes> {};
template <class G> struct compact_bits<G, std::enable_if_t<(G::kCompactBits > 0)>> : std::integral_constant<int, G::kCompactBits> {};
template <class G, bool kOn> struct compact_ops {
  BSG_HD static void load(typename G::State&, const void*, int64_t) {}
  BSG_HD static void store(const typename G::State&, void*, int64_t) {}
};
template <class G> struct compact_ops<G, true> {
  BSG_HD static void load(typename G::State& s, const void* b, int64_t i) { G::load_compact(s, b, i); }
  BSG_HD static void store(const typename G::State& s, void* b, int64_t i) { G::store_compact(s, b, i); }
};

// Indirection so that the kernel body only names the fill functions of a game that has them.
template <class G, bool kOn> struct fill_words : std::integral_constant<int, 4> {};
template <class G> struct fill_words<G, true> : std::integral_constant<int, G::kFillWords> {};
template <class G, bool kOn> struct fill_ops {
  BSG_HD static bool done(const typename G::State&) { return true; }
  BSG_HD static void store(const typename G::State&, u32*) {}
  BSG_HD static u32 step(typename G::State&, u32, const typename G::Tables&, int&) { return 1u; }
};
template <class G> struct fill_ops<G, true> {
  BSG_HD static bool done(const typename G::State& s) { return G::fill_done(s); }
  BSG_HD static void store(const typename G::State& s, u32* out) { G::fill_store(s, out); }
  BSG_HD static u32 step(typename G::State& s, u32 r, const typename G::Tables& t, int& a) { return G::step_fill(s, r, t, a); }
};

// Games with value-only playouts (bsg_playout_io.value_only; see HexT).
template <class G, class = void> struct fillable : std::false_type {};
template <class G> struct fillable<G, std::enable_if_t<G::kFillable>> : std::true_type {};

// kFill: value-only launch of a fillable game — the lanes play plain stones until the board is
// full (no labels, no flood, no waiting) and park the board for fill_resolve_kernel, which
// reads the winner off it and publishes ret0 / per-root sums / totals[2].
// kLog: a lean launch that also logs the packed state before every move (and the action) at a
// fixed stride per playout (bsg_playout_io.state_log_stride) — the one-pass trajectory path;
// everything else the kFull variant offers (step limits, r_table, final states) stays there.
template <class G, bool kFull, bool kBase, bool kFill = false, bool kLog = false>
__global__ void __launch_bounds__(kThreads, kFull ? 1 : (kFill ? 4 : min_blocks<G>::value))
playout_kernel(const PlayoutArgs a, const typename G::Tables* __restrict__ gtab) {
  static_assert(!kFill || !kFull, "value-only playouts are a lean launch");
  static_assert(!kLog || (!kFull && !kBase && !kFill), "the logging variant is a plain lean launch");
  constexpr bool kAsync = G::kAsync && !kFill;
  __shared__ __align__(16) typename G::Tables tab;
  {
    const u32* src = reinterpret_cast<const u32*>(gtab);
    u32* dst = reinterpret_cast<u32*>(&tab);
    for (int i = threadIdx.x; i < (int)(sizeof(typename G::Tables) / 4); i += kThreads) dst[i] = src[i];
  }
  __syncthreads();

  // Work distribution: every warp draws chunks of kChunkPlayouts consecutive playouts from one
  // global counter (the hardware warp scheduler is not fair, so equal static slices finish at
  // very different times and the SMs drain early); inside a chunk its lanes claim playouts
  // from a warp-uniform cursor with one ballot, so the claim itself is convergent code.
  const u32 lane = threadIdx.x & 31;
  const u32 lane_lt = (1u << lane) - 1u;
  u32 begin = 0, count = 0, cursor = 0;
  bool drained = false;
  // kBase (forest leaves, evaluator batches: many rollouts per root): root of a playout without a
  // division per claim and per result.  A chunk holds at most kChunkPlayouts consecutive
  // playouts, so with rollouts >= kChunkPlayouts it spans at most two roots: one warp-uniform
  // division when the chunk is drawn (root and remainder of its first playout), one comparison
  // per claim; the lane keeps its root for the result (ncu: the two divisions were 11 % of the
  // warp instructions of the forest's leaf launches).
  u32 croot = 0, crem = 0;     // root of playout `begin`, and begin - croot * rollouts
  u32 my_root = 0;
  // Root staging (lean launches of single-plane games at one playout per root): when a warp
  // draws a chunk it copies the chunk's roots into shared memory with one coalesced load per
  // lane, so a claim — divergent code that runs in almost every group — reads its root with an
  // LDS instead of stalling on a dependent global load.  Two buffers per warp: lanes may still
  // claim the tail of the old chunk in the very claim that draws the new one.
  constexpr bool kStage = BSG_STAGE_ROOTS && !kFull && !kBase && !kFill && G::kPlanes == 1 &&
                          G::kStateBytes <= 16 && (kChunkPlayouts * G::kStateBytes) % 16 == 0;
  constexpr int kStageBytes = kStage ? kChunkPlayouts * G::kStateBytes : 16;
  __shared__ __align__(16) unsigned char root_sm[kStage ? kThreads / 32 : 1][2][kStageBytes];
  const bool stage = kStage && a.rollouts == 1u && !a.compact;   // launch-uniform
  u32 par = 0;                                      // buffer of the chunk (begin, count)

  typename G::State s;
  G::init(s);
  typename G::State prev;
  if (kFull) G::init(prev);
  // Per playout the kernel records only: its length t (frozen when it ends), whether it ended
  // in a terminal state (fin_term) and whether that was a win (fin_win).  A win is always the
  // last mover's, so the winner follows from the root's mover and the parity of t.
  // Flags are 0/1 words (not bools) so they stay in plain registers.
  // (fin_lose: only in a game whose mover can lose by moving — pentago.)
  u32 live = 0, have = 0, fin_term = 0, fin_win = 0, fin_lose = 0;
  u32 lid = 0, t = 0, limit = 0xFFFFFFFFu, root_mv = 0;
  constexpr bool kLazyDraw = inert_when_full<G>::value && !kFull && !kFill && !kAsync;
  constexpr bool kReplay = replay_load<G>::value && !kFull && !kFill && !kAsync;   // see replay_load
  u32 left0 = 0xFFFFFFFFu;   // kLazyDraw: empty cells of the root = the most moves the playout can make
  unsigned long long pid0 = a.first_playout;   // playout id = pid0 + lid (per root with root_playout_base)
  u32 my_steps = 0, my_done = 0, my_stuck = 0;
  int my_ret = 0;
  u32 rnew[4] = {0, 0, 0, 0}, rold[4] = {0, 0, 0, 0}, nblk = 0xFFFFFFFFu;   // kAsync draws (below)
#if BSG_SMEM_DRAWS
  // kAsync: the two Philox blocks a lane may be inside live in shared memory instead of eight
  // registers — block b at rows 4 * (b & 1) .. + 3, so the draw of move t is row t & 7: one LDS
  // in place of a compare and eight selects per iteration.  [row][thread]: no bank conflicts.
  __shared__ u32 draw_sm[kAsync ? 8 : 1][kAsync ? kThreads : 1];
#endif
  const u32 out = a.out_flags;   // which optional outputs exist (uniform)

  for (;;) {
    const u32 want = __ballot_sync(0xFFFFFFFFu, live == 0);
    const u32 nwant = __popc(want);
    u32 nbegin = 0, ncount = 0, nroot = 0, nrem = 0;
    const bool grab = !drained && cursor + nwant > count;  // warp-uniform: the chunk runs out in this claim
    if (grab) {
      u32 nb = 0;
      if (lane == 0) nb = atomicAdd(a.work_counter, kChunkPlayouts);
      nb = __shfl_sync(0xFFFFFFFFu, nb, 0);
      nbegin = min(nb, a.n_playouts);
      ncount = min((u32)kChunkPlayouts, a.n_playouts - nbegin);
      drained = ncount == 0;
      if constexpr (kBase) {
        nroot = nbegin / a.rollouts;
        nrem = nbegin - nroot * a.rollouts;
      }
      if constexpr (kStage) {
        if (stage && ncount) {
          // chunks start at multiples of kChunkPlayouts: the source is 16-byte aligned whenever
          // the root buffer is (cudaMalloc / torch); fall back to words otherwise
          const char* src = (const char*)a.roots + (size_t)(a.root0 + nbegin) * G::kStateBytes;
          unsigned char* dst = root_sm[threadIdx.x >> 5][par ^ 1u];
          const u32 bytes = ncount * G::kStateBytes;
          if ((((uintptr_t)src) & 15) == 0 && (bytes & 15u) == 0) {
            for (u32 i = lane; i < bytes / 16; i += 32) reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
          } else {
            for (u32 i = lane; i < bytes / 4; i += 32) reinterpret_cast<u32*>(dst)[i] = reinterpret_cast<const u32*>(src)[i];
          }
          __syncwarp();
        }
      }
    }
    if (live == 0) {
      if (have) {  // publish the finished playout
        if constexpr (kFill) {   // park the board; the winner is read off it by the next kernel
          u32 bw[fill_words<G, kFill>::value];
          fill_ops<G, kFill>::store(s, bw);
          uint4* dst = reinterpret_cast<uint4*>(a.fill_boards) + (size_t)lid * (fill_words<G, kFill>::value / 4);
#pragma unroll
          for (int q4 = 0; q4 < fill_words<G, kFill>::value / 4; ++q4)
            dst[q4] = make_uint4(bw[4 * q4], bw[4 * q4 + 1], bw[4 * q4 + 2], bw[4 * q4 + 3]);
        }
        int ret = (fin_term & fin_win) ? (int)(((root_mv ^ t) & 1u) << 1) - 1 : 0;
        if constexpr (kFill) ret = 0;
        if constexpr (G::kMoverCanLose) {
          if (fin_term & fin_lose) ret = 1 - (int)(((root_mv ^ t) & 1u) << 1);
        }
        if (!kFill && (out & kOutRet)) a.io.ret0[lid] = (int8_t)ret;
        if (out & kOutLength) a.io.length[lid] = (uint8_t)t;
        if (out & kOutTerminal) a.io.terminal[lid] = (uint8_t)fin_term;
        if (out & kOutRootSums) {
          u32 root = a.rollouts == 1 ? lid : (kBase ? my_root : lid / a.rollouts);
          if (!kFill && a.io.root_sum_ret0 && ret != 0) atomicAdd(a.io.root_sum_ret0 + root, ret);
          if (a.io.root_sum_steps) atomicAdd(a.io.root_sum_steps + root, (int)t);
        }
        my_steps += t;
        my_ret += ret;
        my_done += 1;
      }
      u32 idx = cursor + __popc(want & lane_lt);
      u32 base = begin, cnt = count, buf = par, broot = croot, brem = crem;
      if (idx >= count) { idx -= count; base = nbegin; cnt = ncount; buf = par ^ 1u; broot = nroot; brem = nrem; }
      have = idx < cnt ? 1u : 0u;
      if (have) {  // claim the next playout
        lid = base + idx;
        u32 root;
        if constexpr (kBase) {
          if (a.rollouts >= (u32)kChunkPlayouts) root = broot + ((brem + idx >= a.rollouts) ? 1u : 0u);
          else root = a.rollouts == 1 ? lid : lid / a.rollouts;
          my_root = root;
        } else {
          root = a.rollouts == 1 ? lid : lid / a.rollouts;
        }
        if (kStage && stage) G::load(s, root_sm[threadIdx.x >> 5][buf], kChunkPlayouts, idx);
        else if ((compact_bytes<G>::value > 0) && a.compact)
          compact_ops<G, (compact_bytes<G>::value > 0)>::load(s, a.roots, a.root0 + root);
        else if constexpr (kReplay) {}
        else G::load(s, a.roots, a.n_roots, a.root0 + root);
        u32 lazy_mv = 0, lazy_oc = kOutcomeNone;
        if constexpr (kReplay) replay_ops<G, kReplay>::load(s, a.roots, a.n_roots, a.root0 + root, lazy_mv, lazy_oc);
        bool skip = false;
        if (kBase) {
          const unsigned long long rb = a.io.root_playout_base[a.root0 + root];
          skip = rb == 0xFFFFFFFFFFFFFFFFull;
          pid0 = rb - (unsigned long long)root * a.rollouts;
        }
        t = 0;
        nblk = 0xFFFFFFFFu;
        root_mv = kReplay ? lazy_mv : (u32)G::mover(s);
        if constexpr (kLazyDraw) left0 = moves_left_of<G, kLazyDraw>::get(s);
        if (kFull) {
          limit = a.io.max_steps > 0 ? (u32)a.io.max_steps : 0xFFFFFFFFu;
          if (a.io.step_limit) limit = min(limit, (u32)a.io.step_limit[lid]);
          prev = s;
        }
        const bool term = (kReplay ? lazy_oc != (u32)kOutcomeNone
                                   : (kFill ? fill_ops<G, kFill>::done(s) : G::terminal(s))) || (kBase && skip);
        live = (!term && (!kFull || limit > 0)) ? 1u : 0u;
        fin_term = (kFull && !term) ? 0u : 1u;   // a lean playout always runs to a terminal state
        if constexpr (kReplay) {                 // the ROOT's result, from its meta word
          fin_win = lazy_oc == (lazy_mv ^ 1u) ? 1u : 0u;
          if constexpr (G::kMoverCanLose) fin_lose = lazy_oc == lazy_mv ? 1u : 0u;
        } else {
          fin_win = G::won(s) ? 1u : 0u;
          if constexpr (G::kMoverCanLose) fin_lose = G::lost(s) ? 1u : 0u;
        }
        if (kBase && skip) { fin_win = 0; fin_lose = 0; }   // a skipped slot holds no state
        if constexpr (G::kCanStick) {
          // non-terminal root without a legal action: only a root can be (see the game's note)
          if (!term && !G::has_move(s)) { live = 0; fin_term = 0; fin_win = 0; my_stuck += 1; }
        }
        if (kFull && !live && a.io.final_states) G::store(s, a.io.final_states, a.n_playouts, lid);
      }
    }
    cursor += nwant;
    if (grab) { cursor = cursor > count ? cursor - count : 0u; begin = nbegin; count = ncount; par ^= 1u; croot = nroot; crem = nrem; }
    if (!__any_sync(0xFFFFFFFFu, live != 0)) {
      if (!__any_sync(0xFFFFFFFFu, have != 0)) break;
      continue;
    }

    const unsigned long long pid = (kBase ? pid0 : a.first_playout) + lid;
    u32 rr[4];
    constexpr int kSteps = (kFull || kAsync || kFill) ? 4 : group_steps<G>::value;
    const u32 live_in = live;   // (kLazyDraw) lanes that play in this group
    // kReplay: a lane that still replays its root's history spends the WHOLE group on it (one
    // history move per iteration, idle once it runs out), so t stays 0 until a group boundary
    u32 replaying = 0;
    if constexpr (kReplay) replaying = replay_ops<G, kReplay>::group_begin(s);
    static_assert(kSteps % 4 == 0, "a group is a whole number of Philox blocks");
    if constexpr (!kAsync) {
      philox4x32_10((u32)pid, (u32)(pid >> 32), t >> 2, 0u, a.keys, rr);
    } else {
      // A lane of an asynchronous game does not move in every iteration, so the lanes of a warp
      // are at different phases of their Philox blocks.  The block computation stays convergent
      // by running every four iterations for ALL lanes on the block a lane will need next,
      // ceil(t / 4); the block it may still be inside, floor(t / 4), is kept beside it.  In the
      // next four iterations a lane draws at most t .. t + 3: always one of the two.
#if BSG_SMEM_DRAWS
      nblk = (t + 3u) >> 2;
      philox4x32_10((u32)pid, (u32)(pid >> 32), nblk, 0u, a.keys, rnew);
      {
        // block ceil(t/4) replaces block ceil(t/4) - 2 (or itself); floor(t/4) stays in the other half
        u32* row = &draw_sm[(nblk & 1u) * 4u][threadIdx.x];
#pragma unroll
        for (int i = 0; i < 4; ++i) row[i * kThreads] = rnew[i];
      }
#else
      if ((t >> 2) == nblk) {
#pragma unroll
        for (int i = 0; i < 4; ++i) rold[i] = rnew[i];
      }
      nblk = (t + 3u) >> 2;
      philox4x32_10((u32)pid, (u32)(pid >> 32), nblk, 0u, a.keys, rnew);
#endif
    }
    // Games with a long step (G::kRolledGroup: go, whose step is ~1300 instructions) keep the
    // group as a real loop: four unrolled copies do not fit the instruction cache, and a warp
    // whose lanes diverge inside a step then waits for instruction fetches most of the time
    // (ncu: "no instruction" was the top stall, 6.9 warps per issue).
    constexpr int kUnroll = rolled_group<G>::value ? 1 : kSteps;
#pragma unroll (kUnroll)
    for (int j = 0; j < kSteps; ++j) {
      u32 r, can = live;   // can: this lane makes a move in this iteration
      if constexpr (kReplay) can = live & (replaying ^ 1u);
      if constexpr (!kAsync && rolled_group<G>::value) {
        if (kSteps > 4 && j > 0 && (j & 3) == 0) philox4x32_10((u32)pid, (u32)(pid >> 32), t >> 2, 0u, a.keys, rr);
        const u32 lo2 = (j & 1) ? rr[1] : rr[0], hi2 = (j & 1) ? rr[3] : rr[2];   // no run-time register index
        r = (j & 2) ? hi2 : lo2;
      } else if constexpr (!kAsync) {
        // a live lane entered the group with t = 0 (mod kSteps) and has moved j times since:
        // its next block is t >> 2 (a finished lane's t is frozen: it draws garbage for its dead steps)
        if (kSteps > 4 && j > 0 && (j & 3) == 0) philox4x32_10((u32)pid, (u32)(pid >> 32), t >> 2, 0u, a.keys, rr);
        r = rr[j & 3];
      } else {
        G::ring(s);
#if BSG_EXTRA_RINGS
        // warp-voted extra rings (A/B switch): while at least BSG_EXTRA_RINGS_T lanes of the warp
        // still wait for their flood, the whole warp runs another ring (a no-op for the others)
#pragma unroll 1
        for (int x = 0; x < BSG_EXTRA_RINGS; ++x) {
          if (__popc(__ballot_sync(0xFFFFFFFFu, G::pending(s) && live)) < BSG_EXTRA_RINGS_T) break;
          G::ring(s);
        }
#endif
        can = G::pending(s) ? 0u : live;
#if BSG_SMEM_DRAWS
        r = draw_sm[t & 7u][threadIdx.x];
#else
        const bool in_new = (t >> 2) == nblk;
        const u32 d0 = in_new ? rnew[0] : rold[0], d1 = in_new ? rnew[1] : rold[1];
        const u32 d2 = in_new ? rnew[2] : rold[2], d3 = in_new ? rnew[3] : rold[3];
        const u32 lo2 = (t & 1u) ? d1 : d0, hi2 = (t & 1u) ? d3 : d2;
        r = (t & 2u) ? hi2 : lo2;
#endif
      }
      if (kFull) {
        if (a.io.r_table && can) r = a.io.r_table[(size_t)lid * a.io.r_stride + min(t, (u32)a.io.r_stride - 1)];
        if (can) {
          prev = s;
          if (a.io.state_log) {
            if (a.io.state_log_off) G::store(s, a.io.state_log, a.io.state_log_cap, a.io.state_log_off[lid] + t);
            else if ((int64_t)t < a.io.state_log_stride)
              G::store(s, a.io.state_log, a.io.state_log_cap, (int64_t)t * a.n_playouts + lid);
          }
        }
      }
      int action = 0;
      u32 nlegal = 1;
      u32 made = 1;   // kLazyEnd: 0 in the iteration that only finds the game over
      if constexpr (kLog) {
        if (can && (int64_t)t < a.io.state_log_stride)
          G::store(s, a.io.state_log, a.io.state_log_cap, (int64_t)t * a.n_playouts + lid);
      }
      if constexpr (kFill) {
        nlegal = fill_ops<G, kFill>::step(s, r, tab, action);
      } else if constexpr (G::kLazyEnd && !kFull) {
        nlegal = G::step_lazy(s, r, tab, action, made);
      } else if constexpr (!kAsync) {
        nlegal = G::step(s, r, tab, action);   // finished lanes run dead steps: no divergence
      } else {
        if (can) nlegal = G::step(s, r, tab, action);
      }
      if (kFull) {
        if (can) {
          if (a.io.actions && t < (u32)a.io.actions_stride)
            a.io.actions[(size_t)lid * a.io.actions_stride + t] = (int16_t)action;
          t += 1;
          const bool term = G::terminal(s);
          if (term || t >= limit) {
            live = 0;
            fin_term = term ? 1u : 0u;
            fin_win = (term && G::won(s)) ? 1u : 0u;
            if constexpr (G::kMoverCanLose) fin_lose = (term && G::lost(s)) ? 1u : 0u;
            if (a.io.final_states) {
              if (term && a.io.final_before_terminal) G::store(prev, a.io.final_states, a.n_playouts, lid);
              else G::store(s, a.io.final_states, a.n_playouts, lid);
            }
          }
        }
      } else if constexpr (kLazyDraw) {
        if constexpr (kLog) {
          if (live && a.io.actions && t < (u32)a.io.actions_stride)
            a.io.actions[(size_t)lid * a.io.actions_stride + t] = (int16_t)action;
        }
        // only a win ends the playout inside the group; the draw is found at the group's end
        t += live;
        live = G::won(s) ? 0u : live;
      } else {
        // lean bookkeeping: t counts the moves of the live playout and freezes when it ends;
        // fin_win was cleared at the claim and can only be set by the (single) ending move
        if constexpr (kLog) {
          if (can && made && a.io.actions && t < (u32)a.io.actions_stride)
            a.io.actions[(size_t)lid * a.io.actions_stride + t] = (int16_t)action;
        }
        const u32 term = (kFill ? fill_ops<G, kFill>::done(s) : G::terminal(s)) ? 1u : 0u;
        const u32 won = (!kFill && G::won(s)) ? 1u : 0u;
        fin_win |= can & won;
        if constexpr (G::kMoverCanLose) fin_lose |= can & (G::lost(s) ? 1u : 0u);
        if constexpr (G::kLazyEnd) t += can & made; else t += can;
        live &= (can & term) ^ 1u;
      }
    }
    if constexpr (kLazyDraw) {
      // a lane that entered the group live and is dead here was stopped by a win; then the draw:
      // the board filled up somewhere in this group and the moves counted past that were dead steps
      fin_win |= live_in & (live ^ 1u);
      if (t >= left0) { t = left0; live = 0u; }
    }
  }

  if (a.io.totals) {
    u32 st = my_steps, dn = my_done, sk = my_stuck;
    int rt = my_ret;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      st += __shfl_xor_sync(0xFFFFFFFFu, st, o);
      rt += __shfl_xor_sync(0xFFFFFFFFu, rt, o);
      dn += __shfl_xor_sync(0xFFFFFFFFu, dn, o);
      sk += __shfl_xor_sync(0xFFFFFFFFu, sk, o);
    }
    if ((threadIdx.x & 31) == 0) {
      atomicAdd((unsigned long long*)a.io.totals + 0, (unsigned long long)dn);
      atomicAdd((unsigned long long*)a.io.totals + 1, (unsigned long long)st);
      atomicAdd((unsigned long long*)a.io.totals + 2, (unsigned long long)(long long)rt);
      if (sk) atomicAdd((unsigned long long*)a.io.totals + 3, (unsigned long long)sk);
    }
  }
}


// Second kernel of a value-only launch: one thread per parked board reads the winner off it
// (G::fill_winner: one or two floods, every lane busy from the first ring) and publishes what
// playout_kernel<.., kFill> could not: ret0, the per-root return sums and totals[2].
template <class G>
__global__ void __launch_bounds__(kThreads)
fill_resolve_kernel(const PlayoutArgs a) {
  int my_ret = 0;
  for (u32 p = blockIdx.x * kThreads + threadIdx.x; p < a.n_playouts; p += gridDim.x * kThreads) {
    u32 bw[G::kFillWords];
    const uint4* src = reinterpret_cast<const uint4*>(a.fill_boards) + (size_t)p * (G::kFillWords / 4);
#pragma unroll
    for (int q4 = 0; q4 < G::kFillWords / 4; ++q4) {
      const uint4 v = src[q4];
      bw[4 * q4] = v.x; bw[4 * q4 + 1] = v.y; bw[4 * q4 + 2] = v.z; bw[4 * q4 + 3] = v.w;
    }
    const u32 root = a.rollouts == 1 ? p : p / a.rollouts;
    // a skipped root (root_playout_base = UINT64_MAX) parked nothing
    if (a.io.root_playout_base && a.io.root_playout_base[a.root0 + root] == 0xFFFFFFFFFFFFFFFFull) continue;
    const int ret = G::fill_winner(bw) == 0 ? 1 : -1;
    if (a.io.ret0) a.io.ret0[p] = (int8_t)ret;
    if (a.io.root_sum_ret0) atomicAdd(a.io.root_sum_ret0 + root, ret);
    my_ret += ret;
  }
  if (a.io.totals) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_ret += __shfl_xor_sync(0xFFFFFFFFu, my_ret, o);
    if ((threadIdx.x & 31) == 0 && my_ret != 0)
      atomicAdd((unsigned long long*)a.io.totals + 2, (unsigned long long)(long long)my_ret);
  }
}


// ---------------------------------------------------------------------------------------
// Batched ObservationTensor / LegalActionsMask writer.  HBM-write bound.
//
// Both outputs are 0/1 expansions of a per-state bit string (G::obs_words: kObsSize bits in
// the reference's flat CHW element order; G::mask_words: NumDistinctActions bits).  A CTA
// works on chunks of kChunk consecutive states:
//   phase 1  one thread per state: load the packed state (coalesced SoA read), build its bit
//            string in registers, park it in shared memory (row stride odd: no bank conflicts);
//   phase 2  the chunk's output is one contiguous, 16-byte aligned range of the [n, size]
//            array; thread j emits the j-th 16-byte store of it — 4 floats from 4 bits, or 16
//            mask bytes from 16 bits — so every warp store instruction covers 512 contiguous
//            bytes.  Bits come from shared memory with one funnel shift; a store that spans
//            two (or, for 7/9-action masks, more) states gathers across rows.
// Algorithmic HBM bytes per state = output + packed state; nothing is re-read from HBM.
// ---------------------------------------------------------------------------------------
#ifndef BSG_OBS_TMA
#define BSG_OBS_TMA 1
#endif
#ifndef BSG_OBS_TILE
#define BSG_OBS_TILE 512   // uint4 per staged tile (8 KB); a multiple of kChunk
#endif
constexpr int kChunk = 256;  // states per CTA chunk (= threads); multiple of 32 keeps stores aligned

template <int kBits, int kRow>
__device__ __forceinline__ u32 gather_bits(const u32* __restrict__ sm, int st, int r, int want, int nst) {
  // `want` (<= 16) bits starting at bit r of state st's string, continuing into later states
  const u32* p = sm + st * kRow;
  u32 x = __funnelshift_r(p[r >> 5], p[(r >> 5) + 1], r & 31);
  int got = kBits - r;
  if (got >= want) return x & ((1u << want) - 1u);
  x &= (1u << got) - 1u;
  for (st += 1; got < want && st < nst; ++st) {
    x |= sm[st * kRow] << got;  // rows hold kBits >= 1 valid bits, the rest is zero
    got += kBits;
  }
  return x & ((1u << want) - 1u);
}

// Dynamic shared memory of expand_kernel: the staging ring + one bit-string row per state.
// Aligned fast path (BSG_OBS_LUT): when a state's bit string is a whole number of 32-bit words
// (breakthrough 192 planes bits / 768 mask bits, othello 192) the rows are stored unpadded, the
// chunk's bits are ONE contiguous string, vector j is bits [4j, 4j+4) (planes) or [16j, 16j+16)
// (mask) of it, and its 16 output bytes come from a shared-memory table: planes 16 x uint4
// (nibble -> four floats), mask 256 x 8 bytes (byte -> eight 0/1 bytes).
#ifndef BSG_OBS_LUT
#define BSG_OBS_LUT 1
#endif
template <class G, bool kMask> struct expand_aligned_t {
  static constexpr bool value = BSG_OBS_LUT && BSG_OBS_TMA && ((kMask ? G::kNumActions : G::kObsSize) % 32 == 0);
};
template <class G, bool kMask> constexpr size_t expand_smem_bytes() {
  constexpr int kWords = kMask ? G::kMaskWords : G::kObsWords;
  constexpr int kRow = expand_aligned_t<G, kMask>::value ? kWords : ((kWords + 1) | 1);
#if BSG_OBS_TMA
  return (size_t)3 * BSG_OBS_TILE * 16 + (size_t)kChunk * kRow * 4 + 16 +
         (expand_aligned_t<G, kMask>::value ? (kMask ? 2048 : 256) : 0);
#else
  return 0;
#endif
}

// kMask = false: float32 observation planes; true: uint8 legal-action mask.
template <class G, bool kMask>
__global__ void __launch_bounds__(kChunk)
expand_kernel(const void* __restrict__ states, int64_t n, const uint8_t* __restrict__ players,
              void* __restrict__ out_) {
  constexpr int kBits = kMask ? G::kNumActions : G::kObsSize;
  constexpr int kWords = kMask ? G::kMaskWords : G::kObsWords;
  constexpr bool kAligned = expand_aligned_t<G, kMask>::value;
  constexpr int kRow = kAligned ? kWords : ((kWords + 1) | 1);   // general: + one zero word for the funnel read; odd
  constexpr int kPer = kMask ? 16 : 4;              // output elements per 16-byte store
#if BSG_OBS_TMA
  // dynamic shared memory: [three staging tiles][kChunk rows]  (expand_smem_bytes below)
  extern __shared__ __align__(128) unsigned char expand_smem[];
  u32* const sm = reinterpret_cast<u32*>(expand_smem + 3 * BSG_OBS_TILE * 16);
#else
  __shared__ u32 sm[kChunk * kRow];
#endif
#if BSG_OBS_TMA
  // Output staging for bulk (TMA) stores: three 8 KB tiles in a ring (dynamic shared memory, so
  // that the writers with long rows — hex / y planes, breakthrough's 768-action mask — fit too).  A tile is written with
  // STS.128, made visible to the async proxy, and one thread hands it to the copy engine; the
  // ring of three needs one barrier per tile (see the wait_group.read below).
  constexpr int kTileVec = BSG_OBS_TILE;
  constexpr bool kTma = true;
  uint4* const stage = reinterpret_cast<uint4*>(expand_smem);
  int tsel = 0;
  // look-up table of the aligned path, behind the rows (16-byte aligned)
  unsigned char* const lut_raw = expand_smem + 3 * BSG_OBS_TILE * 16 + (((size_t)kChunk * kRow * 4 + 15) & ~(size_t)15);
  if constexpr (kAligned) {
    if constexpr (kMask) {
      uint2* lut = reinterpret_cast<uint2*>(lut_raw);
      const u32 b = threadIdx.x;   // kChunk = 256 threads: one entry each
      lut[b] = make_uint2(((b & 15u) * 0x00204081u) & 0x01010101u, ((b >> 4) * 0x00204081u) & 0x01010101u);
    } else {
      uint4* lut = reinterpret_cast<uint4*>(lut_raw);
      if (threadIdx.x < 16) {
        const u32 x = threadIdx.x;
        lut[x] = make_uint4((x & 1u) * 0x3F800000u, ((x >> 1) & 1u) * 0x3F800000u, ((x >> 2) & 1u) * 0x3F800000u,
                            ((x >> 3) & 1u) * 0x3F800000u);
      }
    }
    __syncthreads();
  }
#else
  constexpr bool kTma = false;
#endif
  const int64_t n_chunks = (n + kChunk - 1) / kChunk;
  for (int64_t c = blockIdx.x; c < n_chunks; c += gridDim.x) {
    const int64_t s0 = c * kChunk;
    const int nst = (int)min((int64_t)kChunk, n - s0);
    // ---- phase 1
    if ((int)threadIdx.x < nst) {
      typename G::State s;
      G::load(s, states, n, s0 + threadIdx.x);
      u32 w[kWords];
      if (kMask) {
        G::mask_words(s, w);
      } else {
        int player = players ? (int)players[s0 + threadIdx.x] : G::mover(s);
        G::obs_words(s, player, w);
      }
      u32* row = sm + threadIdx.x * kRow;
#pragma unroll
      for (int i = 0; i < kWords; ++i) row[i] = w[i];
#pragma unroll
      for (int i = kWords; i < kRow; ++i) row[i] = 0u;
    }
    __syncthreads();
    // ---- phase 2
    const int total = nst * kBits;                  // output elements of this chunk
    const int n16 = total / kPer;
    uint4* out16 = reinterpret_cast<uint4*>(reinterpret_cast<char*>(out_) +
                                            s0 * (int64_t)kBits * (kMask ? 1 : 4));
#if BSG_OBS_TMA
    if constexpr (kTma) {
      const int ntiles = (n16 + kTileVec - 1) / kTileVec;
      for (int t = 0; t < ntiles; ++t) {
        uint4* sbuf = stage + tsel * kTileVec;
#pragma unroll
        for (int q = 0; q < kTileVec / kChunk; ++q) {
          const int j = t * kTileVec + q * kChunk + threadIdx.x;
          if constexpr (kAligned) {
            if (j < n16) {
              uint4 v;
              if constexpr (kMask) {
                const u32 x = (sm[j >> 1] >> ((j & 1) * 16)) & 0xFFFFu;
                const uint2* lut = reinterpret_cast<const uint2*>(lut_raw);
                const uint2 lo = lut[x & 255u], hi = lut[x >> 8];
                v = make_uint4(lo.x, lo.y, hi.x, hi.y);
              } else {
                const u32 x = (sm[j >> 3] >> ((j & 7) * 4)) & 15u;
                v = reinterpret_cast<const uint4*>(lut_raw)[x];
              }
              sbuf[q * kChunk + threadIdx.x] = v;
            }
          } else if (j < n16) {
            const int e0 = j * kPer;
            const int st = e0 / kBits, r = e0 - st * kBits;
            u32 x = gather_bits<kBits, kRow>(sm, st, r, kPer, nst);
            uint4 v;
            if (kMask) {
              v.x = ((x & 15u) * 0x00204081u) & 0x01010101u;
              v.y = (((x >> 4) & 15u) * 0x00204081u) & 0x01010101u;
              v.z = (((x >> 8) & 15u) * 0x00204081u) & 0x01010101u;
              v.w = (((x >> 12) & 15u) * 0x00204081u) & 0x01010101u;
            } else {
              v.x = (x & 1u) * 0x3F800000u;
              v.y = ((x >> 1) & 1u) * 0x3F800000u;
              v.z = ((x >> 2) & 1u) * 0x3F800000u;
              v.w = ((x >> 3) & 1u) * 0x3F800000u;
            }
            sbuf[q * kChunk + threadIdx.x] = v;
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        // Before anyone passes this barrier and writes the NEXT tile, the store that last used
        // the next tile's buffer (issued two tiles ago) must have been read out of shared memory:
        // all committed groups but the newest one.
        if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
          const int vecs = min(kTileVec, n16 - t * kTileVec);
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                       :: "l"(out16 + (size_t)t * kTileVec), "r"((u32)__cvta_generic_to_shared(sbuf)), "r"(vecs * 16) : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        tsel = tsel == 2 ? 0 : tsel + 1;
      }
    }
#endif
    if constexpr (!kTma) {
    for (int j = threadIdx.x; j < n16; j += kChunk) {
      const int e0 = j * kPer;
      const int st = e0 / kBits, r = e0 - st * kBits;
      u32 x = gather_bits<kBits, kRow>(sm, st, r, kPer, nst);
      uint4 v;
      if (kMask) {  // 4 bits -> 4 bytes: (nibble * 0x00204081) & 0x01010101
        v.x = ((x & 15u) * 0x00204081u) & 0x01010101u;
        v.y = (((x >> 4) & 15u) * 0x00204081u) & 0x01010101u;
        v.z = (((x >> 8) & 15u) * 0x00204081u) & 0x01010101u;
        v.w = (((x >> 12) & 15u) * 0x00204081u) & 0x01010101u;
      } else {      // bit -> 0.0f / 1.0f
        v.x = (x & 1u) * 0x3F800000u;
        v.y = ((x >> 1) & 1u) * 0x3F800000u;
        v.z = ((x >> 2) & 1u) * 0x3F800000u;
        v.w = ((x >> 3) & 1u) * 0x3F800000u;
      }
      __stcs(out16 + j, v);
    }
    }
    // tail of the very last chunk when its size is not a multiple of 16 bytes
    const int rest = total - n16 * kPer;
    if ((int)threadIdx.x < rest) {
      const int e = n16 * kPer + threadIdx.x;
      const int st = e / kBits, r = e - st * kBits;
      u32 bit = (sm[st * kRow + (r >> 5)] >> (r & 31)) & 1u;
      if (kMask) reinterpret_cast<uint8_t*>(out16)[e] = (uint8_t)bit;
      else reinterpret_cast<float*>(out16)[e] = bit ? 1.f : 0.f;
    }
    __syncthreads();
  }
#if BSG_OBS_TMA
  if (kTma && threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
#endif
}

// ---------------------------------------------------------------------------------------
// Integer-issue microbenchmark (roofline denominator).  8 independent chains per thread.
// ---------------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(kThreads) int_peak_kernel(int iters, u32* sink) {
  u32 x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = threadIdx.x * 2654435761u + i * 40503u + blockIdx.x;
  u32 k = blockIdx.x | 1u;
  const u32 sh = 0x02000000u << (iters & 1);  // opaque to ptxas: keeps mul.hi an IMAD.HI
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int rep = 0; rep < 8; ++rep) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (KIND == 0) x[i] = (x[i] ^ k) + (x[(i + 1) & 7] & 0x55555555u) ;          // LOP3 + IADD3
        else if (KIND == 1) x[i] = x[i] * 0x9E3779B1u + k;                             // IMAD
        else if (KIND == 2) { x[i] = (x[i] ^ (x[(i + 3) & 7] >> 3)); x[i] = x[i] * 0x9E3779B1u + k; } // SHF/LOP + IMAD
        else if (KIND == 3) { x[i] = (x[i] ^ x[(i + 3) & 7]) * 0x9E3779B1u + k; }       // LOP3 + IMAD, 1:1
        else if (KIND == 4) { asm("mul.hi.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(0x9E3779B1u)); x[i] |= 0x80000000u; }  // IMAD.HI (+1 LOP3 keeps the chain alive; counted)
        else if (KIND == 5) {                                                            // IMAD.WIDE + LOP3
          unsigned long long w;
          asm("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(x[i]), "r"(0x9E3779B1u));
          x[i] = (u32)w ^ (u32)(w >> 32) ^ k;
        } else {                                                                         // LOP3 + IMAD.HI (power-of-two multiplier from a register = a right shift)
          u32 y = x[i] ^ x[(i + 3) & 7] ^ k;
          asm("mul.hi.u32 %0, %1, %2;" : "=r"(x[i]) : "r"(y), "r"(sh));
          x[i] ^= y;
        }
      }
    }
  }
  u32 acc = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc ^= x[i];
  if (acc == 0x12345678u) sink[0] = acc;
}

}  // namespace bsg

// =========================================================================================
// C ABI
// =========================================================================================
using namespace bsg;

struct bsg_engine {
  int game = -1;
  int device = -1;
  int sm_count = 0;
  void* tables_dev = nullptr;
  // scratch for the *_host wrappers
  void* d_roots = nullptr; size_t d_roots_bytes = 0;
  void* d_out = nullptr; size_t d_out_bytes = 0;
  void* d_fill = nullptr; size_t d_fill_bytes = 0;   // parked boards of value-only launches
  void* d_scan = nullptr; size_t d_scan_bytes = 0;   // tile sums / offsets of bsg_trajectory_scan
  int value_only = 0;                                // bsg_engine_set_value_only: the *_host playouts
  void* h_pinned = nullptr; size_t h_pinned_bytes = 0;
  uint64_t* d_totals = nullptr;
  uint32_t* d_work = nullptr;          // ring of per-launch work counters (playout_kernel)
  unsigned work_slot = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t aux_stream = nullptr;   // second lane of the chunked host path (copy/compute overlap)
  cudaEvent_t ev_a = nullptr, ev_b = nullptr;
  ncclComm_t comm = nullptr;           // bsg_comm_init: this engine's own NCCL communicator
  int comm_ranks = 1, comm_rank = 0;
  std::string err;
};

namespace {
int fail(bsg_engine* e, int code, const std::string& msg) {
  if (e) e->err = msg;
  return code;
}
int cuda_fail(bsg_engine* e, cudaError_t st, const char* what) {
  return fail(e, BSG_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(st));
}
#define BSG_CUDA(e, call)                                   \
  do {                                                      \
    cudaError_t st__ = (call);                              \
    if (st__ != cudaSuccess) return cuda_fail(e, st__, #call); \
  } while (0)

int ensure(bsg_engine* e, void** p, size_t* have, size_t need, bool pinned) {
  if (*have >= need) return BSG_OK;
  if (*p) { if (pinned) cudaFreeHost(*p); else cudaFree(*p); *p = nullptr; *have = 0; }
  size_t cap = need + need / 4 + 256;
  BSG_CUDA(e, pinned ? cudaMallocHost(p, cap) : cudaMalloc(p, cap));
  *have = cap;
  return BSG_OK;
}

struct FillWordsOp {
  template <class G> int operator()() const {
    if constexpr (fillable<G>::value) return G::kFillWords; else return 0;
  }
};

struct FactOp {
  int which;
  template <class G> int operator()() const {
    return which == 0 ? G::kNumActions : (which == 1 ? G::kMaxLen : (which == 2 ? G::kObsSize : G::kStateBytes));
  }
};

struct CreateOp {
  bsg_engine* e;
  template <class G> int operator()() const {
    typename G::Tables t;
    G::fill_tables(t);
    BSG_CUDA(e, cudaMalloc(&e->tables_dev, sizeof(t)));
    BSG_CUDA(e, cudaMemcpy(e->tables_dev, &t, sizeof(t), cudaMemcpyHostToDevice));
    return BSG_OK;
  }
};


// ---- GPU-resident search forest (bsg_forest.cuh): one thread per tree ---------------------
constexpr int kForestThreads = 64;
template <class G>
__global__ void forest_reset_kernel(ForestView F) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < F.n_trees) ForestTree<G>::reset(F, v);
}
// descend / finish: one WARP per tree (ForestTree::descend_warp); the other forest kernels are
// one thread per tree.
constexpr int kForestWarpThreads = 128;   // 4 trees per CTA
template <class G>
__global__ void __launch_bounds__(kForestWarpThreads)
forest_descend_kernel(ForestView F, const typename G::Tables* __restrict__ tab) {
  const int v = (blockIdx.x * kForestWarpThreads + threadIdx.x) >> 5;
  if (v >= F.n_trees) return;           // whole warps leave
  ForestTree<G>::descend_warp(F, v, *tab);
}
template <class G>
__global__ void __launch_bounds__(kForestWarpThreads) forest_finish_kernel(ForestView F) {
  const int v = (blockIdx.x * kForestWarpThreads + threadIdx.x) >> 5;
  if (v >= F.n_trees) return;
  ForestTree<G>::finish_round_warp(F, v);
  __syncwarp();
  // trees that still have simulations to run after this round (the host reads the counter
  // after the last enqueued round of a search step)
  if ((threadIdx.x & 31) == 0 && F.sims_left[v] > 0) atomicAdd(F.active, 1u);
}
template <class G>
__global__ void forest_expand_root_kernel(ForestView F) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < F.n_trees) ForestTree<G>::expand_root(F, v);
}
// Trajectory log compaction (bsg_trajectory_*).  Exclusive scan of the one-byte lengths in three
// small kernels: per-CTA sums of 4096 lengths (16 per thread, one 16-byte load), one CTA scans the
// CTA sums, every CTA rescans its tile and writes its offsets.
constexpr int kScanTile = 4096;
__device__ __forceinline__ u32 scan_bytes_sum(u32 x) {       // sum of the four bytes of x
  return (x & 0xFFu) + ((x >> 8) & 0xFFu) + ((x >> 16) & 0xFFu) + (x >> 24);
}
__device__ __forceinline__ void scan_load16(const uint8_t* len, int64_t n, int64_t i0, u32 w[4]) {
  if (i0 + 16 <= n && (((uintptr_t)(len + i0)) & 15) == 0) {
    const uint4 v = *reinterpret_cast<const uint4*>(len + i0);
    w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
  } else {
    for (int k = 0; k < 4; ++k) {
      u32 x = 0;
      for (int j = 0; j < 4; ++j) { const int64_t i = i0 + 4 * k + j; if (i < n) x |= (u32)len[i] << (8 * j); }
      w[k] = x;
    }
  }
}
__device__ __forceinline__ u32 scan_block_exclusive(u32 mine, u32* sm, u32* total) {   // 256 threads
  const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  u32 inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const u32 v = __shfl_up_sync(0xFFFFFFFFu, inc, o); if (lane >= (u32)o) inc += v; }
  if (lane == 31) sm[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    u32 w = lane < 8 ? sm[lane] : 0u, winc = w;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) { const u32 v = __shfl_up_sync(0xFFFFFFFFu, winc, o); if (lane >= (u32)o) winc += v; }
    if (lane < 8) sm[lane] = winc - w;
    if (lane == 7) sm[8] = winc;
  }
  __syncthreads();
  if (total) *total = sm[8];
  return sm[warp] + inc - mine;
}
__global__ void __launch_bounds__(256) traj_tile_sums_kernel(const uint8_t* len, int64_t n, u32* tile_sums) {
  __shared__ u32 sm[9];
  u32 w[4];
  scan_load16(len, n, (int64_t)blockIdx.x * kScanTile + threadIdx.x * 16, w);
  const u32 mine = scan_bytes_sum(w[0]) + scan_bytes_sum(w[1]) + scan_bytes_sum(w[2]) + scan_bytes_sum(w[3]);
  u32 total;
  scan_block_exclusive(mine, sm, &total);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
__global__ void __launch_bounds__(256) traj_tile_scan_kernel(const u32* tile_sums, int n_tiles, int64_t* tile_off, int64_t* off_n) {
  __shared__ u32 sm[9];
  __shared__ long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n_tiles; base += 256) {
    const int i = base + threadIdx.x;
    const u32 mine = i < n_tiles ? tile_sums[i] : 0u;
    u32 total;
    const u32 ex = scan_block_exclusive(mine, sm, &total);
    if (i < n_tiles) tile_off[i] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *off_n = carry;
}
__global__ void __launch_bounds__(256) traj_tile_offsets_kernel(const uint8_t* len, int64_t n, const int64_t* tile_off, int64_t* off) {
  __shared__ u32 sm[9];
  u32 w[4];
  const int64_t i0 = (int64_t)blockIdx.x * kScanTile + threadIdx.x * 16;
  scan_load16(len, n, i0, w);
  const u32 mine = scan_bytes_sum(w[0]) + scan_bytes_sum(w[1]) + scan_bytes_sum(w[2]) + scan_bytes_sum(w[3]);
  long long run = tile_off[blockIdx.x] + scan_block_exclusive(mine, sm, nullptr);
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int64_t i = i0 + k;
    if (i < n) off[i] = run;
    run += (w[k >> 2] >> (8 * (k & 3))) & 0xFFu;
  }
}
template <class G>
__global__ void traj_compact_kernel(const void* strided, int64_t n, int64_t stride, const int64_t* off, int64_t total,
                                    const int16_t* acts, int32_t acts_stride, void* out_states,
                                    int64_t* out_actions, int64_t* out_playout, int64_t* out_step) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < total; r += (int64_t)gridDim.x * blockDim.x) {
    int64_t lo = 0, hi = n;                      // the playout p with off[p] <= r < off[p + 1]
    while (hi - lo > 1) {
      const int64_t mid = (lo + hi) >> 1;
      if (off[mid] <= r) lo = mid; else hi = mid;
    }
    const int64_t t = r - off[lo];
    typename G::State s;
    G::load(s, strided, n * stride, t * n + lo);      // move-major log: slot = t * n + playout
    G::store(s, out_states, total, r);
    if (out_actions) out_actions[r] = acts ? (int64_t)acts[lo * acts_stride + t] : -1;
    if (out_playout) out_playout[r] = lo;
    if (out_step) out_step[r] = t;
  }
}
__global__ void forest_fill_alive_kernel(int64_t* dst, int64_t value, const uint8_t* alive, int n) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < n) dst[v] = (!alive || alive[v]) ? value : 0;
}
template <class G>
__global__ void actor_step_kernel(ForestView F, ActorView A, const typename G::Tables* __restrict__ tab) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < A.n_games) Actor<G>::step(F, A, v, *tab);
}
template <class G>
__global__ void actor_emit_kernel(EmitView E) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (int64_t)E.n_games * E.max_moves) actor_emit_one<G>(E, (int)(i % E.n_games), (int)(i / E.n_games));
}
__global__ void forest_fill_kernel(int64_t* dst, int64_t value, const int64_t* per_tree, int n) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < n) dst[v] = per_tree ? per_tree[v] : value;
}
__global__ void forest_pack_kernel(ForestView F, const int64_t* base, unsigned long long* out) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < F.n_trees) forest_pack_root(F, v, base, out);
}
__global__ void forest_install_kernel(ForestView F, const int64_t* base, const int64_t* summed) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < F.n_trees) forest_install_root(F, v, base, summed);
}

__global__ void forest_totals_kernel(const uint64_t* step_totals, uint64_t* acc) {
  if (threadIdx.x < 4) acc[threadIdx.x] += step_totals[threadIdx.x];
}
__global__ void forest_base_add_kernel(int64_t* base, const int64_t* summed, int n_actions) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a < n_actions) {
    base[2 * a] += summed[a * bsg::kMergeFields];
    base[2 * a + 1] += summed[a * bsg::kMergeFields + 1];
  }
}

// NCCL, loaded at run time (a torch process already has libnccl.so.2 mapped; any other host
// gets the system library).  Only the five calls the root merge needs.
struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
  std::string why;
};
const NcclApi& nccl_api() {
  static NcclApi api = [] {
    NcclApi a;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { a.why = std::string("dlopen(libnccl.so.2): ") + dlerror(); return a; }
    a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
    a.AllReduce = (decltype(a.AllReduce))dlsym(h, "ncclAllReduce");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
    a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllReduce && a.GetErrorString;
    if (!a.ok) a.why = "libnccl lacks a required symbol";
    return a;
  }();
  return api;
}

template <class G, bool kFull, bool kBase = false, bool kFill = false, bool kLog = false>
int launch_playouts(bsg_engine* e, const PlayoutArgs& a, cudaStream_t st) {
  // function attributes belong to the current device: one cached answer per device, set under a
  // lock (RootParallelMCTS walks several engines of one process from a thread pool)
  static int per_device[kMaxDevices] = {};
  static std::mutex mu;
  int blocks_per_sm;
  {
    std::lock_guard<std::mutex> lock(mu);
    int& slot = per_device[e->device % kMaxDevices];
    if (slot == 0) {
      int b = 0;
      // all of the L1/shared array as shared memory: the tables of every resident CTA must fit.
      // Games whose state lives in local memory (G::kLocalHeavy: havannah, go) have small tables
      // and want the array as L1 instead (BSG_LOCAL_CARVEOUT per cent of it stay shared memory).
      BSG_CUDA(e, cudaFuncSetAttribute(playout_kernel<G, kFull, kBase, kFill, kLog>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                       local_heavy<G>::value ? BSG_LOCAL_CARVEOUT : (int)cudaSharedmemCarveoutMaxShared));
      BSG_CUDA(e, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, playout_kernel<G, kFull, kBase, kFill, kLog>, kThreads, 0));
      slot = b > 0 ? b : 1;
    }
    blocks_per_sm = slot;
  }
  int64_t max_grid = (int64_t)e->sm_count * blocks_per_sm;
  int64_t need = ((int64_t)a.n_playouts + kThreads - 1) / kThreads;
  int grid = (int)(need < max_grid ? need : max_grid);
  if (grid < 1) grid = 1;
  PlayoutArgs b = a;
  b.out_flags = out_flags_of(a.io);
  if (!b.work_counter) {
    // work counter of this launch: next slot of the engine's ring, zeroed in stream order
    // (a caller that already zeroed a counter of its own passes it in: bsg_playouts_host's
    // small-call path folds it into its one memset)
    b.work_counter = e->d_work + (e->work_slot++ % kWorkSlots);
    BSG_CUDA(e, cudaMemsetAsync(b.work_counter, 0, sizeof(uint32_t), st));
  }
  if constexpr (kFill) {
    // (the chunked host path sized the buffer for the whole call, so this does not reallocate
    // under a chunk that is still running)
    int rc = ensure(e, &e->d_fill, &e->d_fill_bytes, (size_t)(a.fill_base + a.n_playouts) * G::kFillWords * 4, false);
    if (rc != BSG_OK) return rc;
    b.fill_boards = (uint32_t*)e->d_fill + (size_t)a.fill_base * G::kFillWords;
  }
  playout_kernel<G, kFull, kBase, kFill, kLog><<<grid, kThreads, 0, st>>>(b, (const typename G::Tables*)e->tables_dev);
  BSG_CUDA(e, cudaGetLastError());
  if constexpr (kFill) {
    int rgrid = (int)(need < (int64_t)e->sm_count * 8 ? need : (int64_t)e->sm_count * 8);
    fill_resolve_kernel<G><<<rgrid < 1 ? 1 : rgrid, kThreads, 0, st>>>(b);
    BSG_CUDA(e, cudaGetLastError());
  }
  return BSG_OK;
}

template <class G>
int launch_value_only(bsg_engine* e, const PlayoutArgs& a, cudaStream_t st, bool full) {
  if constexpr (fillable<G>::value) {
    if (full)
      return fail(e, BSG_ERR_BAD_ARG, "value_only excludes step limits, r_table, action / state logs and final states");
    return a.io.root_playout_base ? launch_playouts<G, false, true, true>(e, a, st)
                                  : launch_playouts<G, false, false, true>(e, a, st);
  } else {
    return fail(e, BSG_ERR_BAD_ARG, "value_only needs a game whose full board has exactly one winner (hex, y)");
  }
}

struct PlayoutOp {
  bsg_engine* e; PlayoutArgs a; cudaStream_t st;
  template <class G> int operator()() const {
    const bsg_playout_io& io = a.io;
    bool full = io.max_steps > 0 || io.step_limit || io.r_table || io.actions || io.final_states || io.state_log;
    if (io.value_only) return launch_value_only<G>(e, a, st, full);
    // the one-pass trajectory launch: a strided state log (+ actions), nothing else of kFull
    const bool log_only = io.state_log && !io.state_log_off && io.state_log_stride > 0 && !io.root_playout_base &&
                          !(io.max_steps > 0 || io.step_limit || io.r_table || io.final_states);
    if (log_only) return launch_playouts<G, false, false, false, true>(e, a, st);
    if (io.root_playout_base)
      return full ? launch_playouts<G, true, true>(e, a, st) : launch_playouts<G, false, true>(e, a, st);
    return full ? launch_playouts<G, true>(e, a, st) : launch_playouts<G, false>(e, a, st);
  }
};

// Packed SoA plane table of a game: {count, element bytes...}.
struct PlaneOp {
  int* bytes;
  template <class G> int operator()() const {
    for (int p = 0; p < G::kPlanes; ++p) bytes[p] = G::plane_bytes(p);
    return G::kPlanes;
  }
};

struct ObsOp2 {
  bsg_engine* e; const void* states; int64_t n; const uint8_t* players; float* obs; uint8_t* mask;
  cudaStream_t st;
  template <class G, bool kMask> int launch(void* out) const {
    static int per_device[kMaxDevices] = {};
    static std::mutex mu;
    int blocks_per_sm;
    {
      std::lock_guard<std::mutex> lock(mu);
      int& slot = per_device[e->device % kMaxDevices];
      if (slot == 0) {
        int b = 0;
        BSG_CUDA(e, cudaFuncSetAttribute(expand_kernel<G, kMask>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)expand_smem_bytes<G, kMask>()));
        BSG_CUDA(e, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, expand_kernel<G, kMask>, kChunk,
                                                                  expand_smem_bytes<G, kMask>()));
        slot = b > 0 ? b : 1;
      }
      blocks_per_sm = slot;
    }
    int64_t cap = (int64_t)e->sm_count * blocks_per_sm;
    int64_t need = (n + kChunk - 1) / kChunk;
    int grid = (int)(need < cap ? need : cap);
    expand_kernel<G, kMask><<<grid < 1 ? 1 : grid, kChunk, expand_smem_bytes<G, kMask>(), st>>>(states, n, players, out);
    BSG_CUDA(e, cudaGetLastError());
    return BSG_OK;
  }
  template <class G> int operator()() const {
    int rc = BSG_OK;
    if (obs && (rc = launch<G, false>(obs)) != BSG_OK) return rc;
    if (mask && (rc = launch<G, true>(mask)) != BSG_OK) return rc;
    return BSG_OK;
  }
};
}  // namespace

extern "C" {

int bsg_version(void) { return BSG_VERSION; }

const char* bsg_error_string(int code) {
  switch (code) {
    case BSG_OK: return "ok";
    case BSG_ERR_BAD_GAME: return "unsupported game";
    case BSG_ERR_BAD_ARG: return "bad argument";
    case BSG_ERR_ILLEGAL_ACTION: return "illegal action";
    case BSG_ERR_CUDA: return "CUDA error";
    case BSG_ERR_NO_DEVICE: return "no CUDA device";
    case BSG_ERR_STUCK_LANE: return "non-terminal state without legal actions";
    case BSG_ERR_NCCL: return "NCCL error";
    default: return "unknown error";
  }
}

int bsg_game_id(const char* g) {
  if (!g) return BSG_ERR_BAD_GAME;
  std::string s(g);
  if (s == "tic_tac_toe" || s == "tic_tac_toe()") return BSG_TIC_TAC_TOE;
  if (s == "connect_four" || s == "connect_four()") return BSG_CONNECT_FOUR;
  if (s == "breakthrough" || s == "breakthrough()" || s == "breakthrough(rows=8,columns=8)" ||
      s == "breakthrough(columns=8,rows=8)") return BSG_BREAKTHROUGH;
  if (s == "hex" || s == "hex()" || s == "hex(board_size=11)") return BSG_HEX;
  if (s == "hex(board_size=5)") return BSG_HEX5_HOST_ONLY;
  if (s == "pentago" || s == "pentago()" || s == "pentago(ansi_color_output=False)" ||
      s == "pentago(ansi_color_output=false)") return BSG_PENTAGO;
  if (s == "othello" || s == "othello()") return BSG_OTHELLO;
  if (s == "y" || s == "y()" || s == "y(board_size=19)" || s == "y(ansi_color_output=False,board_size=19)" ||
      s == "y(ansi_color_output=false,board_size=19)") return BSG_Y;
  if (s == "y(board_size=9)" || s == "y(ansi_color_output=False,board_size=9)" ||
      s == "y(ansi_color_output=false,board_size=9)") return BSG_Y9_HOST_ONLY;
  if (s == "havannah" || s == "havannah()" || s == "havannah(board_size=8)" ||
      s == "havannah(ansi_color_output=False,board_size=8)" || s == "havannah(ansi_color_output=false,board_size=8)")
    return BSG_HAVANNAH;
  if (s == "go(board_size=9)" || s == "go(board_size=9,komi=7.5)" || s == "go(board_size=9,handicap=0,komi=7.5)" ||
      s == "go(board_size=9,handicap=0)") return BSG_GO;
  if (s == "go(board_size=7,komi=4.5)" || s == "go(board_size=7,handicap=0,komi=4.5)") return BSG_GO7_HOST_ONLY;
  if (s == "havannah(board_size=4)" || s == "havannah(ansi_color_output=False,board_size=4)" ||
      s == "havannah(ansi_color_output=false,board_size=4)") return BSG_HAVANNAH4_HOST_ONLY;
  return BSG_ERR_BAD_GAME;
}

int bsg_num_distinct_actions(int game) { return dispatch_game(game, FactOp{0}); }
int bsg_max_game_length(int game) { return dispatch_game(game, FactOp{1}); }
int bsg_observation_size(int game) { return dispatch_game(game, FactOp{2}); }
int bsg_state_bytes(int game) { return dispatch_game(game, FactOp{3}); }
int bsg_plane_bytes(int game, int* bytes_out, int capacity) {
  int tmp[16];
  const int n = dispatch_game(game, PlaneOp{tmp});
  if (n < 0) return n;
  if (!bytes_out || capacity < n) return BSG_ERR_BAD_ARG;
  for (int p = 0; p < n; ++p) bytes_out[p] = tmp[p];
  return n;
}

int bsg_observation_shape(int game, int shape[3]) {
  switch (game) {
    case BSG_TIC_TAC_TOE: shape[0] = 3; shape[1] = 3; shape[2] = 3; return BSG_OK;   // tic_tac_toe.h:105-107
    case BSG_CONNECT_FOUR: shape[0] = 3; shape[1] = 6; shape[2] = 7; return BSG_OK;  // connect_four.h:112-114
    case BSG_BREAKTHROUGH: shape[0] = 3; shape[1] = 8; shape[2] = 8; return BSG_OK;  // breakthrough.h:107-109
    case BSG_HEX: shape[0] = 9; shape[1] = 11; shape[2] = 11; return BSG_OK;         // hex.h:113-115
    case BSG_HEX5_HOST_ONLY: shape[0] = 9; shape[1] = 5; shape[2] = 5; return BSG_OK;
    case BSG_PENTAGO: shape[0] = 3; shape[1] = 6; shape[2] = 6; return BSG_OK;       // pentago.h:106-108
    case BSG_OTHELLO: shape[0] = 3; shape[1] = 8; shape[2] = 8; return BSG_OK;       // othello.h:139-141
    case BSG_Y: shape[0] = 3; shape[1] = 19; shape[2] = 19; return BSG_OK;           // y.h:189-191
    case BSG_Y9_HOST_ONLY: shape[0] = 3; shape[1] = 9; shape[2] = 9; return BSG_OK;
    case BSG_HAVANNAH: shape[0] = 3; shape[1] = 15; shape[2] = 15; return BSG_OK;     // havannah.h:209-211
    case BSG_HAVANNAH4_HOST_ONLY: shape[0] = 3; shape[1] = 7; shape[2] = 7; return BSG_OK;
    case BSG_GO: shape[0] = 4; shape[1] = 9; shape[2] = 9; return BSG_OK;               // go.h:141-145
    case BSG_GO7_HOST_ONLY: shape[0] = 4; shape[1] = 7; shape[2] = 7; return BSG_OK;
    default: return BSG_ERR_BAD_GAME;
  }
}

int bsg_pack_histories(int game, const int64_t* flat, const int64_t* off, int64_t n, void* out, int64_t cap) {
  if (!off || !out || n < 0 || cap < n) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, PackOp{flat, off, n, out, cap});
}
int bsg_host_state_info(int game, const void* p, int64_t cap, int64_t i, int* cur, int* term, int* ret0) {
  if (!p || i < 0 || i >= cap) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, InfoOp{p, cap, i, cur, term, ret0});
}
int bsg_host_legal_actions(int game, const void* p, int64_t cap, int64_t i, int64_t* out, int capacity) {
  if (!p || !out || i < 0 || i >= cap) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, LegalOp{p, cap, i, out, capacity});
}
int bsg_host_apply_action(int game, void* p, int64_t cap, int64_t i, int64_t a) {
  if (!p || i < 0 || i >= cap) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, ApplyOp{p, cap, i, a});
}
int bsg_host_observation(int game, const void* p, int64_t cap, int64_t i, int player, float* out) {
  if (!p || !out || i < 0 || i >= cap || player < 0 || player > 1) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, ObsOp{p, cap, i, player, out});
}
int bsg_host_legal_mask(int game, const void* p, int64_t cap, int64_t i, uint8_t* out) {
  if (!p || !out || i < 0 || i >= cap) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, MaskOp{p, cap, i, out});
}
int bsg_host_replay_step(int game, void* p, int64_t cap, int64_t i, uint32_t r, int64_t* action) {
  if (!p || i < 0 || i >= cap) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, StepOp{p, cap, i, r, action});
}
int bsg_host_fill_playout(int game, const void* p, int64_t cap, int64_t i, uint64_t seed, uint64_t playout,
                          int* ret0_out, int* moves_out) {
  if (!p || i < 0 || i >= cap) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, FillOp{p, cap, i, seed, playout, ret0_out, moves_out});
}
int bsg_host_to_string(int game, const void* p, int64_t cap, int64_t i, char* buf, int capacity) {
  if (!p || !buf || i < 0 || i >= cap || capacity < 1) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, ToStringOp{p, cap, i, 0, buf, capacity});
}
int bsg_host_serialize(int game, const void* p, int64_t cap, int64_t i, char* buf, int capacity) {
  if (!p || !buf || i < 0 || i >= cap || capacity < 1) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, ToStringOp{p, cap, i, 1, buf, capacity});
}
int bsg_host_deserialize(int game, const char* str, void* p, int64_t cap, int64_t i) {
  if (!p || !str || i < 0 || i >= cap) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, DeserializeOp{str, p, cap, i});
}
uint32_t bsg_draw(uint64_t seed, uint64_t playout, uint32_t t) {
  PhiloxKeys ks;
  philox_schedule(seed, ks);
  u32 out[4];
  philox4x32_10((u32)playout, (u32)(playout >> 32), t >> 2, 0u, ks, out);
  return out[t & 3];
}

int bsg_create(int game, int device, bsg_engine** out) {
  if (!out) return BSG_ERR_BAD_ARG;
  *out = nullptr;
  if (game < 0 || game == BSG_HEX5_HOST_ONLY || game == BSG_Y9_HOST_ONLY || game == BSG_HAVANNAH4_HOST_ONLY || game > BSG_GO)
    return BSG_ERR_BAD_GAME;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return BSG_ERR_NO_DEVICE;
  if (device < 0 || device >= count) return BSG_ERR_BAD_ARG;
  bsg_engine* e = new (std::nothrow) bsg_engine();
  if (!e) return BSG_ERR_BAD_ARG;
  e->game = game;
  e->device = device;
  *out = e;  // returned even on failure so bsg_last_error is readable; caller destroys it
  BSG_CUDA(e, cudaSetDevice(device));
  cudaDeviceProp prop;
  BSG_CUDA(e, cudaGetDeviceProperties(&prop, device));
  e->sm_count = prop.multiProcessorCount;
  BSG_CUDA(e, cudaStreamCreateWithFlags(&e->own_stream, cudaStreamNonBlocking));
  BSG_CUDA(e, cudaStreamCreateWithFlags(&e->aux_stream, cudaStreamNonBlocking));
  BSG_CUDA(e, cudaEventCreateWithFlags(&e->ev_a, cudaEventDisableTiming));
  BSG_CUDA(e, cudaEventCreateWithFlags(&e->ev_b, cudaEventDisableTiming));
  BSG_CUDA(e, cudaMalloc(&e->d_totals, 4 * sizeof(uint64_t)));
  BSG_CUDA(e, cudaMalloc(&e->d_work, kWorkSlots * sizeof(uint32_t)));
  return dispatch_device_game(game, CreateOp{e});
}

void bsg_destroy(bsg_engine* e) {
  if (!e) return;
  if (e->device >= 0) cudaSetDevice(e->device);
  if (e->comm && nccl_api().ok) nccl_api().CommDestroy(e->comm);
  if (e->tables_dev) cudaFree(e->tables_dev);
  if (e->d_roots) cudaFree(e->d_roots);
  if (e->d_out) cudaFree(e->d_out);
  if (e->d_fill) cudaFree(e->d_fill);
  if (e->d_scan) cudaFree(e->d_scan);
  if (e->h_pinned) cudaFreeHost(e->h_pinned);
  if (e->d_totals) cudaFree(e->d_totals);
  if (e->d_work) cudaFree(e->d_work);
  if (e->own_stream) cudaStreamDestroy(e->own_stream);
  if (e->aux_stream) cudaStreamDestroy(e->aux_stream);
  if (e->ev_a) cudaEventDestroy(e->ev_a);
  if (e->ev_b) cudaEventDestroy(e->ev_b);
  delete e;
}
const char* bsg_last_error(bsg_engine* e) { return e ? e->err.c_str() : "null engine"; }
int bsg_engine_game(const bsg_engine* e) { return e ? e->game : BSG_ERR_BAD_ARG; }
int bsg_engine_device(const bsg_engine* e) { return e ? e->device : BSG_ERR_BAD_ARG; }
int bsg_engine_sm_count(const bsg_engine* e) { return e ? e->sm_count : BSG_ERR_BAD_ARG; }
int bsg_engine_set_value_only(bsg_engine* e, int on) {
  if (!e) return BSG_ERR_BAD_ARG;
  if (on && dispatch_device_game(e->game, FillWordsOp{}) <= 0)
    return fail(e, BSG_ERR_BAD_ARG, "value_only needs a game whose full board has exactly one winner (hex, y)");
  e->value_only = on ? 1 : 0;
  return BSG_OK;
}

int bsg_playouts(bsg_engine* e, const void* roots, int64_t n_roots, int32_t rollouts, uint64_t seed,
                 uint64_t first_playout, const bsg_playout_io* io, void* stream) {
  if (!e) return BSG_ERR_BAD_ARG;
  if (!roots || n_roots < 0 || rollouts < 1) return fail(e, BSG_ERR_BAD_ARG, "bad roots/rollouts");
  int64_t total = n_roots * (int64_t)rollouts;
  if (total >= ((int64_t)1 << 32) - (1 << 24)) return fail(e, BSG_ERR_BAD_ARG, "more than 2^32-2^24 playouts in one launch");
  if (total == 0) return BSG_OK;
  PlayoutArgs a;
  memset(&a, 0, sizeof(a));
  a.roots = roots; a.n_roots = n_roots; a.n_playouts = (uint32_t)total; a.rollouts = (uint32_t)rollouts;
  a.first_playout = first_playout;
  philox_schedule(seed, a.keys);
  if (io) a.io = *io;
  if (a.io.r_table && a.io.r_stride < 1) return fail(e, BSG_ERR_BAD_ARG, "r_table needs r_stride >= 1");
  if (a.io.actions && a.io.actions_stride < 1) return fail(e, BSG_ERR_BAD_ARG, "actions needs actions_stride >= 1");
  if (a.io.state_log && ((!a.io.state_log_off && a.io.state_log_stride < 1) || a.io.state_log_cap < 1 ||
                         (a.io.state_log_off && a.io.state_log_stride != 0)))
    return fail(e, BSG_ERR_BAD_ARG, "state_log needs state_log_cap >= 1 and either state_log_off or state_log_stride");
  if (a.io.state_log && !a.io.state_log_off && (int64_t)total * a.io.state_log_stride > a.io.state_log_cap)
    return fail(e, BSG_ERR_BAD_ARG, "state_log_cap is smaller than playouts * state_log_stride");
  BSG_CUDA(e, cudaSetDevice(e->device));
  return dispatch_device_game(e->game, PlayoutOp{e, a, (cudaStream_t)stream});
}

int bsg_observations(bsg_engine* e, const void* states, int64_t n, const uint8_t* players, float* obs,
                     uint8_t* mask, void* stream) {
  if (!e) return BSG_ERR_BAD_ARG;
  if (!states || n < 0 || (!obs && !mask)) return fail(e, BSG_ERR_BAD_ARG, "bad observation arguments");
  if (n == 0) return BSG_OK;
  if (((uintptr_t)obs & 15) || ((uintptr_t)mask & 15)) return fail(e, BSG_ERR_BAD_ARG, "outputs must be 16-byte aligned");
  BSG_CUDA(e, cudaSetDevice(e->device));
  return dispatch_device_game(e->game, ObsOp2{e, states, n, players, obs, mask, (cudaStream_t)stream});
}

// Host-buffer playouts.  The roots are cut into chunks that ping-pong between two streams, so
// the H2D copy of chunk i+1 overlaps the playout kernel of chunk i (the copy is the longer
// leg: 16 B/root over PCIe against < 0.1 ns/root of kernel time).  Totals and per-root sums
// accumulate with integer atomics, so the result does not depend on the chunking.
int bsg_playouts_host(bsg_engine* e, const void* roots_host, int64_t n_roots, int32_t rollouts,
                      uint64_t seed, uint64_t first_playout, int64_t totals_out[4],
                      int32_t* root_ret_host, int32_t* root_steps_host) {
  if (!e) return BSG_ERR_BAD_ARG;
  if (!roots_host || n_roots < 1 || rollouts < 1 || !totals_out)
    return fail(e, BSG_ERR_BAD_ARG, "bad host playout arguments");
  if (n_roots * (int64_t)rollouts >= ((int64_t)1 << 32) - (1 << 24))
    return fail(e, BSG_ERR_BAD_ARG, "more than 2^32-2^24 playouts in one call");
  BSG_CUDA(e, cudaSetDevice(e->device));
  int plane_bytes[8];
  const int n_planes = dispatch_device_game(e->game, PlaneOp{plane_bytes});
  if (n_planes < 1) return fail(e, BSG_ERR_BAD_GAME, "no device kernels for this game");
  size_t in_bytes = (size_t)n_roots * (size_t)bsg_state_bytes(e->game);
  size_t out_elems = (root_ret_host ? n_roots : 0) + (root_steps_host ? n_roots : 0);
  int rc;
  if (in_bytes <= kSmallCallBytes && !e->value_only) {
    // Small call (one leaf, or a batch of leaves of an MCTS round): fixed cost is what matters.
    // One stream, no events, no pointer query (the roots are staged: copying <= 64 KB is
    // cheaper than asking), ONE memset for {totals, work counter, root sums}, ONE copy back.
    // Device layout of d_out here: [0,32) totals, [32,36) work counter, [64, ...) root sums.
    const size_t res_bytes = 64 + out_elems * 4;
    if ((rc = ensure(e, &e->d_roots, &e->d_roots_bytes, in_bytes, false)) != BSG_OK) return rc;
    if ((rc = ensure(e, &e->d_out, &e->d_out_bytes, res_bytes, false)) != BSG_OK) return rc;
    if ((rc = ensure(e, &e->h_pinned, &e->h_pinned_bytes, res_bytes + in_bytes + 64, true)) != BSG_OK) return rc;
    char* h_res = (char*)e->h_pinned;
    char* stage = h_res + ((res_bytes + 63) & ~(size_t)63);
    memcpy(stage, roots_host, in_bytes);
    cudaStream_t st = e->own_stream;
    BSG_CUDA(e, cudaMemcpyAsync(e->d_roots, stage, in_bytes, cudaMemcpyHostToDevice, st));
    BSG_CUDA(e, cudaMemsetAsync(e->d_out, 0, res_bytes, st));
    PlayoutArgs a;
    memset(&a, 0, sizeof(a));
    a.roots = e->d_roots; a.n_roots = n_roots; a.root0 = 0;
    a.n_playouts = (uint32_t)(n_roots * rollouts); a.rollouts = (uint32_t)rollouts;
    a.first_playout = first_playout;
    philox_schedule(seed, a.keys);
    a.io.totals = (uint64_t*)e->d_out;
    a.work_counter = (uint32_t*)((char*)e->d_out + 32);
    int32_t* d_sums = (int32_t*)((char*)e->d_out + 64);
    a.io.root_sum_ret0 = root_ret_host ? d_sums : nullptr;
    a.io.root_sum_steps = root_steps_host ? d_sums + (root_ret_host ? n_roots : 0) : nullptr;
    if ((rc = dispatch_device_game(e->game, PlayoutOp{e, a, st})) != BSG_OK) return rc;
    BSG_CUDA(e, cudaMemcpyAsync(h_res, e->d_out, res_bytes, cudaMemcpyDeviceToHost, st));
    BSG_CUDA(e, cudaStreamSynchronize(st));
    const uint64_t* h_tot = (const uint64_t*)h_res;
    for (int i = 0; i < 4; ++i) totals_out[i] = (int64_t)h_tot[i];
    if (root_ret_host) memcpy(root_ret_host, h_res + 64, (size_t)n_roots * 4);
    if (root_steps_host) memcpy(root_steps_host, h_res + 64 + (root_ret_host ? (size_t)n_roots * 4 : 0), (size_t)n_roots * 4);
    if (totals_out[3] != 0) return fail(e, BSG_ERR_STUCK_LANE, "non-terminal state without legal actions");
    return BSG_OK;
  }
  if ((rc = ensure(e, &e->d_roots, &e->d_roots_bytes, in_bytes, false)) != BSG_OK) return rc;
  if ((rc = ensure(e, &e->d_out, &e->d_out_bytes, out_elems * 4 + 16, false)) != BSG_OK) return rc;
  // Page-locked caller memory (e.g. a torch pin_memory() tensor) is copied from directly;
  // pageable memory is staged through the engine's pinned buffer.
  cudaPointerAttributes attr;
  bool pinned_src = cudaPointerGetAttributes(&attr, roots_host) == cudaSuccess &&
                    attr.type == cudaMemoryTypeHost;
  cudaGetLastError();
  size_t pin_need = (pinned_src ? 0 : in_bytes) + out_elems * 4 + 128;
  if ((rc = ensure(e, &e->h_pinned, &e->h_pinned_bytes, pin_need, true)) != BSG_OK) return rc;
  char* h_res = (char*)e->h_pinned;                      // {totals[4], root sums...}
  const char* src = (const char*)roots_host;
  if (!pinned_src) {
    char* stage = h_res + ((out_elems * 4 + 32 + 63) & ~(size_t)63);
    memcpy(stage, roots_host, in_bytes);
    src = stage;
  }
  cudaStream_t sa = e->own_stream, sb = e->aux_stream;
  BSG_CUDA(e, cudaMemsetAsync(e->d_totals, 0, 4 * sizeof(uint64_t), sa));
  if (out_elems) BSG_CUDA(e, cudaMemsetAsync(e->d_out, 0, out_elems * 4, sa));
  BSG_CUDA(e, cudaEventRecord(e->ev_a, sa));
  BSG_CUDA(e, cudaStreamWaitEvent(sb, e->ev_a, 0));
  int32_t* d_ret = root_ret_host ? (int32_t*)e->d_out : nullptr;
  int32_t* d_steps = root_steps_host ? (int32_t*)e->d_out + (root_ret_host ? n_roots : 0) : nullptr;
  // chunks of >= 2 MiB of roots, at most 16; a small call stays one chunk
  int64_t n_chunks = (int64_t)(in_bytes >> 21);
  n_chunks = n_chunks < 1 ? 1 : (n_chunks > 16 ? 16 : n_chunks);
  PhiloxKeys keys;
  philox_schedule(seed, keys);
  if (e->value_only) {
    const int words = dispatch_device_game(e->game, FillWordsOp{});
    if ((rc = ensure(e, &e->d_fill, &e->d_fill_bytes, (size_t)n_roots * rollouts * words * 4, false)) != BSG_OK) return rc;
  }
  for (int64_t c = 0; c < n_chunks; ++c) {
    const int64_t lo = n_roots * c / n_chunks, hi = n_roots * (c + 1) / n_chunks;
    if (hi <= lo) continue;
    cudaStream_t st = (c & 1) ? sb : sa;
    size_t plane_off = 0;
    for (int p = 0; p < n_planes; ++p) {
      size_t pb = (size_t)plane_bytes[p];
      BSG_CUDA(e, cudaMemcpyAsync((char*)e->d_roots + plane_off + (size_t)lo * pb,
                                  src + plane_off + (size_t)lo * pb, (size_t)(hi - lo) * pb,
                                  cudaMemcpyHostToDevice, st));
      plane_off += (size_t)n_roots * pb;
    }
    PlayoutArgs a;
    memset(&a, 0, sizeof(a));
    a.roots = e->d_roots; a.n_roots = n_roots; a.root0 = lo;
    a.n_playouts = (uint32_t)((hi - lo) * rollouts); a.rollouts = (uint32_t)rollouts;
    a.first_playout = first_playout + (uint64_t)lo * (uint64_t)rollouts;
    a.keys = keys;
    a.io.totals = e->d_totals;
    a.io.root_sum_ret0 = d_ret ? d_ret + lo : nullptr;
    a.io.root_sum_steps = d_steps ? d_steps + lo : nullptr;
    a.io.value_only = e->value_only;
    a.fill_base = (uint64_t)lo * (uint64_t)rollouts;
    if ((rc = dispatch_device_game(e->game, PlayoutOp{e, a, st})) != BSG_OK) return rc;
  }
  BSG_CUDA(e, cudaEventRecord(e->ev_b, sb));
  BSG_CUDA(e, cudaStreamWaitEvent(sa, e->ev_b, 0));
  uint64_t* h_tot = (uint64_t*)h_res;
  BSG_CUDA(e, cudaMemcpyAsync(h_tot, e->d_totals, 32, cudaMemcpyDeviceToHost, sa));
  if (out_elems) BSG_CUDA(e, cudaMemcpyAsync(h_res + 32, e->d_out, out_elems * 4, cudaMemcpyDeviceToHost, sa));
  BSG_CUDA(e, cudaStreamSynchronize(sa));
  for (int i = 0; i < 4; ++i) totals_out[i] = (int64_t)h_tot[i];
  if (root_ret_host) memcpy(root_ret_host, h_res + 32, (size_t)n_roots * 4);
  if (root_steps_host) memcpy(root_steps_host, h_res + 32 + (root_ret_host ? (size_t)n_roots * 4 : 0), (size_t)n_roots * 4);
  if (totals_out[3] != 0) return fail(e, BSG_ERR_STUCK_LANE, "non-terminal state without legal actions");
  return BSG_OK;
}

// Host-buffer playouts that return PER-PLAYOUT results — (int)Returns()[0] and the move count,
// 2 bytes per playout — instead of totals only: what a caller that
// consumes every root's value (evaluate_batch at rollouts = 1, a data pipeline) needs.  Same
// two-stream chunking as bsg_playouts_host; each chunk's results are copied back in its own
// stream right behind its kernel, so the D2H traffic rides under the next chunk's H2D copy (PCIe
// is full duplex).  Page-locked caller buffers are read and written in place.
int bsg_playouts_host_results(bsg_engine* e, const void* roots_host, int64_t n_roots, int32_t rollouts,
                              uint64_t seed, uint64_t first_playout, int64_t totals_out[4],
                              int8_t* ret0_host, uint8_t* length_host) {
  if (!e) return BSG_ERR_BAD_ARG;
  if (!roots_host || n_roots < 1 || rollouts < 1 || !totals_out || (!ret0_host && !length_host))
    return fail(e, BSG_ERR_BAD_ARG, "bad host playout arguments");
  if (e->value_only) return fail(e, BSG_ERR_BAD_ARG, "per-playout results need the exact kernels (value_only is set)");
  const int64_t n_play = n_roots * (int64_t)rollouts;
  if (n_play >= ((int64_t)1 << 32) - (1 << 24))
    return fail(e, BSG_ERR_BAD_ARG, "more than 2^32-2^24 playouts in one call");
  BSG_CUDA(e, cudaSetDevice(e->device));
  int plane_bytes[8];
  const int n_planes = dispatch_device_game(e->game, PlaneOp{plane_bytes});
  if (n_planes < 1) return fail(e, BSG_ERR_BAD_GAME, "no device kernels for this game");
  const size_t in_bytes = (size_t)n_roots * (size_t)bsg_state_bytes(e->game);
  const size_t out_bytes = (size_t)n_play * ((ret0_host ? 1 : 0) + (length_host ? 1 : 0));
  int rc;
  if ((rc = ensure(e, &e->d_roots, &e->d_roots_bytes, in_bytes, false)) != BSG_OK) return rc;
  if ((rc = ensure(e, &e->d_out, &e->d_out_bytes, out_bytes + 64, false)) != BSG_OK) return rc;
  auto is_pinned = [](const void* p) {
    cudaPointerAttributes attr;
    bool pinned = cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    return pinned;
  };
  const bool pin_in = is_pinned(roots_host);
  const bool pin_out = (!ret0_host || is_pinned(ret0_host)) && (!length_host || is_pinned(length_host));
  const size_t pin_need = 128 + (pin_in ? 0 : in_bytes + 64) + (pin_out ? 0 : out_bytes + 64);
  if ((rc = ensure(e, &e->h_pinned, &e->h_pinned_bytes, pin_need, true)) != BSG_OK) return rc;
  char* h_res = (char*)e->h_pinned;
  char* h_in = h_res + 128;
  char* h_out = h_in + (pin_in ? 0 : ((in_bytes + 63) & ~(size_t)63));
  const char* src = (const char*)roots_host;
  if (!pin_in) { memcpy(h_in, roots_host, in_bytes); src = h_in; }
  int8_t* h_ret = ret0_host;
  uint8_t* h_len = length_host;
  if (!pin_out) {
    h_ret = ret0_host ? (int8_t*)h_out : nullptr;
    h_len = length_host ? (uint8_t*)h_out + (ret0_host ? n_play : 0) : nullptr;
  }
  int8_t* d_ret = ret0_host ? (int8_t*)e->d_out : nullptr;
  uint8_t* d_len = length_host ? (uint8_t*)e->d_out + (ret0_host ? n_play : 0) : nullptr;
  cudaStream_t sa = e->own_stream, sb = e->aux_stream;
  BSG_CUDA(e, cudaMemsetAsync(e->d_totals, 0, 4 * sizeof(uint64_t), sa));
  BSG_CUDA(e, cudaEventRecord(e->ev_a, sa));
  BSG_CUDA(e, cudaStreamWaitEvent(sb, e->ev_a, 0));
  int64_t n_chunks = (int64_t)(in_bytes >> 21);
  n_chunks = n_chunks < 1 ? 1 : (n_chunks > 16 ? 16 : n_chunks);
  PhiloxKeys keys;
  philox_schedule(seed, keys);
  for (int64_t c = 0; c < n_chunks; ++c) {
    const int64_t lo = n_roots * c / n_chunks, hi = n_roots * (c + 1) / n_chunks;
    if (hi <= lo) continue;
    cudaStream_t st = (c & 1) ? sb : sa;
    size_t plane_off = 0;
    for (int p = 0; p < n_planes; ++p) {
      const size_t pb = (size_t)plane_bytes[p];
      BSG_CUDA(e, cudaMemcpyAsync((char*)e->d_roots + plane_off + (size_t)lo * pb, src + plane_off + (size_t)lo * pb,
                                  (size_t)(hi - lo) * pb, cudaMemcpyHostToDevice, st));
      plane_off += (size_t)n_roots * pb;
    }
    const int64_t p_lo = lo * rollouts, p_n = (hi - lo) * rollouts;
    PlayoutArgs a;
    memset(&a, 0, sizeof(a));
    a.roots = e->d_roots; a.n_roots = n_roots; a.root0 = lo;
    a.n_playouts = (uint32_t)p_n; a.rollouts = (uint32_t)rollouts;
    a.first_playout = first_playout + (uint64_t)p_lo;
    a.keys = keys;
    a.io.totals = e->d_totals;
    a.io.ret0 = d_ret ? d_ret + p_lo : nullptr;
    a.io.length = d_len ? d_len + p_lo : nullptr;
    if ((rc = dispatch_device_game(e->game, PlayoutOp{e, a, st})) != BSG_OK) return rc;
    if (d_ret) BSG_CUDA(e, cudaMemcpyAsync(h_ret + p_lo, d_ret + p_lo, (size_t)p_n, cudaMemcpyDeviceToHost, st));
    if (d_len) BSG_CUDA(e, cudaMemcpyAsync(h_len + p_lo, d_len + p_lo, (size_t)p_n, cudaMemcpyDeviceToHost, st));
  }
  BSG_CUDA(e, cudaEventRecord(e->ev_b, sb));
  BSG_CUDA(e, cudaStreamWaitEvent(sa, e->ev_b, 0));
  uint64_t* h_tot = (uint64_t*)h_res;
  BSG_CUDA(e, cudaMemcpyAsync(h_tot, e->d_totals, 32, cudaMemcpyDeviceToHost, sa));
  BSG_CUDA(e, cudaStreamSynchronize(sa));
  for (int i = 0; i < 4; ++i) totals_out[i] = (int64_t)h_tot[i];
  if (!pin_out) {
    if (ret0_host) memcpy(ret0_host, h_ret, (size_t)n_play);
    if (length_host) memcpy(length_host, h_len, (size_t)n_play);
  }
  if (totals_out[3] != 0) return fail(e, BSG_ERR_STUCK_LANE, "non-terminal state without legal actions");
  return BSG_OK;
}

// ---- compact roots (connect_four: 7 bytes per root)
}  // extern "C"
namespace {
struct CompactBitsOp {
  template <class G> int operator()() const { return compact_bits<G>::value; }
};
struct CompactConvertOp {
  const void* in; void* out; int64_t n; int to_compact;
  template <class G> int operator()() const {
    if constexpr (compact_bytes<G>::value > 0) {
      // expanding reads three words per root: work on a padded copy of the compact buffer
      std::vector<unsigned char> padded;
      const void* src = in;
      const size_t cbytes = ((size_t)n * compact_bits<G>::value + 7) / 8;
      if (!to_compact) {
        padded.assign(cbytes + 16, 0);
        memcpy(padded.data(), in, cbytes);
        src = padded.data();
      } else {
        memset(out, 0, cbytes);                  // store_compact ORs its bits in
      }
      for (int64_t i = 0; i < n; ++i) {
        typename G::State s;
        if (to_compact) { G::load(s, src, n, i); compact_ops<G, true>::store(s, out, i); }
        else { compact_ops<G, true>::load(s, src, i); G::store(s, out, n, i); }
      }
      return BSG_OK;
    } else {
      memcpy(out, in, (size_t)n * G::kStateBytes);   // no compact format: the packed one
      return BSG_OK;
    }
  }
};
}  // namespace
extern "C" {
int bsg_compact_state_bits(int game) { return dispatch_game(game, CompactBitsOp{}); }
int64_t bsg_compact_buffer_bytes(int game, int64_t n) {
  const int bits = bsg_compact_state_bits(game);
  return bits < 0 || n < 0 ? (int64_t)BSG_ERR_BAD_ARG : (n * bits + 7) / 8;
}
int bsg_compact_roots(int game, const void* packed_host, int64_t n, void* compact_out) {
  if (!packed_host || !compact_out || n < 0) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, CompactConvertOp{packed_host, compact_out, n, 1});
}
int bsg_expand_roots(int game, const void* compact_host, int64_t n, void* packed_out) {
  if (!compact_host || !packed_out || n < 0) return BSG_ERR_BAD_ARG;
  return dispatch_game(game, CompactConvertOp{compact_host, packed_out, n, 0});
}

// bsg_playouts_host_results for roots in the compact format: the same two-stream chunking, an
// eighth less to copy for connect_four.  ret0_host / length_host may both be NULL (totals only).
int bsg_playouts_host_compact(bsg_engine* e, const void* roots_compact_host, int64_t n_roots, int32_t rollouts,
                              uint64_t seed, uint64_t first_playout, int64_t totals_out[4],
                              int8_t* ret0_host, uint8_t* length_host) {
  if (!e) return BSG_ERR_BAD_ARG;
  if (!roots_compact_host || n_roots < 1 || rollouts < 1 || !totals_out)
    return fail(e, BSG_ERR_BAD_ARG, "bad host playout arguments");
  if (e->value_only) return fail(e, BSG_ERR_BAD_ARG, "compact roots run the exact kernels (value_only is set)");
  const int bits = bsg_compact_state_bits(e->game);
  if (bits == 8 * bsg_state_bytes(e->game))   // no compact format for this game: the packed path
    return (ret0_host || length_host)
               ? bsg_playouts_host_results(e, roots_compact_host, n_roots, rollouts, seed, first_playout, totals_out, ret0_host, length_host)
               : bsg_playouts_host(e, roots_compact_host, n_roots, rollouts, seed, first_playout, totals_out, nullptr, nullptr);
  const int64_t n_play = n_roots * (int64_t)rollouts;
  if (n_play >= ((int64_t)1 << 32) - (1 << 24))
    return fail(e, BSG_ERR_BAD_ARG, "more than 2^32-2^24 playouts in one call");
  BSG_CUDA(e, cudaSetDevice(e->device));
  const size_t in_bytes = ((size_t)n_roots * (size_t)bits + 7) / 8;
  const size_t out_bytes = (size_t)n_play * ((ret0_host ? 1 : 0) + (length_host ? 1 : 0));
  int rc;
  if ((rc = ensure(e, &e->d_roots, &e->d_roots_bytes, in_bytes + 16, false)) != BSG_OK) return rc;
  if ((rc = ensure(e, &e->d_out, &e->d_out_bytes, out_bytes + 64, false)) != BSG_OK) return rc;
  auto is_pinned = [](const void* p) {
    cudaPointerAttributes attr;
    bool pinned = cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    return pinned;
  };
  const bool pin_in = is_pinned(roots_compact_host);
  const bool pin_out = (!ret0_host || is_pinned(ret0_host)) && (!length_host || is_pinned(length_host));
  const size_t pin_need = 128 + (pin_in ? 0 : in_bytes + 64) + (pin_out ? 0 : out_bytes + 64);
  if ((rc = ensure(e, &e->h_pinned, &e->h_pinned_bytes, pin_need, true)) != BSG_OK) return rc;
  char* h_res = (char*)e->h_pinned;
  char* h_in = h_res + 128;
  char* h_out = h_in + (pin_in ? 0 : ((in_bytes + 63) & ~(size_t)63));
  const char* src = (const char*)roots_compact_host;
  if (!pin_in) { memcpy(h_in, roots_compact_host, in_bytes); src = h_in; }
  int8_t* h_ret = ret0_host;
  uint8_t* h_len = length_host;
  if (!pin_out) {
    h_ret = ret0_host ? (int8_t*)h_out : nullptr;
    h_len = length_host ? (uint8_t*)h_out + (ret0_host ? n_play : 0) : nullptr;
  }
  int8_t* d_ret = ret0_host ? (int8_t*)e->d_out : nullptr;
  uint8_t* d_len = length_host ? (uint8_t*)e->d_out + (ret0_host ? n_play : 0) : nullptr;
  cudaStream_t sa = e->own_stream, sb = e->aux_stream;
  BSG_CUDA(e, cudaMemsetAsync(e->d_totals, 0, 4 * sizeof(uint64_t), sa));
  BSG_CUDA(e, cudaEventRecord(e->ev_a, sa));
  BSG_CUDA(e, cudaStreamWaitEvent(sb, e->ev_a, 0));
  int64_t n_chunks = (int64_t)(in_bytes >> 21);
  n_chunks = n_chunks < 1 ? 1 : (n_chunks > 16 ? 16 : n_chunks);
  PhiloxKeys keys;
  philox_schedule(seed, keys);
  for (int64_t c = 0; c < n_chunks; ++c) {
    // chunk boundaries at multiples of 32 roots: whatever the bits per root, a chunk then starts
    // on a word of the bit string (the device copy sits at the same offsets as the host buffer)
    int64_t lo = (n_roots * c / n_chunks) & ~(int64_t)31, hi = c + 1 == n_chunks ? n_roots : ((n_roots * (c + 1) / n_chunks) & ~(int64_t)31);
    if (hi <= lo) continue;
    cudaStream_t st = (c & 1) ? sb : sa;
    const size_t b_lo = (size_t)lo * bits / 8, b_hi = c + 1 == n_chunks ? in_bytes : (size_t)hi * bits / 8;
    BSG_CUDA(e, cudaMemcpyAsync((char*)e->d_roots + b_lo, src + b_lo, b_hi - b_lo, cudaMemcpyHostToDevice, st));
    const int64_t p_lo = lo * rollouts, p_n = (hi - lo) * rollouts;
    PlayoutArgs a;
    memset(&a, 0, sizeof(a));
    a.roots = e->d_roots; a.n_roots = n_roots; a.root0 = lo; a.compact = 1;
    a.n_playouts = (uint32_t)p_n; a.rollouts = (uint32_t)rollouts;
    a.first_playout = first_playout + (uint64_t)p_lo;
    a.keys = keys;
    a.io.totals = e->d_totals;
    a.io.ret0 = d_ret ? d_ret + p_lo : nullptr;
    a.io.length = d_len ? d_len + p_lo : nullptr;
    if ((rc = dispatch_device_game(e->game, PlayoutOp{e, a, st})) != BSG_OK) return rc;
    if (d_ret) BSG_CUDA(e, cudaMemcpyAsync(h_ret + p_lo, d_ret + p_lo, (size_t)p_n, cudaMemcpyDeviceToHost, st));
    if (d_len) BSG_CUDA(e, cudaMemcpyAsync(h_len + p_lo, d_len + p_lo, (size_t)p_n, cudaMemcpyDeviceToHost, st));
  }
  BSG_CUDA(e, cudaEventRecord(e->ev_b, sb));
  BSG_CUDA(e, cudaStreamWaitEvent(sa, e->ev_b, 0));
  uint64_t* h_tot = (uint64_t*)h_res;
  BSG_CUDA(e, cudaMemcpyAsync(h_tot, e->d_totals, 32, cudaMemcpyDeviceToHost, sa));
  BSG_CUDA(e, cudaStreamSynchronize(sa));
  for (int i = 0; i < 4; ++i) totals_out[i] = (int64_t)h_tot[i];
  if (!pin_out) {
    if (ret0_host) memcpy(ret0_host, h_ret, (size_t)n_play);
    if (length_host) memcpy(length_host, h_len, (size_t)n_play);
  }
  if (totals_out[3] != 0) return fail(e, BSG_ERR_STUCK_LANE, "non-terminal state without legal actions");
  return BSG_OK;
}

int bsg_observations_host(bsg_engine* e, const void* states_host, int64_t n, const uint8_t* players_host,
                          float* obs_host, uint8_t* mask_host) {
  if (!e) return BSG_ERR_BAD_ARG;
  if (!states_host || n < 1 || (!obs_host && !mask_host)) return fail(e, BSG_ERR_BAD_ARG, "bad host observation arguments");
  BSG_CUDA(e, cudaSetDevice(e->device));
  size_t sb = (size_t)n * (size_t)bsg_state_bytes(e->game);
  size_t in_bytes = ((sb + 15) & ~(size_t)15) + (players_host ? (((size_t)n + 15) & ~(size_t)15) : 0);
  size_t obs_bytes = obs_host ? (((size_t)n * bsg_observation_size(e->game) * 4 + 15) & ~(size_t)15) : 0;
  size_t mask_bytes = mask_host ? (((size_t)n * bsg_num_distinct_actions(e->game) + 15) & ~(size_t)15) : 0;
  int rc;
  if ((rc = ensure(e, &e->d_roots, &e->d_roots_bytes, in_bytes, false)) != BSG_OK) return rc;
  if ((rc = ensure(e, &e->d_out, &e->d_out_bytes, obs_bytes + mask_bytes, false)) != BSG_OK) return rc;
  cudaStream_t st = e->own_stream;
  BSG_CUDA(e, cudaMemcpyAsync(e->d_roots, states_host, sb, cudaMemcpyHostToDevice, st));
  uint8_t* d_players = nullptr;
  if (players_host) {
    d_players = (uint8_t*)e->d_roots + ((sb + 15) & ~(size_t)15);
    BSG_CUDA(e, cudaMemcpyAsync(d_players, players_host, (size_t)n, cudaMemcpyHostToDevice, st));
  }
  float* d_obs = obs_host ? (float*)e->d_out : nullptr;
  uint8_t* d_mask = mask_host ? (uint8_t*)e->d_out + obs_bytes : nullptr;
  if ((rc = bsg_observations(e, e->d_roots, n, d_players, d_obs, d_mask, st)) != BSG_OK) return rc;
  if (obs_host) BSG_CUDA(e, cudaMemcpyAsync(obs_host, d_obs, (size_t)n * bsg_observation_size(e->game) * 4, cudaMemcpyDeviceToHost, st));
  if (mask_host) BSG_CUDA(e, cudaMemcpyAsync(mask_host, d_mask, (size_t)n * bsg_num_distinct_actions(e->game), cudaMemcpyDeviceToHost, st));
  BSG_CUDA(e, cudaStreamSynchronize(st));
  return BSG_OK;
}

// ---- native MCTS tree (bsg_mcts.cuh)
struct bsg_mcts { bsg::MctsBase* t = nullptr; };

int bsg_mcts_create(bsg_engine* e, double uct_c, int solve, int n_rollouts, int leaf_batch,
                    uint64_t eval_seed, uint64_t shuffle_seed, bsg_mcts** out) {
  if (!e || !out || n_rollouts < 1 || leaf_batch < 1) return BSG_ERR_BAD_ARG;
  *out = nullptr;
  bsg::MctsBase* t = nullptr;
  int rc = dispatch_device_game(e->game, MctsCreateOp{&t});
  if (rc != BSG_OK) return rc;
  t->engine = e; t->uct_c = uct_c; t->solve = solve; t->n_rollouts = n_rollouts;
  t->leaf_batch = leaf_batch; t->eval_seed = eval_seed; t->shuffle_seed = shuffle_seed;
  bsg_mcts* m = new (std::nothrow) bsg_mcts();
  if (!m) { delete t; return BSG_ERR_BAD_ARG; }
  m->t = t;
  *out = m;
  return BSG_OK;
}
void bsg_mcts_destroy(bsg_mcts* m) {
  if (!m) return;
  delete m->t;
  delete m;
}
const char* bsg_mcts_last_error(bsg_mcts* m) {
  return (m && m->t) ? (m->t->err.empty() ? bsg_last_error(m->t->engine) : m->t->err.c_str()) : "null tree";
}
int bsg_mcts_search(bsg_mcts* m, const void* root, int64_t simulations, int resume) {
  if (!m || !m->t || !root || simulations < 0) return BSG_ERR_BAD_ARG;
  return m->t->search(root, simulations, resume);
}
int bsg_mcts_expand_root(bsg_mcts* m, const void* root) {
  if (!m || !m->t || !root) return BSG_ERR_BAD_ARG;
  return m->t->expand_root(root);
}
int bsg_mcts_root_children(bsg_mcts* m, int capacity, int64_t* actions, int64_t* visits,
                           double* reward, int32_t* outcome) {
  if (!m || !m->t || m->t->nodes.empty()) return BSG_ERR_BAD_ARG;
  const bsg::MctsNode& root = m->t->nodes[0];
  for (int i = 0; i < root.n_children && i < capacity; ++i) {
    const bsg::MctsNode& c = m->t->nodes[root.first_child + i];
    if (actions) actions[i] = c.action;
    if (visits) visits[i] = c.explore;
    if (reward) reward[i] = c.reward;
    if (outcome) outcome[i] = c.has_outcome ? (int32_t)(c.player == 0 ? c.out0 : -c.out0) : 2;
  }
  return root.n_children;
}
int bsg_mcts_info(bsg_mcts* m, double info[8]) {
  if (!m || !m->t || !info || m->t->nodes.empty()) return BSG_ERR_BAD_ARG;
  const bsg::MctsNode& root = m->t->nodes[0];
  info[0] = (double)root.explore; info[1] = root.has_outcome; info[2] = root.out0;
  info[3] = (double)m->t->nodes.size(); info[4] = (double)m->t->total_game_steps;
  info[5] = (double)m->t->kernel_launches; info[6] = (double)m->t->next_playout;
  info[7] = (double)m->t->root_player;
  return BSG_OK;
}
int bsg_mcts_set_root_child(bsg_mcts* m, int64_t action, int64_t visits, double reward, int32_t outcome) {
  if (!m || !m->t || m->t->nodes.empty()) return BSG_ERR_BAD_ARG;
  const bsg::MctsNode& root = m->t->nodes[0];
  for (int i = 0; i < root.n_children; ++i) {
    bsg::MctsNode& c = m->t->nodes[root.first_child + i];
    if (c.action != action) continue;
    c.explore = visits;
    c.reward = reward;
    if (outcome != 2) { c.has_outcome = 1; c.out0 = c.player == 0 ? (double)outcome : -(double)outcome; }
    return BSG_OK;
  }
  return BSG_ERR_ILLEGAL_ACTION;
}
int bsg_mcts_set_root_explore(bsg_mcts* m, int64_t n) {
  if (!m || !m->t || m->t->nodes.empty() || n < 0) return BSG_ERR_BAD_ARG;
  m->t->nodes[0].explore = n;
  return BSG_OK;
}
int bsg_mcts_set_next_playout(bsg_mcts* m, uint64_t next) {
  if (!m || !m->t) return BSG_ERR_BAD_ARG;
  m->t->next_playout = next;
  return BSG_OK;
}


// ---- GPU-resident search forest -------------------------------------------------------------
}  // extern "C"

struct bsg_forest {
  bsg_engine* e = nullptr;
  bsg::ForestView F;
  std::vector<void*> allocs;
  void* d_root = nullptr;
  int64_t* d_per_tree = nullptr;
  uint64_t* d_totals = nullptr;
  uint64_t eval_seed = 0;
  int64_t rounds = 0, launches = 0, game_steps = 0;
  // bsg_forest_merge: dense per-action buffers {base [A][2], summed [A][5]} and the CUDA events
  // around every all-reduce (read by bsg_forest_allreduce_stats)
  int64_t* d_base = nullptr;
  int64_t* d_merge = nullptr;
  uint64_t* d_step_totals = nullptr;    // {playouts, game steps, sum ret0, stuck lanes} of all searches so far
  int num_actions = 0;
  std::vector<cudaEvent_t> ar_events;   // pairs, in issue order
  size_t ar_used = 0;
  int64_t merges = 0;
  std::string err;
  ~bsg_forest() {
    for (void* p : allocs) cudaFree(p);
    for (cudaEvent_t ev : ar_events) cudaEventDestroy(ev);
  }
};

namespace {
struct ForestGeomOp {
  int* depth;
  template <class G> int operator()() const { *depth = G::kMaxDepth + 2; return BSG_OK; }
};
template <class T> int forest_alloc(bsg_forest* f, T** p, size_t count) {
  void* q = nullptr;
  cudaError_t st = cudaMalloc(&q, count * sizeof(T) + 16);
  if (st != cudaSuccess) { f->err = std::string("cudaMalloc: ") + cudaGetErrorString(st); return BSG_ERR_CUDA; }
  cudaMemset(q, 0, count * sizeof(T) + 16);
  f->allocs.push_back(q);
  *p = (T*)q;
  return BSG_OK;
}
int forest_grid(const bsg_forest* f) { return (f->F.n_trees + kForestThreads - 1) / kForestThreads; }

// Test hook: the forest's per-tree code on host memory against the host tree, both with the
// stand-in leaf values of MctsBase::fake_eval.  Returns 0 when every tree is identical node for
// node, 1 + the first differing tree otherwise.
struct ForestSelfTestOp {
  int n_trees, sims, sims2, leaf_batch, n_rollouts, solve; uint64_t seed; const int64_t* hist; int n_hist;
  template <class G> int operator()() const {
    typedef typename G::State State;
    State root;
    G::init(root);
    for (int i = 0; i < n_hist; ++i)
      if (!apply_any<G>(root, (int)hist[i])) return BSG_ERR_ILLEGAL_ACTION;
    std::vector<unsigned char> rootbuf((size_t)G::kStateBytes + 64, 0);
    G::store(root, rootbuf.data(), 1, 0);
    const size_t T = (size_t)n_trees, B = (size_t)leaf_batch;
    const int cap = (sims + sims2 + 2) * (G::kNumActions < 64 ? G::kNumActions : 64) + 2 * G::kNumActions + 8;
    const int depth = G::kMaxDepth + 2;
    std::vector<bsg::MctsNode> nodes(T * (size_t)cap);
    std::vector<int32_t> n_nodes(T), paths(T * B * (size_t)depth), path_len(T * B), leaf_kind(T * B), n_pending(T), leaf_ret(T * B);
    std::vector<uint64_t> shuffle_ctr(T, 0), next_playout(T), leaf_base(T * B);
    std::vector<int64_t> sims_left(T);
    std::vector<double> leaf_out0(T * B), logtab((size_t)(sims + sims2 + leaf_batch + 4));
    std::vector<unsigned char> leaf_states(T * B * (size_t)G::kStateBytes + 64);
    int32_t error = 0; uint32_t active = 0;
    for (size_t i = 1; i < logtab.size(); ++i) logtab[i] = log((double)i);
    bsg::ForestView F;
    memset(&F, 0, sizeof(F));
    F.nodes = nodes.data(); F.n_nodes = n_nodes.data(); F.shuffle_ctr = shuffle_ctr.data();
    F.next_playout = next_playout.data(); F.sims_left = sims_left.data(); F.paths = paths.data();
    F.path_len = path_len.data(); F.leaf_kind = leaf_kind.data(); F.leaf_out0 = leaf_out0.data();
    F.n_pending = n_pending.data(); F.leaf_states = leaf_states.data(); F.leaf_base = leaf_base.data();
    F.leaf_ret = leaf_ret.data(); F.logtab = logtab.data(); F.error = &error; F.active = &active;
    F.root = rootbuf.data(); F.shuffle_seed0 = seed + 1000; F.n_trees = n_trees; F.cap = cap;
    F.batch = leaf_batch; F.depth = depth; F.n_rollouts = n_rollouts; F.solve = solve;
    F.logtab_n = (int32_t)logtab.size(); F.uct_c = 2.0;
    const uint64_t first = 12345;
    const typename G::Tables& tab = TablesOf<G>::get();
    for (int v = 0; v < n_trees; ++v) { next_playout[(size_t)v] = first + ((uint64_t)v << 40); bsg::ForestTree<G>::reset(F, v); }
    std::vector<bsg::MctsTree<G>> ref((size_t)n_trees);
    for (int v = 0; v < n_trees; ++v) {
      bsg::MctsTree<G>& t = ref[(size_t)v];
      t.fake_eval = true; t.uct_c = 2.0; t.solve = solve; t.n_rollouts = n_rollouts; t.leaf_batch = leaf_batch;
      t.eval_seed = seed; t.shuffle_seed = seed + 1000 + (uint64_t)v; t.next_playout = first + ((uint64_t)v << 40);
    }
    for (int phase = 0; phase < 2; ++phase) {
      const int budget = phase == 0 ? sims : sims2;
      if (phase == 1 && budget == 0) break;
      for (int v = 0; v < n_trees; ++v) sims_left[(size_t)v] = budget + (phase == 0 ? v % 3 : 0);   // uneven shares
      for (;;) {
        active = 0;
        for (int v = 0; v < n_trees; ++v) {
          if (sims_left[(size_t)v] > 0) ++active;
          bsg::ForestTree<G>::descend(F, v, tab);
        }
        if (error) return BSG_ERR_BAD_ARG;
        if (!active) break;
        for (size_t sl = 0; sl < T * B; ++sl)
          leaf_ret[sl] = leaf_base[sl] == bsg::kSkipRoot ? 0 : bsg::fake_leaf_value(seed, leaf_base[sl], n_rollouts);
        for (int v = 0; v < n_trees; ++v) bsg::ForestTree<G>::finish_round(F, v);
      }
      for (int v = 0; v < n_trees; ++v) bsg::ForestTree<G>::expand_root(F, v);
      for (int v = 0; v < n_trees; ++v) {
        bsg::MctsTree<G>& t = ref[(size_t)v];
        int rc = t.search(rootbuf.data(), budget + (phase == 0 ? v % 3 : 0), phase);
        if (rc != BSG_OK) return rc;
        t.expand_root(rootbuf.data());
        if ((int)t.nodes.size() != n_nodes[(size_t)v]) return 1 + v;
        if (t.next_playout != next_playout[(size_t)v] || t.shuffle_ctr != shuffle_ctr[(size_t)v]) return 1 + v;
        for (size_t i = 0; i < t.nodes.size(); ++i) {
          const bsg::MctsNode& a = t.nodes[i];
          const bsg::MctsNode& b = nodes[(size_t)v * (size_t)cap + i];
          if (a.action != b.action || a.player != b.player || a.explore != b.explore || a.reward != b.reward ||
              a.virt != b.virt || a.has_outcome != b.has_outcome || a.out0 != b.out0 ||
              a.first_child != b.first_child || a.n_children != b.n_children)
            return 1 + v;
        }
      }
    }
    return 0;
  }
};

struct ForestSearchOp {
  bsg_forest* f; int resume; int64_t max_left;   // max_left: the largest per-tree budget of this step
  template <class G> int launch_leaves(cudaStream_t st) const {
    bsg_engine* e = f->e;
    const bsg::ForestView& F = f->F;
    const int64_t slots = (int64_t)F.n_trees * F.batch;
    PlayoutArgs a;
    memset(&a, 0, sizeof(a));
    a.roots = F.leaf_states; a.n_roots = slots; a.root0 = 0;
    a.n_playouts = (uint32_t)(slots * F.n_rollouts); a.rollouts = (uint32_t)F.n_rollouts;
    a.first_playout = 0;
    bsg::philox_schedule(f->eval_seed, a.keys);
    a.io.root_sum_ret0 = F.leaf_ret;
    a.io.totals = f->d_totals;
    a.io.root_playout_base = F.leaf_base;
    int rc;
    if constexpr (fillable<G>::value) {   // bsg_engine_set_value_only: same leaf values, fewer instructions
      a.io.value_only = e->value_only;
      rc = e->value_only ? launch_playouts<G, false, true, true>(e, a, st) : launch_playouts<G, false, true>(e, a, st);
    } else {
      rc = launch_playouts<G, false, true>(e, a, st);
    }
    f->launches += 3 + (a.io.value_only ? 1 : 0);   // descend + playouts (+ fill_resolve) + finish
    return rc;
  }
  template <class G> int operator()() const {
    bsg_engine* e = f->e;
    cudaStream_t st = e->own_stream;
    const bsg::ForestView& F = f->F;
    const int grid = forest_grid(f);
    const int wgrid = (F.n_trees * 32 + kForestWarpThreads - 1) / kForestWarpThreads;
    if (!resume) forest_reset_kernel<G><<<grid, kForestThreads, 0, st>>>(F);
    BSG_CUDA(e, cudaMemsetAsync(f->d_totals, 0, 4 * sizeof(uint64_t), st));
    const int64_t slots = (int64_t)F.n_trees * F.batch;
    // A tree with k simulations left needs ceil(k / batch) rounds unless a solved terminal leaf
    // cut one of its rounds short, so that many rounds are enqueued back to back with no host
    // synchronisation (a tree that is done idles through them: its slots are skipped by the
    // playout kernel); the last of them — and every round after it — has the backup kernel count
    // the trees that still have simulations left: normally the first answer is "none".
    const int64_t burst = (max_left + F.batch - 1) / F.batch;
    int rc;
    uint32_t hdr[2] = {0, 0};   // {trees still searching, error}
    for (int64_t r = 0;; ++r) {
      const bool enqueued = r < burst;            // the rounds known to be needed: no host sync between them
      if (!enqueued && hdr[0] == 0) break;
      BSG_CUDA(e, cudaMemsetAsync(F.leaf_ret, 0, (size_t)slots * sizeof(int32_t), st));
      forest_descend_kernel<G><<<wgrid, kForestWarpThreads, 0, st>>>(F, (const typename G::Tables*)e->tables_dev);
      if ((rc = launch_leaves<G>(st)) != BSG_OK) return rc;
      const bool ask = r + 1 >= burst;            // the backup kernel counts the trees with simulations left
      if (ask) BSG_CUDA(e, cudaMemsetAsync(F.active, 0, sizeof(uint32_t), st));
      forest_finish_kernel<G><<<wgrid, kForestWarpThreads, 0, st>>>(F);
      f->rounds += 1;
      if (ask) {
        BSG_CUDA(e, cudaMemcpyAsync(&hdr[0], F.active, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        BSG_CUDA(e, cudaMemcpyAsync(&hdr[1], F.error, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        BSG_CUDA(e, cudaStreamSynchronize(st));
        if (hdr[1]) {
          f->err = hdr[1] == 1 ? "forest: node pool exhausted (raise node_capacity)"
                 : (hdr[1] == 2 ? "forest: log table exhausted (raise max_visits)" : "forest: path deeper than the game allows");
          return BSG_ERR_BAD_ARG;
        }
      }
    }
    forest_expand_root_kernel<G><<<grid, kForestThreads, 0, st>>>(F);
    // totals of this step's playouts (game steps, stuck lanes): accumulated on the device, read
    // by bsg_forest_info / the next synchronising call
    forest_totals_kernel<<<1, 32, 0, st>>>(f->d_totals, f->d_step_totals);
    BSG_CUDA(e, cudaGetLastError());
    return BSG_OK;
  }
};
}  // namespace

extern "C" {

int bsg_forest_create(bsg_engine* e, int32_t n_trees, int32_t node_capacity, int64_t max_visits,
                      double uct_c, int solve, int n_rollouts, int leaf_batch, uint64_t eval_seed,
                      uint64_t shuffle_seed0, uint64_t first_playout, bsg_forest** out) {
  if (!e || !out || n_trees < 1 || node_capacity < 2 || max_visits < 4 || n_rollouts < 1 || leaf_batch < 1)
    return BSG_ERR_BAD_ARG;
  *out = nullptr;
  if ((int64_t)n_trees * leaf_batch * n_rollouts >= ((int64_t)1 << 32) - (1 << 24))
    return fail(e, BSG_ERR_BAD_ARG, "forest: trees * leaf_batch * rollouts exceeds one launch");
  BSG_CUDA(e, cudaSetDevice(e->device));
  int depth = 0;
  int rc = dispatch_device_game(e->game, ForestGeomOp{&depth});
  if (rc != BSG_OK) return rc;
  bsg_forest* f = new (std::nothrow) bsg_forest();
  if (!f) return BSG_ERR_BAD_ARG;
  f->e = e;
  f->eval_seed = eval_seed;
  bsg::ForestView& F = f->F;
  memset(&F, 0, sizeof(F));
  F.n_trees = n_trees; F.cap = node_capacity; F.batch = leaf_batch; F.depth = depth;
  F.n_rollouts = n_rollouts; F.solve = solve; F.uct_c = uct_c; F.shuffle_seed0 = shuffle_seed0;
  F.logtab_n = (int32_t)(max_visits + leaf_batch + 2);
  const size_t T = (size_t)n_trees, B = (size_t)leaf_batch;
  double* logtab = nullptr;
  rc = forest_alloc(f, &F.nodes, T * (size_t)node_capacity);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.n_nodes, T);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.shuffle_ctr, T);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.next_playout, T);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.sims_left, T);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.paths, T * B * (size_t)depth);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.path_len, T * B);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.leaf_kind, T * B);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.leaf_out0, T * B);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.n_pending, T);
  unsigned char* leaf_states = nullptr;
  if (rc == BSG_OK) rc = forest_alloc(f, &leaf_states, T * B * (size_t)bsg_state_bytes(e->game));
  if (rc == BSG_OK) rc = forest_alloc(f, &F.leaf_base, T * B);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.leaf_ret, T * B);
  if (rc == BSG_OK) rc = forest_alloc(f, &logtab, (size_t)F.logtab_n);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.error, 4);
  if (rc == BSG_OK) rc = forest_alloc(f, &F.active, 4);
  unsigned char* d_root = nullptr;
  if (rc == BSG_OK) rc = forest_alloc(f, &d_root, (size_t)bsg_state_bytes(e->game) + 64);
  if (rc == BSG_OK) rc = forest_alloc(f, &f->d_per_tree, T);
  if (rc == BSG_OK) rc = forest_alloc(f, &f->d_totals, 4);
  f->num_actions = bsg_num_distinct_actions(e->game);
  if (rc == BSG_OK) rc = forest_alloc(f, &f->d_step_totals, 4);
  if (rc == BSG_OK) rc = forest_alloc(f, &f->d_base, (size_t)f->num_actions * 2);
  if (rc == BSG_OK) rc = forest_alloc(f, &f->d_merge, (size_t)f->num_actions * bsg::kMergeFields);
  if (rc != BSG_OK) { e->err = f->err; delete f; return rc; }
  F.leaf_states = leaf_states;
  F.root = d_root;
  f->d_root = d_root;
  {  // log(n) from the host's libm — the table the host tree's log() calls amount to
    std::vector<double> h((size_t)F.logtab_n);
    h[0] = 0.0;
    for (int32_t i = 1; i < F.logtab_n; ++i) h[(size_t)i] = log((double)i);
    cudaMemcpy(logtab, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice);
    F.logtab = logtab;
    std::vector<uint64_t> np(T);
    for (size_t v = 0; v < T; ++v) np[v] = first_playout + ((uint64_t)v << 40);
    cudaMemcpy(F.next_playout, np.data(), T * sizeof(uint64_t), cudaMemcpyHostToDevice);
  }
  BSG_CUDA(e, cudaGetLastError());
  *out = f;
  return BSG_OK;
}
void bsg_forest_destroy(bsg_forest* f) {
  if (!f) return;
  if (f->e && f->e->device >= 0) cudaSetDevice(f->e->device);
  delete f;
}
const char* bsg_forest_last_error(bsg_forest* f) {
  return f ? (f->err.empty() ? bsg_last_error(f->e) : f->err.c_str()) : "null forest";
}
int bsg_forest_search(bsg_forest* f, const void* root_host, int64_t simulations_per_tree,
                      const int64_t* per_tree_host, int resume) {
  if (!f || !root_host || simulations_per_tree < 0) return BSG_ERR_BAD_ARG;
  bsg_engine* e = f->e;
  f->err.clear();
  BSG_CUDA(e, cudaSetDevice(e->device));
  cudaStream_t st = e->own_stream;
  BSG_CUDA(e, cudaMemcpyAsync(f->d_root, root_host, (size_t)bsg_state_bytes(e->game), cudaMemcpyHostToDevice, st));
  if (per_tree_host)
    BSG_CUDA(e, cudaMemcpyAsync(f->d_per_tree, per_tree_host, (size_t)f->F.n_trees * sizeof(int64_t),
                                cudaMemcpyHostToDevice, st));
  forest_fill_kernel<<<forest_grid(f), kForestThreads, 0, st>>>(
      f->F.sims_left, simulations_per_tree, per_tree_host ? f->d_per_tree : nullptr, f->F.n_trees);
  int64_t max_left = simulations_per_tree;
  if (per_tree_host) {
    max_left = 0;
    for (int32_t v = 0; v < f->F.n_trees; ++v) if (per_tree_host[v] > max_left) max_left = per_tree_host[v];
  }
  return dispatch_device_game(e->game, ForestSearchOp{f, resume, max_left});
}
int bsg_forest_pack_root(bsg_forest* f, const int64_t* base_dev, int64_t* out_dev) {
  if (!f || !base_dev || !out_dev) return BSG_ERR_BAD_ARG;
  bsg_engine* e = f->e;
  BSG_CUDA(e, cudaSetDevice(e->device));
  forest_pack_kernel<<<forest_grid(f), kForestThreads, 0, e->own_stream>>>(
      f->F, base_dev, (unsigned long long*)out_dev);
  BSG_CUDA(e, cudaStreamSynchronize(e->own_stream));
  return BSG_OK;
}
int bsg_forest_install_root(bsg_forest* f, const int64_t* base_dev, const int64_t* summed_dev) {
  if (!f || !base_dev || !summed_dev) return BSG_ERR_BAD_ARG;
  bsg_engine* e = f->e;
  BSG_CUDA(e, cudaSetDevice(e->device));
  forest_install_kernel<<<forest_grid(f), kForestThreads, 0, e->own_stream>>>(f->F, base_dev, summed_dev);
  BSG_CUDA(e, cudaStreamSynchronize(e->own_stream));
  return BSG_OK;
}
// ---- one-pass trajectories
}  // extern "C"
namespace {
struct TrajCompactOp {
  bsg_engine* e; const void* strided; int64_t n, stride; const int64_t* off; int64_t total;
  const int16_t* acts; int32_t acts_stride; void* out_states; int64_t* oa; int64_t* op; int64_t* os; cudaStream_t st;
  template <class G> int operator()() const {
    const int64_t blocks = (total + 255) / 256;
    const int grid = (int)(blocks < (int64_t)e->sm_count * 16 ? blocks : (int64_t)e->sm_count * 16);
    traj_compact_kernel<G><<<grid < 1 ? 1 : grid, 256, 0, st>>>(strided, n, stride, off, total, acts, acts_stride,
                                                              out_states, oa, op, os);
    BSG_CUDA(e, cudaGetLastError());
    return BSG_OK;
  }
};
}  // namespace
extern "C" {
int bsg_trajectory_scan(bsg_engine* e, const uint8_t* length_dev, int64_t n, int64_t* off_dev,
                        int64_t* total_out, void* stream) {
  if (!e || !length_dev || !off_dev || n < 1) return BSG_ERR_BAD_ARG;
  BSG_CUDA(e, cudaSetDevice(e->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int n_tiles = (int)((n + kScanTile - 1) / kScanTile);
  // tile sums (u32) and tile offsets (int64) live in the engine's result scratch
  int rc = ensure(e, &e->d_scan, &e->d_scan_bytes, (size_t)n_tiles * 12 + 96, false);
  if (rc != BSG_OK) return rc;
  int64_t* tile_off = (int64_t*)e->d_scan;
  u32* tile_sums = (u32*)((char*)e->d_scan + (((size_t)n_tiles * 8 + 15) & ~(size_t)15));
  traj_tile_sums_kernel<<<n_tiles, 256, 0, st>>>(length_dev, n, tile_sums);
  traj_tile_scan_kernel<<<1, 256, 0, st>>>(tile_sums, n_tiles, tile_off, off_dev + n);
  traj_tile_offsets_kernel<<<n_tiles, 256, 0, st>>>(length_dev, n, tile_off, off_dev);
  BSG_CUDA(e, cudaGetLastError());
  if (total_out) {
    BSG_CUDA(e, cudaMemcpyAsync(total_out, off_dev + n, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    BSG_CUDA(e, cudaStreamSynchronize(st));
  }
  return BSG_OK;
}
int bsg_trajectory_compact(bsg_engine* e, const void* strided_states_dev, int64_t n, int64_t stride,
                           const int64_t* off_dev, int64_t total, const int16_t* strided_actions_dev,
                           int32_t actions_stride, void* out_states_dev, int64_t* out_actions_dev,
                           int64_t* out_playout_dev, int64_t* out_step_dev, void* stream) {
  if (!e || !strided_states_dev || !off_dev || !out_states_dev || n < 1 || stride < 1 || total < 0) return BSG_ERR_BAD_ARG;
  if (total == 0) return BSG_OK;
  BSG_CUDA(e, cudaSetDevice(e->device));
  return dispatch_device_game(e->game, TrajCompactOp{e, strided_states_dev, n, stride, off_dev, total, strided_actions_dev,
                                                     actions_stride, out_states_dev, out_actions_dev, out_playout_dev,
                                                     out_step_dev, (cudaStream_t)stream});
}

// ---- self-play actor on the forest (bsg_actor.cuh): one game per tree
}  // extern "C"
namespace {
struct ActorStepOp {
  bsg_forest* f; ActorView A;
  template <class G> int operator()() const {
    bsg_engine* e = f->e;
    actor_step_kernel<G><<<(A.n_games + kForestThreads - 1) / kForestThreads, kForestThreads, 0, e->own_stream>>>(
        f->F, A, (const typename G::Tables*)e->tables_dev);
    BSG_CUDA(e, cudaGetLastError());
    return BSG_OK;
  }
};
struct ActorEmitOp {
  bsg_engine* e; EmitView E;
  template <class G> int operator()() const {
    const int64_t n = (int64_t)E.n_games * E.max_moves;
    actor_emit_kernel<G><<<(unsigned)((n + 127) / 128), 128, 0, e->own_stream>>>(E);
    BSG_CUDA(e, cudaGetLastError());
    return BSG_OK;
  }
};
}  // namespace
extern "C" {

int bsg_forest_search_roots(bsg_forest* f, const void* roots_dev, const uint8_t* alive_dev,
                            int64_t simulations_per_tree) {
  if (!f || !roots_dev || simulations_per_tree < 0) return BSG_ERR_BAD_ARG;
  bsg_engine* e = f->e;
  f->err.clear();
  BSG_CUDA(e, cudaSetDevice(e->device));
  forest_fill_alive_kernel<<<forest_grid(f), kForestThreads, 0, e->own_stream>>>(
      f->F.sims_left, simulations_per_tree, alive_dev, f->F.n_trees);
  f->F.root = roots_dev;
  f->F.root_per_tree = 1;
  const int rc = dispatch_device_game(e->game, ForestSearchOp{f, 0, simulations_per_tree});
  f->F.root = f->d_root;
  f->F.root_per_tree = 0;
  return rc;
}

int bsg_forest_actor_step(bsg_forest* f, const bsg_actor_io* io, uint32_t* n_alive_out) {
  if (!f || !io || !io->states || !io->alive || !io->length || !io->outcome || !io->state_log ||
      !io->policy_log || !io->est_log || !io->action_log || !io->weights || !io->n_alive_dev)
    return BSG_ERR_BAD_ARG;
  if (io->n_games != f->F.n_trees || io->move < 0 || io->move >= io->max_moves || !(io->temperature > 0.0))
    return fail(f->e, BSG_ERR_BAD_ARG, "actor step: one game per tree, 0 <= move < max_moves, temperature > 0");
  bsg_engine* e = f->e;
  BSG_CUDA(e, cudaSetDevice(e->device));
  ActorView A;
  A.states = io->states; A.alive = io->alive; A.length = io->length; A.outcome = io->outcome;
  A.n_alive = io->n_alive_dev; A.state_log = io->state_log; A.policy_log = io->policy_log;
  A.est_log = io->est_log; A.action_log = io->action_log; A.weights = io->weights;
  A.n_games = io->n_games; A.max_moves = io->max_moves; A.move = io->move;
  A.temperature_drop = io->temperature_drop; A.temperature = io->temperature;
  A.sample_seed = io->sample_seed; A.game_id_base = io->game_id_base;
  BSG_CUDA(e, cudaMemsetAsync(A.n_alive, 0, sizeof(uint32_t), e->own_stream));
  int rc = dispatch_device_game(e->game, ActorStepOp{f, A});
  if (rc != BSG_OK) return rc;
  if (n_alive_out) {
    BSG_CUDA(e, cudaMemcpyAsync(n_alive_out, A.n_alive, sizeof(uint32_t), cudaMemcpyDeviceToHost, e->own_stream));
    BSG_CUDA(e, cudaStreamSynchronize(e->own_stream));
  }
  return BSG_OK;
}

int bsg_actor_emit(bsg_engine* e, const bsg_actor_emit_io* io) {
  if (!e || !io || !io->state_log || !io->policy_log || !io->est_log || !io->action_log || !io->length ||
      !io->outcome || !io->offsets || !io->out_states || !io->out_policy || !io->out_value)
    return BSG_ERR_BAD_ARG;
  BSG_CUDA(e, cudaSetDevice(e->device));
  EmitView E;
  E.state_log = io->state_log; E.policy_log = io->policy_log; E.est_log = io->est_log;
  E.action_log = io->action_log; E.length = io->length; E.outcome = io->outcome; E.offsets = io->offsets;
  E.out_states = io->out_states; E.out_policy = io->out_policy; E.out_value = io->out_value;
  E.out_est = io->out_est; E.out_action = io->out_action; E.out_player = io->out_player;
  E.n_games = io->n_games; E.max_moves = io->max_moves; E.total = io->total;
  if (E.total < 1 || E.n_games < 1 || E.max_moves < 1) return BSG_OK;
  return dispatch_device_game(e->game, ActorEmitOp{e, E});
}
void* bsg_engine_stream(bsg_engine* e) { return e ? (void*)e->own_stream : nullptr; }
void* bsg_host_alloc(size_t bytes, int write_combined) {
  void* p = nullptr;
  const unsigned flags = cudaHostAllocPortable | (write_combined ? cudaHostAllocWriteCombined : 0u);
  if (cudaHostAlloc(&p, bytes ? bytes : 1, flags) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}
void bsg_host_free(void* p) { if (p) cudaFreeHost(p); }

// ---- the engine's own NCCL communicator + the root merge on the engine's stream
int bsg_comm_unique_id(uint8_t id_out[128]) {
  if (!id_out) return BSG_ERR_BAD_ARG;
  const NcclApi& api = nccl_api();
  if (!api.ok) return BSG_ERR_NCCL;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  if (api.GetUniqueId(&id) != ncclSuccess) return BSG_ERR_NCCL;
  memcpy(id_out, &id, 128);
  return BSG_OK;
}
int bsg_comm_init(bsg_engine* e, int n_ranks, int rank, const uint8_t id[128]) {
  if (!e || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return BSG_ERR_BAD_ARG;
  const NcclApi& api = nccl_api();
  if (!api.ok) return fail(e, BSG_ERR_NCCL, api.why);
  BSG_CUDA(e, cudaSetDevice(e->device));
  if (e->comm) { api.CommDestroy(e->comm); e->comm = nullptr; }
  ncclUniqueId uid;
  memcpy(&uid, id, 128);
  ncclResult_t st = api.CommInitRank(&e->comm, n_ranks, uid, rank);
  if (st != ncclSuccess) { e->comm = nullptr; return fail(e, BSG_ERR_NCCL, std::string("ncclCommInitRank: ") + api.GetErrorString(st)); }
  e->comm_ranks = n_ranks; e->comm_rank = rank;
  return BSG_OK;
}
int bsg_comm_destroy(bsg_engine* e) {
  if (!e) return BSG_ERR_BAD_ARG;
  if (e->comm && nccl_api().ok) { cudaSetDevice(e->device); nccl_api().CommDestroy(e->comm); }
  e->comm = nullptr; e->comm_ranks = 1; e->comm_rank = 0;
  return BSG_OK;
}
int bsg_comm_ranks(const bsg_engine* e) { return e ? (e->comm ? e->comm_ranks : 1) : BSG_ERR_BAD_ARG; }

// One search step's root merge, entirely on the engine's stream: zero the sum buffer, pack the
// trees' root-child deltas since `base`, ONE int64 ncclAllReduce(sum) over the engine's
// communicator (skipped without one: world size 1), install base + sum into every tree, advance
// base.  No host synchronisation.
int bsg_forest_merge(bsg_forest* f, int first_step) {
  if (!f) return BSG_ERR_BAD_ARG;
  bsg_engine* e = f->e;
  BSG_CUDA(e, cudaSetDevice(e->device));
  cudaStream_t st = e->own_stream;
  const int A = f->num_actions;
  if (first_step) BSG_CUDA(e, cudaMemsetAsync(f->d_base, 0, (size_t)A * 2 * sizeof(int64_t), st));
  BSG_CUDA(e, cudaMemsetAsync(f->d_merge, 0, (size_t)A * bsg::kMergeFields * sizeof(int64_t), st));
  forest_pack_kernel<<<forest_grid(f), kForestThreads, 0, st>>>(f->F, f->d_base, (unsigned long long*)f->d_merge);
  if (e->comm) {
    if (f->ar_used + 2 > f->ar_events.size()) {
      for (int k = 0; k < 2; ++k) {
        cudaEvent_t ev;
        BSG_CUDA(e, cudaEventCreate(&ev));
        f->ar_events.push_back(ev);
      }
    }
    BSG_CUDA(e, cudaEventRecord(f->ar_events[f->ar_used], st));
    ncclResult_t rs = nccl_api().AllReduce(f->d_merge, f->d_merge, (size_t)A * bsg::kMergeFields, ncclInt64, ncclSum, e->comm, st);
    if (rs != ncclSuccess) return fail(e, BSG_ERR_NCCL, std::string("ncclAllReduce: ") + nccl_api().GetErrorString(rs));
    BSG_CUDA(e, cudaEventRecord(f->ar_events[f->ar_used + 1], st));
    f->ar_used += 2;
  }
  forest_install_kernel<<<forest_grid(f), kForestThreads, 0, st>>>(f->F, f->d_base, f->d_merge);
  forest_base_add_kernel<<<(A + 127) / 128, 128, 0, st>>>(f->d_base, f->d_merge, A);
  BSG_CUDA(e, cudaGetLastError());
  f->merges += 1;
  return BSG_OK;
}
// Synchronises the engine's stream and reports the all-reduces issued by bsg_forest_merge since
// the last call: stats = {count, mean microseconds, max microseconds} (CUDA events on the
// engine's stream around each ncclAllReduce).
int bsg_forest_allreduce_stats(bsg_forest* f, double stats[3]) {
  if (!f || !stats) return BSG_ERR_BAD_ARG;
  bsg_engine* e = f->e;
  BSG_CUDA(e, cudaSetDevice(e->device));
  BSG_CUDA(e, cudaStreamSynchronize(e->own_stream));
  double sum = 0, mx = 0;
  const size_t n = f->ar_used / 2;
  for (size_t i = 0; i < n; ++i) {
    float ms = 0;
    BSG_CUDA(e, cudaEventElapsedTime(&ms, f->ar_events[2 * i], f->ar_events[2 * i + 1]));
    sum += ms * 1e3;
    if (ms * 1e3 > mx) mx = ms * 1e3;
  }
  stats[0] = (double)n; stats[1] = n ? sum / (double)n : 0.0; stats[2] = mx;
  f->ar_used = 0;
  return BSG_OK;
}
int bsg_forest_tree_root(bsg_forest* f, int32_t tree, int capacity, int64_t* actions, int64_t* visits,
                         double* reward, int32_t* outcome, double info[4]) {
  if (!f || tree < 0 || tree >= f->F.n_trees) return BSG_ERR_BAD_ARG;
  bsg_engine* e = f->e;
  BSG_CUDA(e, cudaSetDevice(e->device));
  BSG_CUDA(e, cudaStreamSynchronize(e->own_stream));   // a merge may still be in flight
  bsg::MctsNode root;
  const bsg::MctsNode* nodes = f->F.nodes + (size_t)tree * f->F.cap;
  BSG_CUDA(e, cudaMemcpy(&root, nodes, sizeof(root), cudaMemcpyDeviceToHost));
  std::vector<bsg::MctsNode> ch((size_t)(root.n_children > 0 ? root.n_children : 0));
  if (!ch.empty())
    BSG_CUDA(e, cudaMemcpy(ch.data(), nodes + root.first_child, ch.size() * sizeof(bsg::MctsNode), cudaMemcpyDeviceToHost));
  for (int i = 0; i < root.n_children && i < capacity; ++i) {
    const bsg::MctsNode& c = ch[(size_t)i];
    if (actions) actions[i] = c.action;
    if (visits) visits[i] = c.explore;
    if (reward) reward[i] = c.reward;
    if (outcome) outcome[i] = c.has_outcome ? (int32_t)(c.player == 0 ? c.out0 : -c.out0) : 2;
  }
  if (info) {
    int32_t n_nodes = 0;
    BSG_CUDA(e, cudaMemcpy(&n_nodes, f->F.n_nodes + tree, sizeof(n_nodes), cudaMemcpyDeviceToHost));
    info[0] = (double)root.explore; info[1] = root.has_outcome; info[2] = root.out0; info[3] = (double)n_nodes;
  }
  return root.n_children;
}
int bsg_forest_selftest(int game, int n_trees, int sims, int sims_resume, int leaf_batch, int n_rollouts,
                        int solve, uint64_t seed, const int64_t* history, int n_history) {
  if (n_trees < 1 || sims < 0 || sims_resume < 0 || leaf_batch < 1 || n_rollouts < 1 || (n_history > 0 && !history))
    return BSG_ERR_BAD_ARG;
  return dispatch_device_game(game, ForestSelfTestOp{n_trees, sims, sims_resume, leaf_batch, n_rollouts, solve,
                                                     seed, history, n_history});
}
int bsg_forest_info(bsg_forest* f, double info[4]) {
  if (!f || !info) return BSG_ERR_BAD_ARG;
  {
    bsg_engine* e = f->e;
    BSG_CUDA(e, cudaSetDevice(e->device));
    uint64_t tot[4];
    BSG_CUDA(e, cudaMemcpyAsync(tot, f->d_step_totals, sizeof(tot), cudaMemcpyDeviceToHost, e->own_stream));
    BSG_CUDA(e, cudaStreamSynchronize(e->own_stream));
    f->game_steps = (int64_t)tot[1];
    if (tot[3] != 0) { f->err = "forest: non-terminal state without legal actions"; return BSG_ERR_STUCK_LANE; }
  }
  info[0] = (double)f->game_steps; info[1] = (double)f->launches; info[2] = (double)f->rounds;
  info[3] = (double)f->F.n_trees;
  return BSG_OK;
}

// ---- value-only playouts without any allocation (
// no counterpart in the reference).  The scratch of a value-only launch — the launch's work
// counter and one parked board per playout — lives in a file-scope device array of a fixed,
// modest size instead of the engine's cudaMalloc'ed buffers.  Every GPU that loads the library
// holds the array until exit; every call shares it, so calls that overlap on the GPU race.
}  // extern "C"

namespace bsg {
constexpr size_t kStaticScratchHeader = 256;                  // work counter (first word), padded
constexpr size_t kStaticBoardBytes = (size_t)16 << 20;        // 16 MiB of parked boards
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchHeader + kStaticBoardBytes];
}  // namespace bsg

namespace {
// Per device, once, with that device current: the device address of the array (its name is only
// a host-side stand-in in host code) and the launch geometry of the two kernels.
struct StaticScratchSetup {
  unsigned char* ptr = nullptr;
  int blocks_per_sm = 0;
};
template <class G>
int static_scratch_setup(bsg_engine* e, StaticScratchSetup* out) {
  static StaticScratchSetup per_device[kMaxDevices];
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  if (e->device < 0 || e->device >= kMaxDevices) return fail(e, BSG_ERR_BAD_ARG, "device index beyond the per-device table");
  StaticScratchSetup& slot = per_device[e->device];
  if (!slot.ptr) {
    void* p = nullptr;
    BSG_CUDA(e, cudaGetSymbolAddress(&p, g_static_scratch));
    int b = 0;
    BSG_CUDA(e, cudaFuncSetAttribute(playout_kernel<G, false, false, true>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    BSG_CUDA(e, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, playout_kernel<G, false, false, true>, kThreads, 0));
    slot.blocks_per_sm = b > 0 ? b : 1;
    slot.ptr = (unsigned char*)p;
  }
  *out = slot;
  return BSG_OK;
}

struct NoAllocMaxOp {
  template <class G> int operator()() const {
    if constexpr (fillable<G>::value) return (int)(kStaticBoardBytes / ((size_t)G::kFillWords * 4));
    else return 0;
  }
};

struct NoAllocPlayoutOp {
  bsg_engine* e; PlayoutArgs a; cudaStream_t st;
  template <class G> int operator()() const {
    if constexpr (fillable<G>::value) {
      const size_t need = (size_t)a.n_playouts * G::kFillWords * 4;
      if (need > kStaticBoardBytes)
        return fail(e, BSG_ERR_BAD_ARG, "bsg_playouts_value_only_noalloc: " + std::to_string(a.n_playouts) +
                    " playouts exceed the static scratch (" + std::to_string(kStaticBoardBytes / ((size_t)G::kFillWords * 4)) +
                    " playouts of this game; bsg_noalloc_max_playouts)");
      StaticScratchSetup su;
      int rc = static_scratch_setup<G>(e, &su);
      if (rc != BSG_OK) return rc;
      PlayoutArgs b = a;
      b.out_flags = out_flags_of(a.io);
      // the array keeps the last call's contents: the work counter must start at zero (stream
      // order); every playout of this launch overwrites its own board before it is read
      b.work_counter = reinterpret_cast<uint32_t*>(su.ptr);
      b.fill_boards = reinterpret_cast<uint32_t*>(su.ptr + kStaticScratchHeader);
      b.fill_base = 0;
      BSG_CUDA(e, cudaMemsetAsync(b.work_counter, 0, sizeof(uint32_t), st));
      const int64_t need_blocks = ((int64_t)a.n_playouts + kThreads - 1) / kThreads;
      const int64_t max_grid = (int64_t)e->sm_count * su.blocks_per_sm;
      int grid = (int)(need_blocks < max_grid ? need_blocks : max_grid);
      playout_kernel<G, false, false, true><<<grid < 1 ? 1 : grid, kThreads, 0, st>>>(b, (const typename G::Tables*)e->tables_dev);
      BSG_CUDA(e, cudaGetLastError());
      const int rgrid = (int)(need_blocks < (int64_t)e->sm_count * 8 ? need_blocks : (int64_t)e->sm_count * 8);
      fill_resolve_kernel<G><<<rgrid < 1 ? 1 : rgrid, kThreads, 0, st>>>(b);
      BSG_CUDA(e, cudaGetLastError());
      return BSG_OK;
    } else {
      return fail(e, BSG_ERR_BAD_ARG, "value_only needs a game whose full board has exactly one winner (hex, y)");
    }
  }
};
}  // namespace

extern "C" {

int bsg_noalloc_max_playouts(int game) {
  const int n = dispatch_device_game(game, NoAllocMaxOp{});
  return n;
}

int bsg_playouts_value_only_noalloc(bsg_engine* e, const void* roots, int64_t n_roots, int32_t rollouts,
                                    uint64_t seed, uint64_t first_playout, int8_t* ret0, uint8_t* length,
                                    int32_t* root_sum_ret0, int32_t* root_sum_steps, uint64_t* totals,
                                    void* stream) {
  if (!e) return BSG_ERR_BAD_ARG;
  if (!roots || n_roots < 0 || rollouts < 1) return fail(e, BSG_ERR_BAD_ARG, "bad roots/rollouts");
  const int64_t total = n_roots * (int64_t)rollouts;
  if (total >= ((int64_t)1 << 32) - (1 << 24)) return fail(e, BSG_ERR_BAD_ARG, "more than 2^32-2^24 playouts in one launch");
  if (total == 0) return BSG_OK;
  PlayoutArgs a;
  memset(&a, 0, sizeof(a));
  a.roots = roots; a.n_roots = n_roots; a.n_playouts = (uint32_t)total; a.rollouts = (uint32_t)rollouts;
  a.first_playout = first_playout;
  philox_schedule(seed, a.keys);
  a.io.ret0 = ret0; a.io.length = length; a.io.root_sum_ret0 = root_sum_ret0;
  a.io.root_sum_steps = root_sum_steps; a.io.totals = totals; a.io.value_only = 1;
  BSG_CUDA(e, cudaSetDevice(e->device));
  return dispatch_device_game(e->game, NoAllocPlayoutOp{e, a, (cudaStream_t)stream});
}

int bsg_int_peak(bsg_engine* e, int kind, int iters, double* out) {
  if (!e || !out || iters < 1 || kind < 0 || kind > 6) return BSG_ERR_BAD_ARG;
  BSG_CUDA(e, cudaSetDevice(e->device));
  int grid = e->sm_count * 8;
  cudaEvent_t a, b;
  BSG_CUDA(e, cudaEventCreate(&a));
  BSG_CUDA(e, cudaEventCreate(&b));
  u32* sink = (u32*)e->d_totals;
  double best = 0;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(a, e->own_stream);
    if (kind == 0) int_peak_kernel<0><<<grid, kThreads, 0, e->own_stream>>>(iters, sink);
    else if (kind == 1) int_peak_kernel<1><<<grid, kThreads, 0, e->own_stream>>>(iters, sink);
    else if (kind == 2) int_peak_kernel<2><<<grid, kThreads, 0, e->own_stream>>>(iters, sink);
    else if (kind == 3) int_peak_kernel<3><<<grid, kThreads, 0, e->own_stream>>>(iters, sink);
    else if (kind == 4) int_peak_kernel<4><<<grid, kThreads, 0, e->own_stream>>>(iters, sink);
    else if (kind == 5) int_peak_kernel<5><<<grid, kThreads, 0, e->own_stream>>>(iters, sink);
    else int_peak_kernel<6><<<grid, kThreads, 0, e->own_stream>>>(iters, sink);
    cudaEventRecord(b, e->own_stream);
    BSG_CUDA(e, cudaEventSynchronize(b));
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    double per_thread = (double)iters * 64.0 * (kind == 0 ? 2.0 : (kind == 1 ? 1.0 : (kind == 2 ? 3.0 : (kind == 6 ? 4.0 : 2.0))));
    double rate = per_thread * (double)grid * kThreads / (ms * 1e-3);
    if (rep > 0 && rate > best) best = rate;
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  *out = best;
  return BSG_OK;
}

}  // extern "C"
