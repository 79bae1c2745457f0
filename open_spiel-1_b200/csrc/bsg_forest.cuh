// GPU-resident search forest (second half: "then a GPU-resident
// tree").
//
// T independent MCTS trees live in device memory and are searched together: per round one
// kernel descends every tree (one thread per tree, up to `batch` leaves kept apart by virtual
// visits), ONE launch of the playout kernel evaluates all leaves of all trees (n_rollouts
// playouts each; the trees' disjoint playout-id ranges ride in bsg_playout_io.root_playout_base),
// and one kernel backs the results up.  Nothing crosses PCIe inside a round except a 4-byte
// "trees still searching" counter.
//
// Each tree runs EXACTLY the algorithm of the host tree (csrc/bsg_mcts.cuh, itself the
// reference's MCTSBot: python/algorithms/mcts.py:292-443) — same shuffles, same playout ids,
// same UCT doubles (log from a host-built table, IEEE divide / sqrt / add / multiply without
// contraction), same solver backups.  Tree v of the forest is therefore the tree that
// mcts.NativeMCTSBot builds with shuffle_seed + v and next_playout = base + (v << 40), and the
// merged root statistics equal parallel.RootParallelMCTS(trees_per_rank = T) bit for bit
// (tests/test_gpu_forest.py).  What the forest buys is throughput: one host thread walks one
// tree at ~1 us per simulation and keeps a B200 ~1 % busy; thousands of trees walked by the
// GPU itself make the playout kernel the bound.
//
// The per-tree code below is __host__ __device__ so the CPU test-suite can run the very same
// functions on host memory (bsg_forest_selftest).
#pragma once
#include "bsg_mcts.cuh"

namespace bsg {

constexpr uint64_t kSkipRoot = 0xFFFFFFFFFFFFFFFFull;

struct ForestView {
  MctsNode* nodes;          // [T][cap]
  int32_t* n_nodes;         // [T]
  uint64_t* shuffle_ctr;    // [T]
  uint64_t* next_playout;   // [T]
  int64_t* sims_left;       // [T] simulations still to run in this search step
  int32_t* paths;           // [T][batch][depth]
  int32_t* path_len;        // [T][batch]
  int32_t* leaf_kind;       // [T][batch]: 1 = evaluated leaf, 2 = terminal leaf
  double* leaf_out0;        // [T][batch]: returns[0] of a terminal leaf
  int32_t* n_pending;       // [T]
  void* leaf_states;        // packed SoA, capacity T * batch
  uint64_t* leaf_base;      // [T * batch] playout-id base per leaf slot (kSkipRoot = unused)
  int32_t* leaf_ret;        // [T * batch] sum of ret0 over the rollouts of the slot
  const double* logtab;     // log(n), n < logtab_n (built on the host: the host tree's libm)
  int32_t* error;           // != 0: node pool or log table exhausted
  uint32_t* active;         // trees that descended at least once in this round
  const void* root;         // packed root state (capacity 1), or one root per tree (capacity n_trees)
  int32_t root_per_tree;    // != 0: tree v searches root v of `root` (self-play actors: one game per tree)
  uint64_t shuffle_seed0;   // tree v shuffles with seed shuffle_seed0 + v
  int32_t n_trees, cap, batch, depth, n_rollouts, solve, logtab_n;
  double uct_c;
};

// Products without contraction into fused multiply-adds: the host tree and the Python mirror
// round after every operation.
BSG_HD double f_add(double a, double b) {
#if defined(__CUDA_ARCH__)
  return __dadd_rn(a, b);
#else
  return a + b;
#endif
}
BSG_HD double f_mul(double a, double b) {
#if defined(__CUDA_ARCH__)
  return __dmul_rn(a, b);
#else
  return a * b;
#endif
}
BSG_HD double f_div(double a, double b) {
#if defined(__CUDA_ARCH__)
  return __ddiv_rn(a, b);
#else
  return a / b;
#endif
}
BSG_HD double f_sqrt(double a) {
#if defined(__CUDA_ARCH__)
  return __dsqrt_rn(a);
#else
  return sqrt(a);
#endif
}

BSG_HD uint32_t forest_draw(uint64_t seed, uint64_t playout) {   // bsg_draw(seed, playout, 0)
  PhiloxKeys ks;
  philox_schedule(seed, ks);
  u32 out[4];
  philox4x32_10((u32)playout, (u32)(playout >> 32), 0u, 0u, ks, out);
  return out[0];
}

// G::apply with or without the game's tables (hex and pentago take them).
template <class G>
BSG_HD auto forest_apply(typename G::State& s, int a, const typename G::Tables& t, int)
    -> decltype(G::apply(s, a, t)) {
  return G::apply(s, a, t);
}
template <class G>
BSG_HD bool forest_apply(typename G::State& s, int a, const typename G::Tables&, long) {
  return G::apply(s, a);
}

template <class G>
struct ForestTree {
  typedef typename G::State State;
  typedef typename G::Tables Tables;

  BSG_HD static double ret_of(double out0, int player) { return player == 0 ? out0 : -out0; }

  // mcts.py:308-324 / MctsTree::expand
  BSG_HD static void expand(const ForestView& F, int v, MctsNode* nodes, int& n_nodes, int32_t node,
                            const State& st) {
    int legal[kMaxLegal];
    const int n = G::legal(st, legal);
    if (n_nodes + n > F.cap) { *F.error = 1; return; }
    uint64_t ctr = F.shuffle_ctr[v];
    const uint64_t seed = F.shuffle_seed0 + (uint64_t)v;
    for (int i = n - 1; i >= 1; --i) {
      const uint32_t r = forest_draw(seed, ctr++);
      const int j = (int)(r % (uint32_t)(i + 1));
      const int tmp = legal[i]; legal[i] = legal[j]; legal[j] = tmp;
    }
    F.shuffle_ctr[v] = ctr;
    const int player = G::mover(st);
    for (int i = 0; i < n; ++i) {
      MctsNode c;
      c.action = legal[i]; c.player = player; c.explore = 0; c.reward = 0.0; c.virt = 0;
      c.has_outcome = 0; c.out0 = 0.0; c.first_child = -1; c.n_children = 0;
      nodes[n_nodes + i] = c;
    }
    nodes[node].first_child = n_nodes;
    nodes[node].n_children = n;
    n_nodes += n;
  }

  // SearchNode.uct_value with the virtual-visit rule / MctsTree::value_of
  BSG_HD static double value_of(const ForestView& F, const MctsNode& ch, double log_pn, double log_pn2,
                                bool batched, bool& inf) {
    inf = false;
    if (batched && ch.virt && !ch.has_outcome) {
      const double n = (double)(ch.explore + ch.virt);
      return f_add(f_div(ch.reward - (double)ch.virt, n), f_mul(F.uct_c, f_sqrt(f_div(log_pn2, n))));
    }
    if (ch.has_outcome) return ret_of(ch.out0, ch.player);
    if (ch.explore == 0) { inf = true; return 0.0; }
    return f_add(f_div(ch.reward, (double)ch.explore),
                 f_mul(F.uct_c, f_sqrt(f_div(log_pn, (double)ch.explore))));
  }

  // One batched round, first half: descents (MctsTree::round up to evaluate()).
  BSG_HD static void descend(const ForestView& F, int v, const Tables& tab) {
    MctsNode* nodes = F.nodes + (size_t)v * F.cap;
    int n_nodes = F.n_nodes[v];
    const int B = F.batch;
    const int64_t left = F.sims_left[v];   // (a solved root ends the search AFTER a round, as on the host)
    const int k = (int)(left < B ? left : B);
    State root_state;
    G::load(root_state, F.root, F.root_per_tree ? F.n_trees : 1, F.root_per_tree ? v : 0);
    const bool batched = B > 1;
    int n_pending = 0, n_leaves = 0;
    const uint64_t base = F.next_playout[v];
    for (int i = 0; i < k; ++i) {
      int32_t* path = F.paths + ((size_t)v * B + i) * F.depth;
      int len = 0;
      path[len++] = 0;
      State working = root_state;
      int32_t node = 0;
      while (!G::terminal(working) && (nodes[node].explore + nodes[node].virt) > 0) {
        if (nodes[node].n_children == 0 && nodes[node].first_child < 0) {
          expand(F, v, nodes, n_nodes, node, working);
          if (*F.error) break;
        }
        const MctsNode nd = nodes[node];
        const int64_t parent_n = nd.explore + nd.virt;
        if (parent_n >= F.logtab_n) { *F.error = 2; break; }
        const double log_pn = F.logtab[parent_n];
        const double log_pn2 = parent_n > 2 ? log_pn : F.logtab[2];
        int32_t best = -1;
        double best_v = 0.0;
        bool best_inf = false;
        for (int c = 0; c < nd.n_children; ++c) {
          bool inf;
          const double val = value_of(F, nodes[nd.first_child + c], log_pn, log_pn2, batched, inf);
          // v > best_v with +infinity for unvisited children: the first infinite one wins
          const bool better = best < 0 || (!best_inf && (inf || val > best_v));
          if (better) { best = nd.first_child + c; best_v = val; best_inf = inf; }
        }
        if (best < 0) break;
        forest_apply<G>(working, nodes[best].action, tab, 0);
        node = best;
        if (len < F.depth) path[len++] = node; else { *F.error = 3; break; }
      }
      for (int p = 0; p < len; ++p) nodes[path[p]].virt += 1;
      F.path_len[(size_t)v * B + i] = len;
      const bool term = G::terminal(working);
      if (term) {
        F.leaf_kind[(size_t)v * B + i] = 2;
        F.leaf_out0[(size_t)v * B + i] = (double)outcome_ret0(G::outcome(working));
      } else {
        F.leaf_kind[(size_t)v * B + i] = 1;
        const size_t slot = (size_t)v * B + n_leaves;
        G::store(working, F.leaf_states, (int64_t)F.n_trees * B, (int64_t)slot);
        F.leaf_base[slot] = base + (uint64_t)n_leaves * (uint64_t)F.n_rollouts;
        ++n_leaves;
      }
      ++n_pending;
      if (term && F.solve) break;
    }
    for (int i = n_leaves; i < B; ++i) F.leaf_base[(size_t)v * B + i] = kSkipRoot;
    F.next_playout[v] = base + (uint64_t)n_leaves * (uint64_t)F.n_rollouts;
    F.n_pending[v] = n_pending;
    F.n_nodes[v] = n_nodes;
  }

  // mcts.py:408-439 / MctsTree::backup
  BSG_HD static void backup(MctsNode* nodes, const int32_t* path, int len, double ret0, bool solved) {
    for (int k = len - 1; k >= 0; --k) {
      MctsNode& nd = nodes[path[k]];
      nd.reward += ret_of(ret0, nd.player);
      nd.explore += 1;
      if (solved && nd.n_children > 0) {
        const int mover = nodes[nd.first_child].player;
        int32_t best = -1;
        bool all_solved = true;
        for (int i = 0; i < nd.n_children; ++i) {
          const MctsNode& c = nodes[nd.first_child + i];
          if (!c.has_outcome) all_solved = false;
          else if (best < 0 || ret_of(c.out0, mover) > ret_of(nodes[best].out0, mover)) best = nd.first_child + i;
        }
        if (best >= 0 && (all_solved || ret_of(nodes[best].out0, mover) == 1.0)) {
          nd.has_outcome = 1;
          nd.out0 = nodes[best].out0;
        } else {
          solved = false;
        }
      }
    }
  }

  // One batched round, second half (MctsTree::round after evaluate()).
  BSG_HD static void finish_round(const ForestView& F, int v) {
    MctsNode* nodes = F.nodes + (size_t)v * F.cap;
    const int B = F.batch;
    const int n_pending = F.n_pending[v];
    int vi = 0;
    for (int i = 0; i < n_pending; ++i) {
      const int32_t* path = F.paths + ((size_t)v * B + i) * F.depth;
      const int len = F.path_len[(size_t)v * B + i];
      for (int p = 0; p < len; ++p) nodes[path[p]].virt -= 1;
      if (F.leaf_kind[(size_t)v * B + i] == 2) {
        const double r0 = F.leaf_out0[(size_t)v * B + i];
        MctsNode& last = nodes[path[len - 1]];
        last.has_outcome = 1;
        last.out0 = r0;
        backup(nodes, path, len, r0, F.solve != 0);
      } else {
        const double r0 = (double)F.leaf_ret[(size_t)v * B + vi] / (double)F.n_rollouts;
        backup(nodes, path, len, r0, false);
        ++vi;
      }
    }
    F.sims_left[v] -= n_pending;
    if (nodes[0].has_outcome) F.sims_left[v] = 0;
    F.n_pending[v] = 0;
  }

#if defined(__CUDACC__)
  // ---- the same two halves of a round with one WARP per tree (device only).  The search is
  // unchanged — the lanes only share the work inside a step of it: the children of a node are
  // scored in parallel (one child per lane, strided when a node has more than 32) and an ordered
  // butterfly reduction returns what the sequential scan of descend() returns — the first
  // unvisited child if there is one, else the lowest-index child of maximal value; the virtual
  // visits of a path and the plain (unsolved) backups touch distinct nodes and run one node per
  // lane.  Expansion (the Philox shuffle) and solver backups stay on lane 0.  Every branch below
  // is warp-uniform: all lanes read the same memory after a __syncwarp().
  struct Pick { double val; int32_t idx; int32_t inf; };
  __device__ static bool pick_beats(const Pick& a, const Pick& b) {   // a before b in the scan's order
    if (b.idx < 0) return a.idx >= 0;
    if (a.idx < 0) return false;
    if (a.inf != b.inf) return a.inf != 0;
    if (a.inf) return a.idx < b.idx;
    if (a.val != b.val) return a.val > b.val;
    return a.idx < b.idx;
  }
  __device__ static void descend_warp(const ForestView& F, int v, const Tables& tab) {
    const unsigned full = 0xFFFFFFFFu;
    const int lane = (int)(threadIdx.x & 31);
    MctsNode* nodes = F.nodes + (size_t)v * F.cap;
    int n_nodes = F.n_nodes[v];
    const int B = F.batch;
    const int64_t left = F.sims_left[v];
    const int k = (int)(left < B ? left : B);
    State root_state;
    G::load(root_state, F.root, F.root_per_tree ? F.n_trees : 1, F.root_per_tree ? v : 0);
    const bool batched = B > 1;
    int n_pending = 0, n_leaves = 0;
    const uint64_t base = F.next_playout[v];
    for (int i = 0; i < k; ++i) {
      int32_t* path = F.paths + ((size_t)v * B + i) * F.depth;
      int len = 1;
      if (lane == 0) path[0] = 0;
      State working = root_state;
      int32_t node = 0;
      for (;;) {
        MctsNode nd = nodes[node];
        if (G::terminal(working) || (nd.explore + nd.virt) <= 0) break;
        if (nd.n_children == 0 && nd.first_child < 0) {
          if (lane == 0) expand(F, v, nodes, n_nodes, node, working);
          __syncwarp();
          n_nodes = __shfl_sync(full, n_nodes, 0);
          nd = nodes[node];
          if (nd.first_child < 0) break;          // node pool exhausted (*F.error is set)
        }
        const int64_t parent_n = nd.explore + nd.virt;
        if (parent_n >= F.logtab_n) { if (lane == 0) *F.error = 2; break; }
        const double log_pn = F.logtab[parent_n];
        const double log_pn2 = parent_n > 2 ? log_pn : F.logtab[2];
        Pick mine;
        mine.val = 0.0; mine.idx = -1; mine.inf = 0;
        for (int c = lane; c < nd.n_children; c += 32) {
          bool inf;
          const double val = value_of(F, nodes[nd.first_child + c], log_pn, log_pn2, batched, inf);
          Pick cand;
          cand.val = val; cand.idx = nd.first_child + c; cand.inf = inf ? 1 : 0;
          if (pick_beats(cand, mine)) mine = cand;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          Pick other;
          other.val = __shfl_xor_sync(full, mine.val, o);
          other.idx = __shfl_xor_sync(full, mine.idx, o);
          other.inf = __shfl_xor_sync(full, mine.inf, o);
          if (pick_beats(other, mine)) mine = other;
        }
        const int32_t best = mine.idx;
        if (best < 0) break;
        forest_apply<G>(working, nodes[best].action, tab, 0);
        node = best;
        if (len < F.depth) { if (lane == 0) path[len] = node; ++len; }
        else { if (lane == 0) *F.error = 3; break; }
      }
      __syncwarp();
      for (int p = lane; p < len; p += 32) nodes[path[p]].virt += 1;   // distinct nodes
      __syncwarp();
      const bool term = G::terminal(working);
      if (lane == 0) {
        F.path_len[(size_t)v * B + i] = len;
        if (term) {
          F.leaf_kind[(size_t)v * B + i] = 2;
          F.leaf_out0[(size_t)v * B + i] = (double)outcome_ret0(G::outcome(working));
        } else {
          F.leaf_kind[(size_t)v * B + i] = 1;
          const size_t slot = (size_t)v * B + n_leaves;
          G::store(working, F.leaf_states, (int64_t)F.n_trees * B, (int64_t)slot);
          F.leaf_base[slot] = base + (uint64_t)n_leaves * (uint64_t)F.n_rollouts;
        }
      }
      if (!term) ++n_leaves;
      ++n_pending;
      if (term && F.solve) break;
    }
    for (int i = n_leaves + lane; i < B; i += 32) F.leaf_base[(size_t)v * B + i] = kSkipRoot;
    if (lane == 0) {
      F.next_playout[v] = base + (uint64_t)n_leaves * (uint64_t)F.n_rollouts;
      F.n_pending[v] = n_pending;
      F.n_nodes[v] = n_nodes;
    }
  }

  __device__ static void finish_round_warp(const ForestView& F, int v) {
    const int lane = (int)(threadIdx.x & 31);
    MctsNode* nodes = F.nodes + (size_t)v * F.cap;
    const int B = F.batch;
    const int n_pending = F.n_pending[v];
    int vi = 0;
    for (int i = 0; i < n_pending; ++i) {
      const int32_t* path = F.paths + ((size_t)v * B + i) * F.depth;
      const int len = F.path_len[(size_t)v * B + i];
      if (F.leaf_kind[(size_t)v * B + i] == 2) {
        if (lane == 0) {
          for (int p = 0; p < len; ++p) nodes[path[p]].virt -= 1;
          const double r0 = F.leaf_out0[(size_t)v * B + i];
          MctsNode& last = nodes[path[len - 1]];
          last.has_outcome = 1;
          last.out0 = r0;
          backup(nodes, path, len, r0, F.solve != 0);
        }
      } else {
        // backup(.., solved = false): every node of the path gets the same three updates
        const double r0 = (double)F.leaf_ret[(size_t)v * B + vi] / (double)F.n_rollouts;
        for (int p = lane; p < len; p += 32) {
          MctsNode& nd = nodes[path[p]];
          nd.virt -= 1;
          nd.reward += ret_of(r0, nd.player);
          nd.explore += 1;
        }
        ++vi;
      }
      __syncwarp();
    }
    if (lane == 0) {
      F.sims_left[v] -= n_pending;
      if (nodes[0].has_outcome) F.sims_left[v] = 0;
      F.n_pending[v] = 0;
    }
  }
#endif

  // A tree that ran no simulation still joins the merge with an expanded root
  // (parallel.RootParallelMCTS._search_tree / MctsTree::expand_root).
  BSG_HD static void expand_root(const ForestView& F, int v) {
    MctsNode* nodes = F.nodes + (size_t)v * F.cap;
    State root_state;
    G::load(root_state, F.root, F.root_per_tree ? F.n_trees : 1, F.root_per_tree ? v : 0);
    if (nodes[0].first_child < 0 && !G::terminal(root_state)) {
      int n_nodes = F.n_nodes[v];
      expand(F, v, nodes, n_nodes, 0, root_state);
      F.n_nodes[v] = n_nodes;
    }
  }

  BSG_HD static void reset(const ForestView& F, int v) {
    MctsNode* nodes = F.nodes + (size_t)v * F.cap;
    State root_state;
    G::load(root_state, F.root, F.root_per_tree ? F.n_trees : 1, F.root_per_tree ? v : 0);
    MctsNode r;
    r.action = -1; r.player = G::mover(root_state); r.explore = 0; r.reward = 0.0; r.virt = 0;
    r.has_outcome = 0; r.out0 = 0.0; r.first_child = -1; r.n_children = 0;
    nodes[0] = r;
    F.n_nodes[v] = 1;
    F.n_pending[v] = 0;
  }
};

// ---- root merge (parallel.RootParallelMCTS._pack / _merge), dense per action id ----------
// fields per action: visit delta, scaled reward delta, then how many trees hold a solver
// proof "mover loses / draws / wins" for that child.
constexpr int kMergeFields = 5;

BSG_HD void forest_pack_root(const ForestView& F, int v, const int64_t* base, unsigned long long* out) {
  const MctsNode* nodes = F.nodes + (size_t)v * F.cap;
  const MctsNode& root = nodes[0];
  for (int i = 0; i < root.n_children; ++i) {
    const MctsNode& c = nodes[root.first_child + i];
    const int a = c.action;
#if defined(__CUDA_ARCH__)
    const long long scaled = __double2ll_rn(c.reward * (double)F.n_rollouts);
#else
    const long long scaled = llrint(c.reward * (double)F.n_rollouts);
#endif
    const long long dv = (long long)c.explore - (long long)base[2 * a];
    const long long dr = scaled - (long long)base[2 * a + 1];
    const int proven = c.has_outcome ? (int)(c.player == 0 ? c.out0 : -c.out0) + 1 : -1;
#if defined(__CUDA_ARCH__)
    atomicAdd(out + a * kMergeFields + 0, (unsigned long long)dv);
    atomicAdd(out + a * kMergeFields + 1, (unsigned long long)dr);
    if (proven >= 0) atomicAdd(out + a * kMergeFields + 2 + proven, 1ull);
#else
    out[a * kMergeFields + 0] += (unsigned long long)dv;
    out[a * kMergeFields + 1] += (unsigned long long)dr;
    if (proven >= 0) out[a * kMergeFields + 2 + proven] += 1ull;
#endif
  }
}

BSG_HD void forest_install_root(const ForestView& F, int v, const int64_t* base, const int64_t* summed) {
  MctsNode* nodes = F.nodes + (size_t)v * F.cap;
  MctsNode& root = nodes[0];
  int64_t visits = 0;
  for (int i = 0; i < root.n_children; ++i) {
    MctsNode& c = nodes[root.first_child + i];
    const int a = c.action;
    const int64_t n = base[2 * a] + summed[a * kMergeFields + 0];
    const int64_t r = base[2 * a + 1] + summed[a * kMergeFields + 1];
    c.explore = n;
    c.reward = (double)r / (double)F.n_rollouts;
    visits += n;
    const int64_t* proofs = summed + a * kMergeFields + 2;
    if (!c.has_outcome && (proofs[0] | proofs[1] | proofs[2]) != 0) {
      int arg = 0;                                    // numpy.argmax: the first maximum
      if (proofs[1] > proofs[arg]) arg = 1;
      if (proofs[2] > proofs[arg]) arg = 2;
      const double val = (double)(arg - 1);           // for the mover
      c.has_outcome = 1;
      c.out0 = c.player == 0 ? val : -val;
    }
  }
  if (root.n_children > 0 && visits > root.explore) root.explore = visits;
}

}  // namespace bsg
