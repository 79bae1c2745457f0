// Native host tree for MCTS with batched leaf evaluation on the GPU.
//
// Semantics are those of the reference's MCTSBot (open_spiel/python/algorithms/mcts.py:199-443;
// C++ twin algorithms/mcts.cc:255-428): UCT descent, uniform priors in legal-action order
// shuffled once per expansion, MCTS-Solver backups.  What differs is where time goes: the
// reference calls the evaluator once per leaf; here up to `leaf_batch` leaves are gathered with
// virtual visits (exactly the rule of the Python mirror mcts.MCTSBot._batched_round in this
// repo) and all their rollouts run in ONE launch of the playout kernel through
// bsg_playouts_host.  With the same seeds the native tree and the Python mirror build the
// same tree, visit for visit (tests/test_gpu_mcts.py) — the Python mirror is the specification.
//
// The tree lives on the host because the reference's does (mcts.cc); nodes are a flat vector,
// children of a node are contiguous, states are not stored (a descent replays the actions on a
// copy of the packed root state with the same G::apply the CPU helpers use).
#pragma once
#include <math.h>

#include <limits>
#include <vector>

#include "bsg_host.cuh"

struct bsg_engine;
extern "C" int bsg_playouts_host(bsg_engine* e, const void* roots_host, int64_t n_roots,
                                 int32_t rollouts_per_root, uint64_t seed, uint64_t first_playout,
                                 int64_t totals_out[4], int32_t* root_sum_ret0_host,
                                 int32_t* root_sum_steps_host);
extern "C" uint32_t bsg_draw(uint64_t seed, uint64_t playout, uint32_t t);

namespace bsg {

struct MctsNode {
  int32_t action = -1;
  int32_t player = 0;        // the player who moves INTO this node (mcts.py:94-99)
  int64_t explore = 0;       // explore_count
  double reward = 0.0;       // total_reward, from `player`'s point of view
  int32_t virt = 0;          // virtual visits of in-flight batched simulations
  int32_t has_outcome = 0;   // proven (MCTS-Solver): out0 = returns[0], returns[1] = -out0
  double out0 = 0.0;
  int32_t first_child = -1;
  int32_t n_children = 0;
};

struct MctsBase {
  virtual ~MctsBase() {}
  virtual int search(const void* root_packed, int64_t simulations, int resume) = 0;
  virtual int expand_root(const void* root_packed) = 0;
  std::vector<MctsNode> nodes;
  bsg_engine* engine = nullptr;
  double uct_c = 2.0;
  int solve = 1, n_rollouts = 1, leaf_batch = 1;
  uint64_t eval_seed = 0, shuffle_seed = 0, shuffle_ctr = 0, next_playout = 0;
  int64_t total_game_steps = 0, kernel_launches = 0;
  int root_player = 0;
  // Test hook (bsg_forest_selftest): leaf values from a hash of the playout-id base instead of
  // rollouts, so tree logic can be compared on a machine without a GPU.
  bool fake_eval = false;
  std::string err;
};

// The stand-in leaf value of the test hook: an integer in [-n_rollouts, n_rollouts].
inline int32_t fake_leaf_value(uint64_t seed, uint64_t playout_base, int n_rollouts) {
  PhiloxKeys ks;
  philox_schedule(seed ^ 0x9E3779B97F4A7C15ull, ks);
  u32 out[4];
  philox4x32_10((u32)playout_base, (u32)(playout_base >> 32), 7u, 0u, ks, out);
  return (int32_t)(out[0] % (uint32_t)(2 * n_rollouts + 1)) - n_rollouts;
}

template <class G>
struct MctsTree : MctsBase {
  typedef typename G::State State;
  State root_state;
  // scratch of one batched round
  std::vector<std::vector<int32_t>> paths;
  std::vector<State> leaf_states;
  std::vector<int> leaf_terminal;
  std::vector<unsigned char> packed;
  std::vector<int32_t> root_ret;

  static double ret_of(double out0, int player) { return player == 0 ? out0 : -out0; }

  // Fisher-Yates with the engine's Philox draws: the Python mirror's PhiloxShuffle does the same.
  void shuffle(std::vector<int>& a) {
    for (int i = (int)a.size() - 1; i >= 1; --i) {
      uint32_t r = bsg_draw(shuffle_seed, shuffle_ctr++, 0);
      int j = (int)(r % (uint32_t)(i + 1));
      int tmp = a[i]; a[i] = a[j]; a[j] = tmp;
    }
  }
  // mcts.py:308-324: children from the evaluator's prior (uniform, legal order), shuffled.
  void expand(int32_t node, const State& st) {
    int legal[kMaxLegal > 128 * 6 + 8 ? kMaxLegal : 128 * 6 + 8];
    int n = G::legal(st, legal);
    std::vector<int> acts(legal, legal + n);
    shuffle(acts);
    int32_t first = (int32_t)nodes.size();
    int player = G::mover(st);
    for (int i = 0; i < n; ++i) {
      MctsNode c;
      c.action = acts[i];
      c.player = player;
      nodes.push_back(c);
    }
    nodes[node].first_child = first;
    nodes[node].n_children = n;
  }
  // SearchNode.uct_value (mcts.py:124-133) with the virtual-visit rule of the batched mirror.
  // log_pn = log(parent visits) (log(max(parent, 2)) for a child with virtual visits) is the
  // same for every child of a node, so the caller computes both once per node visit.
  double value_of(const MctsNode& ch, double log_pn, double log_pn2, bool batched) const {
    if (batched && ch.virt && !ch.has_outcome) {
      double n = (double)(ch.explore + ch.virt);
      return (ch.reward - ch.virt) / n + uct_c * sqrt(log_pn2 / n);
    }
    if (ch.has_outcome) return ret_of(ch.out0, ch.player);
    if (ch.explore == 0) return std::numeric_limits<double>::infinity();
    return ch.reward / (double)ch.explore + uct_c * sqrt(log_pn / (double)ch.explore);
  }
  // mcts.py:292-345: descend to a leaf; `working` ends as the leaf's state.
  void tree_policy(std::vector<int32_t>& path, State& working) {
    path.clear();
    path.push_back(0);
    working = root_state;
    int32_t node = 0;
    const bool batched = leaf_batch > 1;
    while (!G::terminal(working) && (nodes[node].explore + nodes[node].virt) > 0) {
      if (nodes[node].n_children == 0 && nodes[node].first_child < 0) expand(node, working);
      const MctsNode& nd = nodes[node];
      int64_t parent_n = nd.explore + nd.virt;
      const double log_pn = log((double)parent_n);
      const double log_pn2 = parent_n > 2 ? log_pn : log(2.0);
      int32_t best = -1;
      double best_v = -std::numeric_limits<double>::infinity();
      const MctsNode* ch = &nodes[nd.first_child];
      for (int i = 0; i < nd.n_children; ++i) {
        double v = value_of(ch[i], log_pn, log_pn2, batched);
        if (best < 0 || v > best_v) { best = nd.first_child + i; best_v = v; }
      }
      if (best < 0) break;  // no legal action in a non-terminal state (blocked breakthrough)
      apply_any<G>(working, nodes[best].action);
      node = best;
      path.push_back(node);
    }
  }
  // mcts.py:408-439 (MCTS-Solver, two players, zero sum).
  void backup(const std::vector<int32_t>& path, double ret0, bool solved) {
    for (int k = (int)path.size() - 1; k >= 0; --k) {
      MctsNode& nd = nodes[path[k]];
      nd.reward += ret_of(ret0, nd.player);
      nd.explore += 1;
      if (solved && nd.n_children > 0) {
        int mover = nodes[nd.first_child].player;
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
  int evaluate(int n_leaves) {
    if (n_leaves == 0) return BSG_OK;
    packed.assign((size_t)n_leaves * G::kStateBytes, 0);
    int li = 0;
    for (size_t i = 0; i < leaf_states.size(); ++i)
      if (!leaf_terminal[i]) G::store(leaf_states[i], packed.data(), n_leaves, li++);
    root_ret.assign(n_leaves, 0);
    if (fake_eval) {
      for (int i = 0; i < n_leaves; ++i)
        root_ret[i] = fake_leaf_value(eval_seed, next_playout + (uint64_t)i * (uint64_t)n_rollouts, n_rollouts);
      next_playout += (uint64_t)n_leaves * (uint64_t)n_rollouts;
      return BSG_OK;
    }
    int64_t totals[4];
    int rc = bsg_playouts_host(engine, packed.data(), n_leaves, n_rollouts, eval_seed, next_playout,
                               totals, root_ret.data(), nullptr);
    if (rc != BSG_OK) return rc;
    next_playout += (uint64_t)n_leaves * (uint64_t)n_rollouts;
    total_game_steps += totals[1];
    kernel_launches += 1;
    return BSG_OK;
  }
  // One round: up to leaf_batch descents kept apart by virtual visits, ONE kernel launch for
  // all non-terminal leaves, then the backups (mcts.MCTSBot._batched_round of this repo).
  int round(int64_t budget, int64_t* done) {
    int64_t k = leaf_batch < budget ? leaf_batch : budget;
    if ((int64_t)paths.size() < k) paths.resize(k);
    leaf_states.clear();
    leaf_terminal.clear();
    int n_pending = 0, n_leaves = 0;
    State working;
    for (int64_t i = 0; i < k; ++i) {
      tree_policy(paths[i], working);
      for (int32_t id : paths[i]) nodes[id].virt += 1;
      bool term = G::terminal(working);
      leaf_states.push_back(working);
      leaf_terminal.push_back(term ? 1 : 0);
      ++n_pending;
      if (!term) ++n_leaves;
      if (term && solve) break;  // a proven leaf may solve the root; resolve it before descending again
    }
    int rc = evaluate(n_leaves);
    if (rc != BSG_OK) return rc;
    int vi = 0;
    for (int i = 0; i < n_pending; ++i) {
      for (int32_t id : paths[i]) nodes[id].virt -= 1;
      if (leaf_terminal[i]) {
        double r0 = (double)outcome_ret0(G::outcome(leaf_states[i]));
        MctsNode& last = nodes[paths[i].back()];
        last.has_outcome = 1;
        last.out0 = r0;
        backup(paths[i], r0, solve != 0);
      } else {
        backup(paths[i], (double)root_ret[vi] / (double)n_rollouts, false);
        ++vi;
      }
    }
    *done += n_pending;
    return BSG_OK;
  }
  void reset(const void* root_packed) {
    G::load(root_state, root_packed, 1, 0);
    nodes.clear();
    MctsNode root;
    root.player = G::mover(root_state);
    nodes.push_back(root);
    root_player = root.player;
  }
  int search(const void* root_packed, int64_t simulations, int resume) override {
    if (!resume || nodes.empty()) reset(root_packed);
    int64_t done = 0;
    while (done < simulations) {
      int rc = round(simulations - done, &done);
      if (rc != BSG_OK) return rc;
      if (nodes[0].has_outcome) break;
    }
    return BSG_OK;
  }
  int expand_root(const void* root_packed) override {
    if (nodes.empty()) reset(root_packed);
    if (nodes[0].first_child < 0 && !G::terminal(root_state)) expand(0, root_state);
    return BSG_OK;
  }
};

struct MctsCreateOp {
  MctsBase** out;
  template <class G> int operator()() const {
    *out = new MctsTree<G>();
    return BSG_OK;
  }
};

}  // namespace bsg
