// Self-play actor on the search forest.
//
// The reference's actor (python/algorithms/alpha_zero/alpha_zero.py:196-228, _play_game) plays
// one game at a time: per move one MCTS search, then
//     policy[a]  = root.children[a].explore_count ** (1 / temperature), normalised
//     action     = best_child() once temperature_drop moves have been made, else sampled from policy
//     record     = (observation_tensor(), current_player, legal_actions_mask(), action, policy,
//                   root.total_reward / root.explore_count)
// and the learner files every record with value = the game's final returns[0]
// (alpha_zero.py:366-369).  Here MANY games are in flight on one GPU: game v is tree v of a
// forest with per-tree roots (ForestView::root_per_tree), one bsg_forest search advances all of
// them by a move, and the kernels below do the rest on the device — no record crosses PCIe.
//
// What is logged per move is the PACKED state (8-244 B); observation planes and legal masks are
// expanded from the log at the end by the batched writer, straight into the replay buffer.
//
// Differences from the reference, both on its unpinned side: the sampled action comes from the
// engine's Philox contract, draw r(sample_seed, game id, move), through the inverse CDF over
// ascending action ids (the reference calls np.random.choice on the global numpy RNG), and no
// Dirichlet noise is mixed into the priors (the rollout evaluator's priors are uniform).
#pragma once
#include "bsg_forest.cuh"

namespace bsg {

struct ActorView {
  void* states;            // packed SoA, capacity n_games: the current position of every game (in/out)
  uint8_t* alive;          // [n_games] 1 while the game is running
  int32_t* length;         // [n_games] moves made (final when the game is over)
  int32_t* outcome;        // [n_games] (int)returns[0] of a finished game
  uint32_t* n_alive;       // device counter: games still running after this step
  // move-major logs, record (m, v) at index m * n_games + v
  void* state_log;         // packed SoA, capacity max_moves * n_games: the state BEFORE move m
  float* policy_log;       // [max_moves * n_games][num_actions]
  float* est_log;          // [max_moves * n_games] root.total_reward / root.explore_count
  int32_t* action_log;     // [max_moves * n_games]
  double* weights;         // scratch [n_games][num_actions]: visit counts ** (1 / temperature)
  int32_t n_games, max_moves, move, temperature_drop;
  double temperature;
  uint64_t sample_seed, game_id_base;
};

// The final-move rule of the reference, SearchNode.best_child / sort_key (mcts.py:144-162):
// (proven outcome for the mover or 0, explore_count, total_reward), first maximum in child order.
BSG_HD int actor_best_child(const MctsNode* nodes, const MctsNode& root) {
  int best = -1;
  double bo = 0.0, br = 0.0;
  int64_t bn = 0;
  for (int i = 0; i < root.n_children; ++i) {
    const MctsNode& c = nodes[root.first_child + i];
    const double o = c.has_outcome ? (c.player == 0 ? c.out0 : -c.out0) : 0.0;
    const bool better = best < 0 || o > bo || (o == bo && (c.explore > bn || (c.explore == bn && c.reward > br)));
    if (better) { best = i; bo = o; bn = c.explore; br = c.reward; }
  }
  return best;
}

template <class G>
struct Actor {
  typedef typename G::State State;

  // One move of game v: record, choose, apply.  Tree v of F holds the finished search of the
  // game's current position.
  BSG_HD static void step(const ForestView& F, const ActorView& A, int v, const typename G::Tables& tab) {
    if (!A.alive[v]) return;
    const int NA = G::kNumActions;
    const MctsNode* nodes = F.nodes + (size_t)v * F.cap;
    const MctsNode& root = nodes[0];
    const size_t rec = (size_t)A.move * A.n_games + v;
    State s;
    G::load(s, A.states, A.n_games, v);
    G::store(s, A.state_log, (int64_t)A.max_moves * A.n_games, (int64_t)rec);
    // policy target: visit counts ** (1 / temperature), normalised, indexed by action id
    double* w = A.weights + (size_t)v * NA;
    float* pol = A.policy_log + rec * NA;
    for (int a = 0; a < NA; ++a) w[a] = 0.0;
    int64_t total = 0;
    double wsum = 0.0;
    const bool unit = A.temperature == 1.0;
    for (int i = 0; i < root.n_children; ++i) {
      const MctsNode& c = nodes[root.first_child + i];
      const double x = unit ? (double)c.explore : pow((double)c.explore, 1.0 / A.temperature);
      w[c.action] = x;
      total += c.explore;
    }
    for (int a = 0; a < NA; ++a) wsum += w[a];
    for (int a = 0; a < NA; ++a) pol[a] = (float)f_div(w[a], wsum);
    A.est_log[rec] = (float)f_div(root.reward, (double)root.explore);
    // action
    int action = -1;
    if (A.move >= A.temperature_drop) {
      const int b = actor_best_child(nodes, root);
      action = nodes[root.first_child + b].action;
    } else {
      PhiloxKeys ks;
      philox_schedule(A.sample_seed, ks);
      u32 out[4];
      const uint64_t gid = A.game_id_base + (uint64_t)v;
      philox4x32_10((u32)gid, (u32)(gid >> 32), (u32)A.move >> 2, 0u, ks, out);
      const u32 r = out[A.move & 3];
      if (unit) {
        // exact: the k-th visit in ascending action order, k = floor(r * total / 2^32)
        const uint64_t k = ((uint64_t)r * (uint64_t)total) >> 32;
        uint64_t cum = 0;
        for (int a = 0; a < NA; ++a) {
          cum += (uint64_t)w[a];
          if (cum > k) { action = a; break; }
        }
      } else {
        const double u = f_mul(f_mul((double)r + 0.5, 1.0 / 4294967296.0), wsum);
        double cum = 0.0;
        for (int a = 0; a < NA; ++a) {
          cum = f_add(cum, w[a]);
          if (w[a] > 0.0) action = a;       // the last positive weight catches rounding at the top
          if (cum > u) break;
        }
      }
    }
    A.action_log[rec] = action;
    forest_apply<G>(s, action, tab, 0);
    G::store(s, A.states, A.n_games, v);
    A.length[v] = A.move + 1;
    if (G::terminal(s)) {
      A.alive[v] = 0;
      A.outcome[v] = outcome_ret0(G::outcome(s));
    } else {
#if defined(__CUDA_ARCH__)
      atomicAdd(A.n_alive, 1u);
#else
      *A.n_alive += 1u;
#endif
    }
  }
};

// Records laid out game-major: record (m, v) goes to row offsets[v] + m (offsets = exclusive
// prefix sum of the lengths), value = the game's final returns[0] (alpha_zero.py:366-369).
struct EmitView {
  const void* state_log; const float* policy_log; const float* est_log; const int32_t* action_log;
  const int32_t* length; const int32_t* outcome; const int64_t* offsets;
  void* out_states;        // packed SoA, capacity total
  float* out_policy;       // [total][num_actions]
  float* out_value;        // [total] final returns[0]
  float* out_est;          // [total] (may be NULL)
  int32_t* out_action;     // [total] (may be NULL)
  int32_t* out_player;     // [total] current player of the record (may be NULL)
  int32_t n_games, max_moves;
  int64_t total;
};

template <class G>
BSG_HD void actor_emit_one(const EmitView& E, int v, int m) {
  if (m >= E.length[v]) return;
  const size_t rec = (size_t)m * E.n_games + v;
  const int64_t dst = E.offsets[v] + m;
  typename G::State s;
  G::load(s, E.state_log, (int64_t)E.max_moves * E.n_games, (int64_t)rec);
  G::store(s, E.out_states, E.total, dst);
  const float* src = E.policy_log + rec * G::kNumActions;
  float* pol = E.out_policy + (size_t)dst * G::kNumActions;
  for (int a = 0; a < G::kNumActions; ++a) pol[a] = src[a];
  E.out_value[dst] = (float)E.outcome[v];
  if (E.out_est) E.out_est[dst] = E.est_log[rec];
  if (E.out_action) E.out_action[dst] = E.action_log[rec];
  if (E.out_player) E.out_player[dst] = G::mover(s);
}

}  // namespace bsg
