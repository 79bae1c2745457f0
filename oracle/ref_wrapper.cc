// TEST INFRASTRUCTURE ONLY — C-ABI veneer over the UNMODIFIED reference C++ classes.
//
// Compiled by oracle/Makefile together with the reference's own sources, read in place
// from /root/reference (never copied), into oracle/_ref/libspiel_ref.so.  Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs load it.
//
// Everything here goes through the reference's public API exactly as a user would:
//   open_spiel::LoadGame            open_spiel/spiel.h:794-803
//   Game::NewInitialState           open_spiel/spiel.h:622
//   State::ApplyAction              open_spiel/spiel.h:205-210
//   State::LegalActions/IsTerminal/Returns/ObservationTensor/CurrentPlayer
//   algorithms::MCTSBot + RandomRolloutEvaluator   open_spiel/algorithms/mcts.h:92-208
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "open_spiel/algorithms/mcts.h"
#include "open_spiel/spiel.h"
#include "philox_ref.h"

namespace {

std::shared_ptr<const open_spiel::Game> GetGame(const char* name) {
  static std::mutex mu;
  static std::map<std::string, std::shared_ptr<const open_spiel::Game>> cache;
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(name);
  if (it != cache.end()) return it->second;
  auto g = open_spiel::LoadGame(name);
  cache[name] = g;
  return g;
}

std::unique_ptr<open_spiel::State> FromHistory(const open_spiel::Game& game,
                                               const int64_t* hist, int n) {
  auto s = game.NewInitialState();
  for (int i = 0; i < n; ++i) s->ApplyAction(hist[i]);
  return s;
}

}  // namespace

extern "C" {

// Static facts of a game: NumDistinctActions, MaxGameLength, ObservationTensorSize.
int ref_game_info(const char* game, int* num_actions, int* max_len, int* obs_size) {
  auto g = GetGame(game);
  *num_actions = g->NumDistinctActions();
  *max_len = g->MaxGameLength();
  *obs_size = g->ObservationTensorSize();
  return 0;
}

// Full dump of the state reached by `hist`: legal actions, current player, terminal flag,
// returns, and the observation tensor for `player` (as the reference's doubles).
int ref_state_dump(const char* game, const int64_t* hist, int n_hist, int player,
                   double* obs_out, int64_t* legal_out, int* n_legal, int* cur_player,
                   int* terminal, double* returns_out) {
  auto g = GetGame(game);
  auto s = FromHistory(*g, hist, n_hist);
  std::vector<open_spiel::Action> legal = s->LegalActions();
  *n_legal = static_cast<int>(legal.size());
  if (legal_out) std::copy(legal.begin(), legal.end(), legal_out);
  *cur_player = s->CurrentPlayer();
  *terminal = s->IsTerminal() ? 1 : 0;
  std::vector<double> ret = s->Returns();
  returns_out[0] = ret[0];
  returns_out[1] = ret[1];
  if (obs_out) {
    std::vector<double> obs;
    s->ObservationTensor(player, &obs);
    std::copy(obs.begin(), obs.end(), obs_out);
  }
  return 0;
}

// Replay contract: from the state reached by `hist`, repeat
//   a = LegalActions()[r_t mod n]; ApplyAction(a)
// until terminal or `max_steps` moves.  r_t comes from `r_table` when given, otherwise from
// the Philox definition in philox_ref.h with (seed, playout).
// Outputs: the actions played, their count, IsTerminal() and (int)Returns()[0] at the end.
// Returns -1 if a non-terminal state offered no legal action.
int ref_replay(const char* game, const int64_t* hist, int n_hist, const uint32_t* r_table,
               uint64_t seed, uint64_t playout, int max_steps, int64_t* actions_out,
               int* len_out, int* terminal_out, int* ret0_out) {
  auto g = GetGame(game);
  auto s = FromHistory(*g, hist, n_hist);
  int t = 0;
  while (!s->IsTerminal() && t < max_steps) {
    std::vector<open_spiel::Action> legal = s->LegalActions();
    if (legal.empty()) return -1;
    uint32_t r = r_table ? r_table[t] : oracle_draw(seed, playout, (uint32_t)t);
    open_spiel::Action a = legal[r % legal.size()];
    if (actions_out) actions_out[t] = a;
    s->ApplyAction(a);
    ++t;
  }
  *len_out = t;
  *terminal_out = s->IsTerminal() ? 1 : 0;
  *ret0_out = static_cast<int>(s->Returns()[0]);
  return 0;
}

// Batched replay over many roots given as a ragged history array; playout id of root i,
// rollout j is (first_playout + i * rollouts + j).  Used by parity tests at >= 1e5 roots.
int ref_replay_batch(const char* game, const int64_t* hist_flat, const int64_t* hist_off,
                     int64_t n_roots, int rollouts, uint64_t seed, uint64_t first_playout,
                     int max_steps, int threads, int32_t* len_out, int8_t* ret0_out,
                     uint8_t* terminal_out, int16_t* actions_out, int actions_stride) {
  auto g = GetGame(game);
  if (threads < 1) threads = 1;
  std::atomic<int64_t> next{0};
  std::atomic<int> bad{0};
  auto work = [&]() {
    std::vector<int64_t> acts(max_steps > 0 ? max_steps : 1);
    for (;;) {
      int64_t i = next.fetch_add(1);
      if (i >= n_roots) break;
      const int64_t* h = hist_flat + hist_off[i];
      int n = static_cast<int>(hist_off[i + 1] - hist_off[i]);
      for (int j = 0; j < rollouts; ++j) {
        int64_t p = i * rollouts + j;
        int len, term, ret0;
        if (ref_replay(game, h, n, nullptr, seed, first_playout + p, max_steps, acts.data(),
                       &len, &term, &ret0) != 0) {
          bad.store(1);
          continue;
        }
        len_out[p] = len;
        ret0_out[p] = static_cast<int8_t>(ret0);
        terminal_out[p] = static_cast<uint8_t>(term);
        if (actions_out) {
          for (int t = 0; t < len && t < actions_stride; ++t)
            actions_out[p * actions_stride + t] = static_cast<int16_t>(acts[t]);
        }
      }
    }
  };
  std::vector<std::thread> pool;
  for (int k = 0; k < threads; ++k) pool.emplace_back(work);
  for (auto& th : pool) th.join();
  return bad.load() ? -1 : 0;
}

// The same replay for roots given as a dense int16 action log (row i holds the root_len[i]
// moves that lead to root i) — what the device root generator returns — so that BASELINE-size
// batches (16 Mi roots) cross without a ragged int64 copy.  Per playout: moves played,
// (int)Returns()[0], IsTerminal().  One std::thread per requested core, playouts handed out in
// blocks.
int ref_replay_log(const char* game, const int16_t* log, int log_stride, const int32_t* root_len,
                   int64_t n_roots, int rollouts, uint64_t seed, uint64_t first_playout,
                   int threads, int32_t* len_out, int8_t* ret0_out, uint8_t* terminal_out) {
  auto g = GetGame(game);
  if (threads < 1) threads = 1;
  std::atomic<int64_t> next{0};
  std::atomic<int> bad{0};
  const int64_t kBlock = 256;
  auto work = [&]() {
    for (;;) {
      const int64_t lo = next.fetch_add(kBlock);
      if (lo >= n_roots) break;
      const int64_t hi = std::min(lo + kBlock, n_roots);
      for (int64_t i = lo; i < hi; ++i) {
        auto root = g->NewInitialState();
        const int16_t* h = log + i * log_stride;
        for (int k = 0; k < root_len[i]; ++k) root->ApplyAction(h[k]);
        for (int j = 0; j < rollouts; ++j) {
          const int64_t p = i * rollouts + j;
          std::unique_ptr<open_spiel::State> s = rollouts == 1 ? std::move(root) : root->Clone();
          int t = 0;
          while (!s->IsTerminal()) {
            std::vector<open_spiel::Action> legal = s->LegalActions();
            if (legal.empty()) { bad.store(1); break; }
            const uint32_t r = oracle_draw(seed, first_playout + (uint64_t)p, (uint32_t)t);
            s->ApplyAction(legal[r % legal.size()]);
            ++t;
          }
          len_out[p] = t;
          ret0_out[p] = static_cast<int8_t>(s->Returns()[0]);
          terminal_out[p] = s->IsTerminal() ? 1 : 0;
        }
      }
    }
  };
  std::vector<std::thread> pool;
  for (int k = 0; k < threads; ++k) pool.emplace_back(work);
  for (auto& th : pool) th.join();
  return bad.load() ? -1 : 0;
}

// State dumps of many states given as a dense int16 action log: ObservationTensor(players[i])
// as bytes (the reference's doubles are 0.0 / 1.0), LegalActionsMask as bytes, CurrentPlayer,
// IsTerminal.  Any out pointer may be NULL.
int ref_state_dump_log(const char* game, const int16_t* log, int log_stride, const int32_t* hist_len,
                       int64_t n, const uint8_t* players, int threads, uint8_t* obs_out,
                       uint8_t* mask_out, int32_t* cur_player_out, uint8_t* terminal_out) {
  auto g = GetGame(game);
  const int obs_size = g->ObservationTensorSize();
  const int n_actions = g->NumDistinctActions();
  if (threads < 1) threads = 1;
  std::atomic<int64_t> next{0};
  std::atomic<int> bad{0};
  auto work = [&]() {
    std::vector<double> obs;
    for (;;) {
      const int64_t lo = next.fetch_add(64);
      if (lo >= n) break;
      const int64_t hi = std::min(lo + (int64_t)64, n);
      for (int64_t i = lo; i < hi; ++i) {
        auto s = g->NewInitialState();
        const int16_t* h = log + i * log_stride;
        for (int k = 0; k < hist_len[i]; ++k) s->ApplyAction(h[k]);
        if (obs_out) {
          s->ObservationTensor(players ? players[i] : 0, &obs);
          for (int e = 0; e < obs_size; ++e) {
            if (obs[e] != 0.0 && obs[e] != 1.0) bad.store(1);
            obs_out[i * obs_size + e] = obs[e] != 0.0 ? 1 : 0;
          }
        }
        if (mask_out) {
          std::memset(mask_out + i * n_actions, 0, n_actions);
          if (!s->IsTerminal())
            for (open_spiel::Action a : s->LegalActions()) mask_out[i * n_actions + a] = 1;
        }
        if (cur_player_out) cur_player_out[i] = s->CurrentPlayer();
        if (terminal_out) terminal_out[i] = s->IsTerminal() ? 1 : 0;
      }
    }
  };
  std::vector<std::thread> pool;
  for (int k = 0; k < threads; ++k) pool.emplace_back(work);
  for (auto& th : pool) th.join();
  return bad.load() ? -1 : 0;
}

// CPU baseline, shaped like open_spiel/examples/benchmark_game.cc:31-97: from each root
// (history), Clone-free replay of LegalActions -> uniform pick -> ApplyAction until terminal,
// optionally filling ObservationTensor(CurrentPlayer()) every step (benchmark_game.cc:46-47).
// One std::thread per requested core, each with its own std::mt19937 (as the reference's
// RandomRolloutEvaluator owns one, mcts.h:105).  Roots are pre-built State objects and every
// playout starts with State::Clone(), the reference's own rollout entry (mcts.cc:45).
// Returns wall seconds; totals through the out pointers.
double ref_random_playouts(const char* game, const int64_t* hist_flat, const int64_t* hist_off,
                           int64_t n_roots, int64_t n_playouts, uint64_t seed, int threads,
                           int with_obs, int64_t* total_steps_out, int64_t* sum_ret0_out) {
  auto g = GetGame(game);
  std::vector<std::unique_ptr<open_spiel::State>> roots;
  roots.reserve(n_roots);
  for (int64_t i = 0; i < n_roots; ++i) {
    roots.push_back(FromHistory(*g, hist_flat + hist_off[i],
                                static_cast<int>(hist_off[i + 1] - hist_off[i])));
  }
  if (threads < 1) threads = 1;
  std::atomic<int64_t> next{0};
  std::atomic<int64_t> steps{0}, sum_ret{0};
  const int64_t kChunk = 64;
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&](int tid) {
    std::mt19937 rng(static_cast<uint32_t>(seed) + 7919u * tid);
    std::vector<double> obs;
    int64_t my_steps = 0, my_ret = 0;
    for (;;) {
      int64_t base = next.fetch_add(kChunk);
      if (base >= n_playouts) break;
      int64_t end = std::min(base + kChunk, n_playouts);
      for (int64_t p = base; p < end; ++p) {
        std::unique_ptr<open_spiel::State> s = roots[p % n_roots]->Clone();
        while (!s->IsTerminal()) {
          if (with_obs) s->ObservationTensor(s->CurrentPlayer(), &obs);
          std::vector<open_spiel::Action> legal = s->LegalActions();
          std::uniform_int_distribution<int> dis(0, static_cast<int>(legal.size()) - 1);
          s->ApplyAction(legal[dis(rng)]);
          ++my_steps;
        }
        my_ret += static_cast<int64_t>(s->Returns()[0]);
      }
    }
    steps += my_steps;
    sum_ret += my_ret;
  };
  std::vector<std::thread> pool;
  for (int k = 0; k < threads; ++k) pool.emplace_back(work, k);
  for (auto& th : pool) th.join();
  double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  *total_steps_out = steps.load();
  *sum_ret0_out = sum_ret.load();
  return secs;
}

// BASELINE config 1: the reference C++ MCTSBot + RandomRolloutEvaluator (mcts.cc) playing
// `n_games` self-play games per thread.  Uses MCTSBot::MCTSearch (mcts.h:192) so the real
// number of simulations (root explore_count; the solver may stop early, mcts.cc:398-401) is
// counted.  Returns wall seconds.
double ref_mcts_selfplay(const char* game, int n_games, int max_sims, int n_rollouts,
                         double uct_c, int solve, uint64_t seed, int threads,
                         int64_t* total_moves_out, int64_t* total_sims_out,
                         int64_t* sum_ret0_out) {
  auto g = GetGame(game);
  if (threads < 1) threads = 1;
  std::atomic<int64_t> moves{0}, sims{0}, sum_ret{0};
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&](int tid) {
    auto evaluator = std::make_shared<open_spiel::algorithms::RandomRolloutEvaluator>(
        n_rollouts, static_cast<int>(seed) + tid);
    open_spiel::algorithms::MCTSBot bot(*g, evaluator, uct_c, max_sims,
                                        /*max_memory_mb=*/1000, solve != 0,
                                        static_cast<int>(seed) + 101 * tid + 1,
                                        /*verbose=*/false);
    int64_t my_moves = 0, my_sims = 0, my_ret = 0;
    for (int i = 0; i < n_games; ++i) {
      auto s = g->NewInitialState();
      while (!s->IsTerminal()) {
        std::unique_ptr<open_spiel::algorithms::SearchNode> root = bot.MCTSearch(*s);
        my_sims += root->explore_count;
        s->ApplyAction(root->BestChild().action);
        ++my_moves;
      }
      my_ret += static_cast<int64_t>(s->Returns()[0]);
    }
    moves += my_moves;
    sims += my_sims;
    sum_ret += my_ret;
  };
  std::vector<std::thread> pool;
  for (int k = 0; k < threads; ++k) pool.emplace_back(work, k);
  for (auto& th : pool) th.join();
  double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  *total_moves_out = moves.load();
  *total_sims_out = sims.load();
  *sum_ret0_out = sum_ret.load();
  return secs;
}

// Text formats: State::ToString / State::Serialize of the state reached by `hist`
// (spiel.h:228, spiel.cc:375-384 and the per-game overrides).  Returns the length written.
int ref_to_string(const char* game, const int64_t* hist, int n_hist, int serialize, char* buf,
                  int capacity) {
  auto g = GetGame(game);
  auto s = FromHistory(*g, hist, n_hist);
  std::string str = serialize ? s->Serialize() : s->ToString();
  if (static_cast<int>(str.size()) + 1 > capacity) return -1;
  std::memcpy(buf, str.c_str(), str.size() + 1);
  return static_cast<int>(str.size());
}

// Game::DeserializeState(str) followed by the same dump as ref_state_dump, plus the
// deserialised state's own ToString().
int ref_deserialize_dump(const char* game, const char* str, int player, double* obs_out,
                         int64_t* legal_out, int* n_legal, int* cur_player, int* terminal,
                         double* returns_out, char* text_out, int text_capacity) {
  auto g = GetGame(game);
  std::unique_ptr<open_spiel::State> s = g->DeserializeState(str);
  std::vector<open_spiel::Action> legal = s->LegalActions();
  *n_legal = static_cast<int>(legal.size());
  if (legal_out) std::copy(legal.begin(), legal.end(), legal_out);
  *cur_player = s->CurrentPlayer();
  *terminal = s->IsTerminal() ? 1 : 0;
  std::vector<double> ret = s->Returns();
  returns_out[0] = ret[0];
  returns_out[1] = ret[1];
  if (obs_out && player >= 0) {
    std::vector<double> obs;
    s->ObservationTensor(player, &obs);
    std::copy(obs.begin(), obs.end(), obs_out);
  }
  std::string text = s->ToString();
  if (text_out && static_cast<int>(text.size()) + 1 <= text_capacity)
    std::memcpy(text_out, text.c_str(), text.size() + 1);
  return 0;
}

// SerializeGameAndState (spiel.cc:428-447) of the state reached by `hist`.
int ref_serialize_game_and_state(const char* game, const int64_t* hist, int n_hist, char* buf,
                                 int capacity) {
  auto g = GetGame(game);
  auto s = FromHistory(*g, hist, n_hist);
  std::string str = open_spiel::SerializeGameAndState(*g, *s);
  if (static_cast<int>(str.size()) + 1 > capacity) return -1;
  std::memcpy(buf, str.c_str(), str.size() + 1);
  return static_cast<int>(str.size());
}

// One MCTS search from a history: returns the chosen action (for solver known-answer tests).
int64_t ref_mcts_step(const char* game, const int64_t* hist, int n_hist, int max_sims,
                      int n_rollouts, double uct_c, int solve, int seed) {
  auto g = GetGame(game);
  auto s = FromHistory(*g, hist, n_hist);
  auto evaluator =
      std::make_shared<open_spiel::algorithms::RandomRolloutEvaluator>(n_rollouts, seed);
  open_spiel::algorithms::MCTSBot bot(*g, evaluator, uct_c, max_sims, 1000, solve != 0, seed + 1,
                                      false);
  return bot.Step(*s);
}

}  // extern "C"
