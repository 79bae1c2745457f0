/* bsg.h — C ABI of the B200 batched game-simulation engine ("bsg").
 *
 * Drop-in boundary for OpenSpiel's rollout hot path.  The reference has no C ABI of its own
 * (its FFI is the pybind11 module open_spiel/python/pybind11/pyspiel.cc); each entry point
 * below names the reference interface it stands in for.  Plain pointers and sizes only; no
 * C++/torch types.  Every function returns BSG_OK (0) or a negative error code and never
 * throws or exits (the reference's SpielFatalError -> RuntimeError convention,
 * pyspiel.cc:71-78,839-840, is reproduced by the Python wrapper).
 *
 * Ownership: the caller owns every buffer.  "dev" pointers are CUDA device pointers on the
 * engine's device (e.g. torch tensor data_ptr()); "host" pointers are ordinary memory.
 * `stream` is a cudaStream_t passed as void* (NULL = default stream); all device work is
 * stream-ordered on it and the call returns without synchronising unless stated.
 * Threading: one engine per (thread, GPU); calls on one engine must be serialised by the
 * caller — the same contract as the reference evaluators, which own a mutable RNG
 * (open_spiel/algorithms/mcts.h:105, python/algorithms/mcts.py:56).
 *
 * Replay contract (bit-exact parity with the reference):
 *   r(seed, playout, t) = philox4x32_10(ctr = {playout_lo, playout_hi, t >> 2, 0},
 *                                       key = {seed_lo, seed_hi})[t & 3]
 *   action_t = State::LegalActions()[ r mod n ]        (ascending action ids)
 * or, when io.r_table is given, r = r_table[playout_local * r_stride + t].
 */
#ifndef BSG_H_
#define BSG_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BSG_VERSION 1

/* Game ids.  Strings accepted by bsg_game_id are the reference's game strings
 * (open_spiel/spiel.h:792-803): "tic_tac_toe", "connect_four", "breakthrough",
 * "breakthrough(rows=8,columns=8)", "hex", "hex(board_size=11)", "pentago", "othello", "y",
 * "y(board_size=19)", "havannah", "havannah(board_size=8)", "go(board_size=9)". */
#define BSG_TIC_TAC_TOE 0
#define BSG_CONNECT_FOUR 1
#define BSG_BREAKTHROUGH 2
#define BSG_HEX 3
/* Host-helper-only id used to check the hex logic against the reference's 5x5 golden
 * playthrough; no device kernels exist for it. */
#define BSG_HEX5_HOST_ONLY 4
/* open_spiel/games/pentago.cc — the first game of SURVEY.md section 8f row 4 (more
 * perfect-information board games on the same kernels). */
#define BSG_PENTAGO 5
#define BSG_OTHELLO 6 /* open_spiel/games/othello.cc, same row */
#define BSG_Y 7       /* open_spiel/games/y.cc, board_size = 19 (the reference's default), same row */
/* Host-helper-only id for the reference's y(board_size=9) golden playthrough; no kernels. */
#define BSG_Y9_HOST_ONLY 8
#define BSG_HAVANNAH 9 /* open_spiel/games/havannah.cc, board_size = 8 (the reference's default), same row */
/* Host-helper-only id for the reference's havannah(board_size=4) golden playthrough; no kernels. */
#define BSG_HAVANNAH4_HOST_ONLY 10
/* open_spiel/games/go.cc at board_size = 9, komi 7.5, no handicap (same row).  The reference's
 * default board is 19 x 19: its games run to 722 moves (the playout ABI counts moves in uint8) and
 * its pseudo-liberty sums are 16-bit, exact only on small boards — see csrc/bsg_go.cuh. */
#define BSG_GO 11
/* Host-helper-only id for the reference's go(board_size=7,komi=4.5) golden playthrough. */
#define BSG_GO7_HOST_ONLY 12

#define BSG_OK 0
#define BSG_ERR_BAD_GAME (-1)
#define BSG_ERR_BAD_ARG (-2)
#define BSG_ERR_ILLEGAL_ACTION (-3)
#define BSG_ERR_CUDA (-4)
#define BSG_ERR_NO_DEVICE (-5)
#define BSG_ERR_STUCK_LANE (-6) /* a non-terminal state without legal actions was met */
#define BSG_ERR_NCCL (-7)       /* libnccl could not be loaded, or an NCCL call failed */

#define BSG_TERMINAL_PLAYER (-4) /* kTerminalPlayerId, open_spiel/spiel.h:48 */

typedef struct bsg_engine bsg_engine;

int bsg_version(void);
const char* bsg_error_string(int code);

/* ---- static game facts: Game::NumDistinctActions / MaxGameLength / ObservationTensorSize
 *      (open_spiel/spiel.h:601-750), plus the packed-state size of this engine. */
int bsg_game_id(const char* game_string);
int bsg_num_distinct_actions(int game);
int bsg_max_game_length(int game);
int bsg_observation_size(int game);
int bsg_observation_shape(int game, int shape_out[3]);
int bsg_state_bytes(int game);
/* Element sizes of the packed structure-of-arrays planes of a game (their sum is
 * bsg_state_bytes): fills bytes_out[0..count) and returns the count, or BSG_ERR_BAD_ARG when
 * `capacity` is too small.  What a caller needs to slice or interleave packed buffers. */
int bsg_plane_bytes(int game, int* bytes_out, int capacity);

/* ---- host helpers (pure CPU, usable without a GPU): moving reference states in and out.
 * Packed states are plane-major structure-of-arrays buffers with a capacity `cap`:
 * plane p of state i lives at base + plane_offset[p] * cap + i * plane_elem[p]
 * (layouts per game in DESIGN.md).  A single state is the cap = 1 case. */

/* Replays action histories (State::history(), pyspiel.cc:322) from the initial state into
 * packed states; stands in for Game::NewInitialState + State::ApplyAction
 * (open_spiel/spiel.h:205-210,622).  hist_off has n + 1 entries. */
int bsg_pack_histories(int game, const int64_t* hist_flat, const int64_t* hist_off, int64_t n,
                       void* packed_host, int64_t cap);
/* State::CurrentPlayer / IsTerminal / Returns()[0] of packed state i. */
int bsg_host_state_info(int game, const void* packed_host, int64_t cap, int64_t i,
                        int* current_player, int* is_terminal, int* ret0);
/* State::LegalActions (ascending).  Returns the count (>= 0) or an error. */
int bsg_host_legal_actions(int game, const void* packed_host, int64_t cap, int64_t i,
                           int64_t* actions_out, int capacity);
/* State::ApplyAction on packed state i, in place. */
int bsg_host_apply_action(int game, void* packed_host, int64_t cap, int64_t i, int64_t action);
/* State::ObservationTensor(player) of one packed state as float 0/1, flat CHW. */
int bsg_host_observation(int game, const void* packed_host, int64_t cap, int64_t i, int player,
                         float* obs_out);
/* State::LegalActionsMask (spiel.h:244-252) of one packed state: NumDistinctActions bytes. */
int bsg_host_legal_mask(int game, const void* packed_host, int64_t cap, int64_t i,
                        uint8_t* mask_out);
/* State::ToString (open_spiel/spiel.h:228; per game tic_tac_toe.cc:126-137, connect_four.cc:166-175,
 * breakthrough.cc:262-282, hex.cc:229-246) of packed state i into buf (NUL-terminated).
 * Returns the length, or BSG_ERR_BAD_ARG when `capacity` is too small. */
int bsg_host_to_string(int game, const void* packed_host, int64_t cap, int64_t i, char* buf,
                       int capacity);
/* State::Serialize for the games whose serialisation is the board (connect_four.cc:227 =
 * ToString(); breakthrough.cc:389-397 = 64 cell characters).  tic_tac_toe and hex serialise
 * as their action history (spiel.cc:375-384), which a packed state does not carry:
 * BSG_ERR_BAD_ARG (the Python mirror writes those from State.history()). */
int bsg_host_serialize(int game, const void* packed_host, int64_t cap, int64_t i, char* buf,
                       int capacity);
/* Game::DeserializeState for the board formats (connect_four.cc:229-279: side to move from the
 * stone counts, outcome from the lines; breakthrough.cc:399-435: side to move stays black, no
 * winner is set — the reference's own quirk, reproduced).  Histories deserialise through
 * bsg_pack_histories. */
int bsg_host_deserialize(int game, const char* str, void* packed_host, int64_t cap, int64_t i);
/* Debug/test hook: applies LegalActions()[r mod n] to packed state i through the very step
 * function the kernels run, so the selection logic can be checked on a CPU.  No product
 * path calls it; playouts exist only as CUDA kernels. */
int bsg_host_replay_step(int game, void* packed_host, int64_t cap, int64_t i, uint32_t r,
                         int64_t* action_out);
/* Debug/test hook for value-only playouts (bsg_playout_io.value_only): plays packed state i on
 * with the draws r(seed, playout, t), t = 0, 1, ..., through the kernels' step_fill until the
 * board is full and reads the winner off it with the kernels' fill_winner.  *ret0_out is the
 * return to player 0, *moves_out the moves played.  hex / y only.  No product path calls it. */
int bsg_host_fill_playout(int game, const void* packed_host, int64_t cap, int64_t i, uint64_t seed,
                          uint64_t playout, int* ret0_out, int* moves_out);
/* The t-th replay draw r(seed, playout, t). */
uint32_t bsg_draw(uint64_t seed, uint64_t playout, uint32_t t);

/* ---- device engine */
int bsg_create(int game, int device, bsg_engine** out);
void bsg_destroy(bsg_engine* e);
const char* bsg_last_error(bsg_engine* e);
int bsg_engine_game(const bsg_engine* e);
int bsg_engine_device(const bsg_engine* e);
int bsg_engine_sm_count(const bsg_engine* e);
/* on != 0: the host-buffer playouts of this engine (bsg_playouts_host) run value-only (see
 * bsg_playout_io.value_only below): same returns, fewer instructions per move, move counts to
 * the full board.  hex / y only (BSG_ERR_BAD_ARG otherwise).  Value-only launches of one engine
 * share one scratch buffer: keep them stream-ordered. */
int bsg_engine_set_value_only(bsg_engine* e, int on);

/* Optional inputs/outputs of one playout launch; all pointers are device pointers and may be
 * NULL.  Playout p (0 <= p < n_roots * rollouts_per_root) starts from root p / rollouts. */
typedef struct bsg_playout_io {
  int32_t max_steps;            /* <= 0: play until terminal */
  const uint8_t* step_limit;    /* per-playout step limit [n_playouts] (0 = no moves) */
  const uint32_t* r_table;      /* pre-generated draws [n_playouts, r_stride] instead of Philox */
  int32_t r_stride;
  int8_t* ret0;                 /* per playout: (int)Returns()[0] at the stop state */
  uint8_t* length;              /* per playout: moves played */
  uint8_t* terminal;            /* per playout: IsTerminal() at the stop state */
  int16_t* actions;             /* action log [n_playouts, actions_stride] */
  int32_t actions_stride;
  int32_t* root_sum_ret0;       /* per root: += sum of ret0 (caller zeroes) */
  int32_t* root_sum_steps;      /* per root: += sum of moves played (caller zeroes) */
  void* final_states;           /* packed SoA (capacity n_playouts) of the stop states */
  int32_t final_before_terminal;/* != 0: a playout that ended writes its last non-terminal state */
  uint64_t* totals;             /* u64[4] += {playouts, steps, sum ret0 (two's complement),
                                   stuck lanes}; caller zeroes */
  /* Trajectory log (the AlphaZero actor's record of every decision point, alpha_zero.py:196-228):
   * the state BEFORE move t of playout p is written as packed state number
   * state_log_off[p] + t of a packed SoA buffer with capacity state_log_cap.  Offsets are the
   * exclusive prefix sum of the lengths of an earlier identical launch (playouts are
   * deterministic), so the log is dense; bsg_observations turns it into observation planes
   * and legal masks. */
  void* state_log;
  int64_t state_log_cap;
  const int64_t* state_log_off;
  /* Per-root playout ids: when non-NULL, rollout j of root i draws with the global id
   * root_playout_base[i] + j instead of first_playout + i * rollouts + j, and a base of
   * UINT64_MAX skips the root (no playout; nothing is added to its sums).  Lets many search
   * trees, each with its own id range, share one launch (bsg_forest_*). */
  const uint64_t* root_playout_base;
  /* Value-only playouts, for the games whose full board has exactly one winner and whose stones
   * never leave the board (hex, y; BSG_ERR_BAD_ARG otherwise): the same draws and the same
   * moves are played on until the board is full — no connectivity bookkeeping per move — and a
   * second kernel reads the winner off the full board.  ret0, root_sum_ret0 and totals[2] are
   * identical to an exact launch (the side that connects first is the side connected at the
   * end); length, root_sum_steps and totals[1] count the moves to the full board instead of to
   * the win.  What RandomRolloutEvaluator::Evaluate (mcts.cc:42-71) needs is only the returns.
   * Excludes max_steps / step_limit / r_table / actions / final_states / state_log. */
  int32_t value_only;
  /* Alternative to state_log_off: a fixed number of moves per playout — the state before move t
   * of playout p goes to slot t * n_playouts + p (move-major: lanes at the same move number write
   * adjacent elements) of a log with capacity n_playouts * state_log_stride; state_log_off must
   * then be NULL and moves t >= state_log_stride are not logged.  One pass, no earlier launch
   * needed; bsg_trajectory_* compacts the log. */
  int64_t state_log_stride;
} bsg_playout_io;

/* The hot path.  Stands in for RandomRolloutEvaluator::Evaluate's playout loop
 * (open_spiel/algorithms/mcts.cc:42-71; python/algorithms/mcts.py:58-74) and for the raw
 * loop of examples/benchmark_game.cc:31-97, for n_roots * rollouts_per_root playouts at once.
 * Draws of playout p use the global id first_playout + p. */
int bsg_playouts(bsg_engine* e, const void* roots_dev, int64_t n_roots, int32_t rollouts_per_root,
                 uint64_t seed, uint64_t first_playout, const bsg_playout_io* io, void* stream);

/* Batched State::ObservationTensor (open_spiel/spiel.h:438-453; per-game writers cited in
 * DESIGN.md): writes [n, C, H, W] float32 (0/1, reference element order) for n packed states.
 * players_dev[i] is the observer of state i; NULL means the player to move.  When
 * legal_mask_dev is non-NULL it also writes State::LegalActionsMask (spiel.h:244-252) as
 * [n, NumDistinctActions] bytes. */
int bsg_observations(bsg_engine* e, const void* states_dev, int64_t n, const uint8_t* players_dev,
                     float* obs_dev, uint8_t* legal_mask_dev, void* stream);

/* Host-buffer convenience wrappers (the reference-facing call with host memory: copies in,
 * launches, copies out, synchronises).  roots_host is a packed SoA buffer of capacity n_roots.
 * totals_out: {playouts, steps, sum ret0, stuck lanes}.  root_sum_* may be NULL. */
int bsg_playouts_host(bsg_engine* e, const void* roots_host, int64_t n_roots,
                      int32_t rollouts_per_root, uint64_t seed, uint64_t first_playout,
                      int64_t totals_out[4], int32_t* root_sum_ret0_host,
                      int32_t* root_sum_steps_host);
/* bsg_playouts_host with PER-PLAYOUT results in host memory: ret0_host[p] = (int)Returns()[0]
 * and length_host[p] = moves played by playout p (n_roots * rollouts_per_root entries each;
 * either may be NULL) — the 2 bytes per playout a consumer of every root's value reads, e.g.
 * RandomRolloutEvaluator.evaluate for a batch of leaves (python/algorithms/mcts.py:58-74) or a
 * data pipeline.  Results of a chunk are copied back behind its kernel while the next chunk's
 * roots are still arriving; page-locked caller buffers are used in place. */
int bsg_playouts_host_results(bsg_engine* e, const void* roots_host, int64_t n_roots,
                              int32_t rollouts_per_root, uint64_t seed, uint64_t first_playout,
                              int64_t totals_out[4], int8_t* ret0_host, uint8_t* length_host);
/* Compact root format for host -> device transfers.  The copy of the roots over PCIe is the
 * end-to-end bound of the path; a connect_four position needs 49 bits (seven rows of seven
 * columns: the mover's stones | one marker per column), so its roots also travel as 49-bit
 * strings back to back, root i at bit 49 * i of the buffer, little endian (empty-cell count and
 * win flag are derived on the device).  bsg_compact_state_bits: bits per compact root (= 8 *
 * bsg_state_bytes for a game without a compact format, whose "compact" buffers are the packed
 * ones); bsg_compact_buffer_bytes: bytes of a buffer of n compact roots, (n * bits + 7) / 8.
 * bsg_compact_roots / bsg_expand_roots convert on the host (expand runs the very decode
 * the kernel runs).  bsg_playouts_host_compact is bsg_playouts_host_results for compact roots;
 * ret0_host and length_host may both be NULL (totals only).  Stands in for the same reference
 * loop (algorithms/mcts.cc:42-71, examples/benchmark_game.cc:31-97). */
int bsg_compact_state_bits(int game);
int64_t bsg_compact_buffer_bytes(int game, int64_t n);
int bsg_compact_roots(int game, const void* packed_host, int64_t n, void* compact_out);
int bsg_expand_roots(int game, const void* compact_host, int64_t n, void* packed_out);
int bsg_playouts_host_compact(bsg_engine* e, const void* roots_compact_host, int64_t n_roots,
                              int32_t rollouts_per_root, uint64_t seed, uint64_t first_playout,
                              int64_t totals_out[4], int8_t* ret0_host, uint8_t* length_host);
int bsg_observations_host(bsg_engine* e, const void* states_host, int64_t n,
                          const uint8_t* players_host, float* obs_host,
                          uint8_t* legal_mask_host);

/* ---- native MCTS tree with batched leaf evaluation.
 * Stands in for MCTSBot::MCTSearch (open_spiel/algorithms/mcts.cc:255-428;
 * python/algorithms/mcts.py:347-443) when the evaluator is the rollout evaluator of this
 * engine: UCT descent (SearchNode::UCTValue, mcts.cc:95-105), uniform shuffled priors,
 * MCTS-Solver backups — on the host, as in the reference — with up to `leaf_batch` leaves
 * gathered by virtual visits and evaluated by ONE launch of the playout kernel
 * (`n_rollouts` playouts per leaf).  leaf_batch = 1 is the reference's one-leaf-per-call
 * order.  Playout ids advance exactly as in the Python GpuRolloutEvaluator (next_playout),
 * priors are shuffled by Fisher-Yates over bsg_draw(shuffle_seed, counter, 0): with the same
 * seeds the Python mirror (mcts.MCTSBot + PhiloxShuffle) builds the identical tree. */
typedef struct bsg_mcts bsg_mcts;
int bsg_mcts_create(bsg_engine* e, double uct_c, int solve, int n_rollouts, int leaf_batch,
                    uint64_t eval_seed, uint64_t shuffle_seed, bsg_mcts** out);
void bsg_mcts_destroy(bsg_mcts* m);
const char* bsg_mcts_last_error(bsg_mcts* m);
/* Runs `simulations` simulations from the packed root state (capacity-1 packed state on the
 * host).  resume != 0 continues the tree of the previous call (same root). */
int bsg_mcts_search(bsg_mcts* m, const void* root_packed_host, int64_t simulations, int resume);
/* Creates the root's children without a simulation (root-parallel merge on a rank with no budget). */
int bsg_mcts_expand_root(bsg_mcts* m, const void* root_packed_host);
/* Root statistics after a search: returns the number of root children (in the tree's shuffled
 * order) and fills up to `capacity` entries.  outcome[i] is the proven value for the root
 * mover (-1, 0, +1) or 2 when not proven. */
int bsg_mcts_root_children(bsg_mcts* m, int capacity, int64_t* actions, int64_t* visits,
                           double* total_reward, int32_t* outcome);
/* info_out: {root explore_count, root proven (0/1), root proven value for player 0, nodes,
 * game steps played by the evaluator, kernel launches, next playout id, root mover}. */
int bsg_mcts_info(bsg_mcts* m, double info_out[8]);
/* Overwrites the statistics of the root child with this action (root-parallel merge,
 * SURVEY.md section 8e); outcome as above.  root_explore < 0 leaves the root count alone. */
int bsg_mcts_set_root_child(bsg_mcts* m, int64_t action, int64_t visits, double total_reward,
                            int32_t outcome);
int bsg_mcts_set_root_explore(bsg_mcts* m, int64_t root_explore);
int bsg_mcts_set_next_playout(bsg_mcts* m, uint64_t next_playout);

/* ---- GPU-resident search forest ("then a GPU-resident tree").
 * n_trees independent MCTS trees in device memory, searched together: per round one kernel
 * descends every tree (up to leaf_batch leaves each), one launch of the playout kernel
 * evaluates all leaves of all trees, one kernel backs the results up.  Tree v runs exactly the
 * search of bsg_mcts_create(.., shuffle_seed0 + v) with playout ids from first_playout +
 * (v << 40): the algorithm of MCTSBot::MCTSearch (algorithms/mcts.cc:255-428,
 * python/algorithms/mcts.py:292-443), visit for visit.  node_capacity bounds the nodes per
 * tree, max_visits the visit count any node can reach (after root merges: the total of all
 * trees of all ranks). */
typedef struct bsg_forest bsg_forest;
int bsg_forest_create(bsg_engine* e, int32_t n_trees, int32_t node_capacity, int64_t max_visits,
                      double uct_c, int solve, int n_rollouts, int leaf_batch, uint64_t eval_seed,
                      uint64_t shuffle_seed0, uint64_t first_playout, bsg_forest** out);
void bsg_forest_destroy(bsg_forest* f);
const char* bsg_forest_last_error(bsg_forest* f);
/* Runs simulations_per_tree (or per_tree_host[v]) simulations in every tree from the packed
 * host root; resume != 0 continues the trees of the previous call.  A tree left without
 * children gets its root expanded (so that it joins a merge). */
int bsg_forest_search(bsg_forest* f, const void* root_host, int64_t simulations_per_tree,
                      const int64_t* per_tree_host, int resume);
/* Root merge of root-parallel search, dense per action id (device buffers of the caller):
 * base_dev int64[NumDistinctActions][2] = {visits, reward * n_rollouts} after the last merge;
 * pack adds into out_dev int64[NumDistinctActions][5] (caller zeroes) the sum over the trees of
 * {visit delta, scaled reward delta, proofs "mover loses / draws / wins"}; install writes
 * base + summed into every tree's root children (summed = out_dev after the caller's
 * all-reduce over ranks). */
int bsg_forest_pack_root(bsg_forest* f, const int64_t* base_dev, int64_t* out_dev);
int bsg_forest_install_root(bsg_forest* f, const int64_t* base_dev, const int64_t* summed_dev);
/* ---- trajectories of random playouts in ONE pass (the record of every decision point,
 * alpha_zero.py:196-228 / 214-217, for uniform-random self-play): bsg_playouts with
 * io.state_log_stride (and io.actions) logs at a fixed stride per playout; bsg_trajectory_scan
 * turns the lengths into dense offsets (off_dev has n + 1 entries, off[n] = total records; the
 * total is also returned through *total_out after a stream synchronisation);
 * bsg_trajectory_compact gathers record r = off[p] + t: packed state -> out_states (capacity
 * total, ready for bsg_observations), action, owner playout and move number (each may be NULL). */
int bsg_trajectory_scan(bsg_engine* e, const uint8_t* length_dev, int64_t n, int64_t* off_dev,
                        int64_t* total_out, void* stream);
int bsg_trajectory_compact(bsg_engine* e, const void* strided_states_dev, int64_t n, int64_t stride,
                           const int64_t* off_dev, int64_t total, const int16_t* strided_actions_dev,
                           int32_t actions_stride, void* out_states_dev, int64_t* out_actions_dev,
                           int64_t* out_playout_dev, int64_t* out_step_dev, void* stream);

/* ---- self-play actor on the forest: MANY games in flight, game v
 * = tree v.  Stands in for _play_game of python/algorithms/alpha_zero/alpha_zero.py:196-228 —
 * per move one search, policy target = root children's explore_count ** (1 / temperature)
 * normalised, action = best_child() (mcts.py:144-162) once `temperature_drop` moves were made,
 * else sampled from the policy; record = (state, action, policy, root.total_reward /
 * root.explore_count) — and for the learner's filing of the records with value = the game's
 * final returns[0] (alpha_zero.py:366-369).  Everything stays in device memory.
 *
 * bsg_forest_search_roots: searches tree v from root v of a packed DEVICE buffer (capacity
 * n_trees), `simulations_per_tree` each; trees whose alive_dev[v] is 0 (NULL = all alive) get
 * none.  Shuffle counters and playout ids of a tree run on from move to move, as a bot object
 * reused for a whole game does.
 * bsg_forest_actor_step: one move of every running game from the finished search (all pointers
 * device memory of the caller; logs are move-major, record (m, v) at m * n_games + v).  The
 * sampled action of move m of game v uses the draw r(sample_seed, game_id_base + v, m) through
 * the inverse CDF over ascending action ids; at temperature 1 exactly (integer visit counts).
 * *n_alive_out (host, may be NULL = no synchronisation) = games still running.
 * bsg_actor_emit: the records game-major (row offsets[v] + m), packed states compacted to
 * capacity `total` (feed them to bsg_observations for the planes and masks), value = final
 * returns[0] of the record's game. */
typedef struct bsg_actor_io {
  void* states;             /* packed SoA, capacity n_games: current positions, updated in place */
  uint8_t* alive;           /* [n_games] */
  int32_t* length;          /* [n_games] moves made */
  int32_t* outcome;         /* [n_games] (int)returns[0] of finished games */
  uint32_t* n_alive_dev;    /* one device counter */
  void* state_log;          /* packed SoA, capacity max_moves * n_games: the state before each move */
  float* policy_log;        /* [max_moves * n_games][NumDistinctActions] */
  float* est_log;           /* [max_moves * n_games] root value estimate */
  int32_t* action_log;      /* [max_moves * n_games] */
  double* weights;          /* scratch [n_games][NumDistinctActions] */
  int32_t n_games, max_moves, move, temperature_drop;
  double temperature;
  uint64_t sample_seed, game_id_base;
} bsg_actor_io;
typedef struct bsg_actor_emit_io {
  const void* state_log; const float* policy_log; const float* est_log; const int32_t* action_log;
  const int32_t* length; const int32_t* outcome;
  const int64_t* offsets;   /* [n_games] exclusive prefix sum of length */
  void* out_states;         /* packed SoA, capacity total */
  float* out_policy;        /* [total][NumDistinctActions] */
  float* out_value;         /* [total] */
  float* out_est;           /* [total] or NULL */
  int32_t* out_action;      /* [total] or NULL */
  int32_t* out_player;      /* [total] or NULL: current player of the record */
  int32_t n_games, max_moves;
  int64_t total;
} bsg_actor_emit_io;
int bsg_forest_search_roots(bsg_forest* f, const void* roots_dev, const uint8_t* alive_dev,
                            int64_t simulations_per_tree);
int bsg_forest_actor_step(bsg_forest* f, const bsg_actor_io* io, uint32_t* n_alive_out);
int bsg_actor_emit(bsg_engine* e, const bsg_actor_emit_io* io);
/* Page-locked host memory for root / result buffers of the *_host entry points (cudaHostAlloc,
 * portable).  write_combined != 0 asks for write-combined pages: the CPU must only WRITE them
 * (reads are uncached and very slow), in exchange the device's DMA reads skip cache snooping —
 * the right kind for roots that the host fills once and the GPU consumes.  NULL on failure. */
void* bsg_host_alloc(size_t bytes, int write_combined);
void bsg_host_free(void* p);
/* The engine's own CUDA stream (cudaStream_t as void*): where bsg_forest_* and bsg_actor_* work
 * is ordered, for callers that want to order their own work behind it. */
void* bsg_engine_stream(bsg_engine* e);

/* Root-parallel search over several GPUs without torch ("one
 * ncclAllReduce per search step").  The engine owns an NCCL communicator of its own; NCCL is
 * loaded with dlopen("libnccl.so.2") at the first call, so the library has no link-time
 * dependency on it.  bsg_comm_unique_id is ncclGetUniqueId (call it on one rank and hand the 128
 * bytes to the others by any means: MPI, a file, torch.distributed, a socket); bsg_comm_init is
 * ncclCommInitRank with the engine's device current (collective: every rank calls it).
 * bsg_comm_ranks is 1 without a communicator. */
int bsg_comm_unique_id(uint8_t id_out[128]);
int bsg_comm_init(bsg_engine* e, int n_ranks, int rank, const uint8_t id[128]);
int bsg_comm_destroy(bsg_engine* e);
int bsg_comm_ranks(const bsg_engine* e);
/* The merge of one search step (python/algorithms/mcts.py:116-122 statistics of the root's
 * children, summed over all trees of all ranks), stream-ordered on the engine's stream with no
 * host synchronisation: pack the deltas since the last merge, ONE int64 ncclAllReduce(sum) over
 * the engine's communicator (none = world size 1 = no collective), install the global
 * statistics into every tree.  first_step != 0 starts a new move (the base is zeroed).  The
 * forest keeps base and sum buffers itself; bsg_forest_pack_root / _install_root below remain
 * for callers that run the collective themselves. */
int bsg_forest_merge(bsg_forest* f, int first_step);
/* Synchronises and reports the all-reduces issued by bsg_forest_merge since the last call:
 * {count, mean microseconds, max microseconds}, timed with CUDA events on the engine's stream. */
int bsg_forest_allreduce_stats(bsg_forest* f, double stats_out[3]);
/* Root children of one tree (as bsg_mcts_root_children); info = {root visits, root solved,
 * root returns[0], nodes used}.  Returns the number of children. */
int bsg_forest_tree_root(bsg_forest* f, int32_t tree, int capacity, int64_t* actions, int64_t* visits,
                         double* reward, int32_t* outcome, double info[4]);
/* info = {game steps played, kernel launches, rounds, trees}. */
int bsg_forest_info(bsg_forest* f, double info[4]);
/* Test hook, no GPU needed and not used by any product path: runs the forest's per-tree code
 * on host memory next to the host tree of bsg_mcts_*, both fed stand-in leaf values hashed from
 * the playout ids, for `sims` simulations and a resumed search of `sims_resume` more, from the
 * state reached by `history`.  0 = every tree identical node for node; k > 0 = tree k - 1
 * differs; < 0 = error. */
int bsg_forest_selftest(int game, int n_trees, int sims, int sims_resume, int leaf_batch, int n_rollouts,
                        int solve, uint64_t seed, const int64_t* history, int n_history);

/* Value-only playouts (bsg_playout_io.value_only above: hex, y) that ALLOCATE NOTHING.  No
 * counterpart in the reference; added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The launch's scratch — its work counter and one parked board per playout (16 B hex, 48 B y) —
 * lives in a file-scope `static __device__` array of 16 MiB (+ 256 B) that every GPU loading the
 * library holds until exit, so one call runs at most bsg_noalloc_max_playouts(game) playouts
 * (hex 1 048 576, y 349 525); more is refused with BSG_ERR_BAD_ARG and a message in
 * bsg_last_error.  Same results as bsg_playouts with io.value_only = 1: ret0_dev [n_playouts],
 * length_dev [n_playouts] (moves to the full board), root_sum_*_dev [n_roots] (+=, caller
 * zeroes), totals_dev u64[4] (+=, caller zeroes); each may be NULL.  Calls no cudaMalloc /
 * cudaFree / allocating library routine; stream-ordered on `stream`, no synchronisation.  Every
 * call shares the one array: calls that overlap on the GPU (two streams, two engines of one
 * device) race — keep them stream-ordered. */
int bsg_noalloc_max_playouts(int game);
int bsg_playouts_value_only_noalloc(bsg_engine* e, const void* roots_dev, int64_t n_roots,
                                    int32_t rollouts_per_root, uint64_t seed, uint64_t first_playout,
                                    int8_t* ret0_dev, uint8_t* length_dev, int32_t* root_sum_ret0_dev,
                                    int32_t* root_sum_steps_dev, uint64_t* totals_dev, void* stream);

/* Integer-issue microbenchmark used for the roofline denominator (bench.py): runs `iters`
 * dependent-chain iterations of `ilp` independent 32-bit ops per thread on every SM and
 * returns thread-instructions per second.  kind: 0 = LOP3/IADD3 mix (ALU pipe), 1 = IMAD
 * (FMA pipe), 2 = SHF + LOP3 + IMAD (2 ALU : 1 FMA), 3 = LOP3 + IMAD (1 : 1, both pipes busy), 4 = IMAD.HI, 5 = IMAD.WIDE + LOP3,
 * 6 = LOP3 + IMAD.HI (1 : 1). */
int bsg_int_peak(bsg_engine* e, int kind, int iters, double* thread_instr_per_s);

#ifdef __cplusplus
}
#endif
#endif /* BSG_H_ */
