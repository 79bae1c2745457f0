/* TEST INFRASTRUCTURE ONLY — CPU oracle for the rollout hot path.
 *
 * A plain-C restatement of the reference's game rules (cell arrays, linear scans — the
 * reference's own algorithm, NOT the bitboard algorithm the CUDA engine uses), so that the
 * two implementations are independent.  Parity is PINNED: tests/test_oracle_golden.py checks
 * this file against the reference's playthrough files and known-answer tests (fixtures under
 * tests/golden/), and tests/test_oracle_vs_reference.py checks it against the unmodified
 * reference compiled into oracle/_ref/libspiel_ref.so.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this.  The product path never does.
 */
#ifndef BSG_ORACLE_H_
#define BSG_ORACLE_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_TTT = 0, ORC_C4 = 1, ORC_BT = 2, ORC_HEX = 3, ORC_PENTAGO = 5, ORC_OTHELLO = 6, ORC_Y = 7, ORC_HAVANNAH = 9, ORC_GO = 11 };
#define ORC_MAX_LEGAL 320 /* pentago: 36 cells x 8 rotations = 288 */
#define ORC_MAX_CELLS 361 /* y: 19 x 19 grid (y.h:37) */
#define ORC_TERMINAL_PLAYER (-4) /* kTerminalPlayerId, open_spiel/spiel.h:48 */

typedef struct {
  int game;
  int cells[ORC_MAX_CELLS]; /* per-game cell codes, see bsg_oracle.c */
  int cur_player;           /* side to move (not masked by terminal) */
  int outcome;              /* -1 none/unknown, 0 player 0 won, 1 player 1 won, 2 draw */
  int num_moves;
  int pieces[2];            /* breakthrough piece counters */
  int board_size;           /* hex N (11 default; 5 for the playthrough); y N (19 default; 9); havannah radius (8 default; 4); go N (9; 7) */
  /* y: the union-find embedded in the cell array (y.h:96-117) */
  int16_t parent[ORC_MAX_CELLS], size[ORC_MAX_CELLS];
  uint8_t edge[ORC_MAX_CELLS];
  /* havannah: corner sets at the leaders and the ring search's marks (havannah.h:96-131) */
  uint8_t corner[ORC_MAX_CELLS], mark[ORC_MAX_CELLS];
  /* go (board sizes up to 9): komi * 2, ko point (-1: none), the last two actions, superko flag,
   * and every earlier position of the game, cell by cell (go.cc:204-211 keeps hashes of them) */
  int komi2, ko_point, last_action, prev_action, superko;
  uint8_t past[2 * 81 + 1][81];
} orc_state;

int orc_num_distinct_actions(int game, int board_size);
int orc_max_game_length(int game, int board_size);
int orc_obs_size(int game, int board_size);

void orc_init(orc_state* s, int game, int board_size);
int orc_is_terminal(const orc_state* s);
int orc_current_player(const orc_state* s);
/* Writes legal actions ascending into out (capacity >= ORC_MAX_LEGAL); returns the count. */
int orc_legal_actions(const orc_state* s, int* out);
/* Returns 0, or -1 when the action is illegal. */
int orc_apply(orc_state* s, int action);
void orc_returns(const orc_state* s, int ret[2]);
/* Flat CHW observation as 0/1 bytes, reference element order. */
void orc_observation(const orc_state* s, int player, uint8_t* out);

/* Replay contract: a_t = legal[r_t mod n]; r_t from r_table or from Philox(seed, playout, t). */
int orc_replay(int game, int board_size, const int64_t* hist, int n_hist, const uint32_t* r_table,
               uint64_t seed, uint64_t playout, int max_steps, int64_t* actions_out, int* len_out,
               int* terminal_out, int* ret0_out);

int orc_replay_batch(int game, int board_size, const int64_t* hist_flat, const int64_t* hist_off,
                     int64_t n_roots, int rollouts, uint64_t seed, uint64_t first_playout,
                     int max_steps, int32_t* len_out, int8_t* ret0_out, uint8_t* terminal_out,
                     int16_t* actions_out, int actions_stride);

/* State dump mirroring ref_state_dump in ref_wrapper.cc. */
int orc_state_dump(int game, int board_size, const int64_t* hist, int n_hist, int player,
                   uint8_t* obs_out, int64_t* legal_out, int* n_legal, int* cur_player,
                   int* terminal, int* returns_out);

/* Timed scalar playout loop (bench.py cpu_baseline kind "port"); returns total steps. */
int64_t orc_random_playouts(int game, int board_size, const int64_t* hist_flat,
                            const int64_t* hist_off, int64_t n_roots, int64_t n_playouts,
                            uint64_t seed, int64_t* sum_ret0_out);

uint32_t orc_draw(uint64_t seed, uint64_t playout, uint32_t t);

#ifdef __cplusplus
}
#endif
#endif
