/* TEST INFRASTRUCTURE ONLY — see bsg_oracle.h.  Plain-C restatement of the reference rules.
 *
 * Cell codes follow the reference enums so that observation planes index the same way:
 *   tic_tac_toe / connect_four: 0 empty, 1 nought (player 1), 2 cross (player 0)
 *       (open_spiel/games/tic_tac_toe.h:44-48, connect_four.h:54-58)
 *   breakthrough: 0 empty, 1 black (player 0), 2 white (player 1)  (breakthrough.h:45-49)
 *   hex: the signed 9-value enum of hex.h:51-61 (-4..4)
 *   othello: 0 empty, 1 black (player 0), 2 white (player 1)  (othello.h:40-44)
 *   y: 0 player 0 (kPlayer1), 1 player 1 (kPlayer2), 2 empty (kPlayerNone), 3 off the board
 *       (kPlayerInvalid)  (y.h:43-48)
 *   pentago: 0 empty, 1 player 0, 2 player 1 (pentago.h:44-49 orders kPlayer1, kPlayer2, kPlayerNone;
 *       only the observation plane order depends on it and that is player-relative)
 */
#include "bsg_oracle.h"

#include <string.h>

#include "philox_ref.h"

enum { E = 0, O = 1, X = 2 };           /* ttt / c4 cell codes */
enum { BT_E = 0, BT_B = 1, BT_W = 2 };  /* breakthrough cell codes */
enum {                                   /* hex.h:51-61 */
  HX_EMPTY = 0, HX_WWEST = -3, HX_WEAST = -2, HX_WWIN = -4, HX_WHITE = -1,
  HX_BNORTH = 3, HX_BSOUTH = 2, HX_BWIN = 4, HX_BLACK = 1
};

uint32_t orc_draw(uint64_t seed, uint64_t playout, uint32_t t) {
  return oracle_draw(seed, playout, t);
}

/* Static facts (tic_tac_toe.h:100-110, connect_four.h:105-118,
 * breakthrough.cc:383-385 + breakthrough.h:100-115, hex.h:98-118). */
int orc_num_distinct_actions(int game, int n) {
  switch (game) {
    case ORC_TTT: return 9;
    case ORC_C4: return 7;
    case ORC_BT: return 8 * 8 * 6 * 2;
    case ORC_PENTAGO: return 36 * 8; /* pentago.h:40-42 */
    case ORC_OTHELLO: return 65;     /* othello.h:128: 64 cells + pass */
    case ORC_Y: return n * n;        /* y.h:174-178: the cut-off corner is never legal */
    case ORC_HAVANNAH: return (2 * n - 1) * (2 * n - 1);   /* havannah.h:196-200 */
    case ORC_GO: return n * n + 1;   /* go.h:52-54 */
    default: return n * n;
  }
}
int orc_max_game_length(int game, int n) {
  switch (game) {
    case ORC_TTT: return 9;
    case ORC_C4: return 42;
    case ORC_BT: return 2 * (2 * 8 - 3) * 8 + 1; /* breakthrough.h:116-118 */
    case ORC_PENTAGO: return 36; /* pentago.h:109-113 */
    case ORC_OTHELLO: return 64; /* othello.h:142 */
    case ORC_Y: return n * (n + 1) / 2; /* y.h:192-197 */
    case ORC_HAVANNAH: return (2 * n - 1) * (2 * n - 1) - n * (n - 1); /* havannah.h:212-217 */
    case ORC_GO: return n * n * 2;   /* go.h:59 */
    default: return n * n;
  }
}
int orc_obs_size(int game, int n) {
  switch (game) {
    case ORC_TTT: return 27;
    case ORC_C4: return 126;
    case ORC_BT: return 192;
    case ORC_PENTAGO: return 3 * 36; /* pentago.h:106-108 */
    case ORC_OTHELLO: return 3 * 64; /* othello.h:139-141 */
    case ORC_Y: return 3 * n * n;    /* y.h:189-191 */
    case ORC_HAVANNAH: return 3 * (2 * n - 1) * (2 * n - 1); /* havannah.h:209-211 */
    case ORC_GO: return 4 * n * n;   /* go.h:141-143, go.cc:108 */
    default: return 9 * n * n;
  }
}

/* ---------------------------------------------------------------- tic_tac_toe */
/* player 0 = cross, player 1 = nought (tic_tac_toe.cc:55-65). */
static int ttt_mark(int player) { return player == 0 ? X : O; }

/* tic_tac_toe.cc:108-118: the eight lines, tested for one player. */
static int ttt_has_line(const int* b, int player) {
  static const int L[8][3] = {{0, 1, 2}, {3, 4, 5}, {6, 7, 8}, {0, 3, 6},
                              {1, 4, 7}, {2, 5, 8}, {0, 4, 8}, {2, 4, 6}};
  int c = ttt_mark(player);
  for (int i = 0; i < 8; ++i)
    if (b[L[i][0]] == c && b[L[i][1]] == c && b[L[i][2]] == c) return 1;
  return 0;
}

/* tic_tac_toe.cc:80-88 */
static int ttt_apply(orc_state* s, int a) {
  if (a < 0 || a >= 9 || s->cells[a] != E) return -1;
  s->cells[a] = ttt_mark(s->cur_player);
  if (ttt_has_line(s->cells, s->cur_player)) s->outcome = s->cur_player;
  s->cur_player = 1 - s->cur_player;
  s->num_moves += 1;
  return 0;
}

/* ---------------------------------------------------------------- connect_four */
#define C4_ROWS 6
#define C4_COLS 7
static int c4_at(const orc_state* s, int r, int c) { return s->cells[r * C4_COLS + c]; }

/* connect_four.cc:130-142 */
static int c4_line_dir(const orc_state* s, int player, int row, int col, int dr, int dc) {
  if (row + 3 * dr >= C4_ROWS || col + 3 * dc >= C4_COLS || row + 3 * dr < 0 || col + 3 * dc < 0)
    return 0;
  int c = ttt_mark(player);
  for (int i = 0; i < 4; ++i) {
    if (c4_at(s, row, col) != c) return 0;
    row += dr;
    col += dc;
  }
  return 1;
}
/* connect_four.cc:123-128, 144-152 */
static int c4_has_line(const orc_state* s, int player) {
  int c = ttt_mark(player);
  for (int col = 0; col < C4_COLS; ++col)
    for (int row = 0; row < C4_ROWS; ++row)
      if (c4_at(s, row, col) == c &&
          (c4_line_dir(s, player, row, col, 0, 1) || c4_line_dir(s, player, row, col, -1, -1) ||
           c4_line_dir(s, player, row, col, -1, 0) || c4_line_dir(s, player, row, col, -1, 1)))
        return 1;
  return 0;
}
/* connect_four.cc:154-159 */
static int c4_is_full(const orc_state* s) {
  for (int col = 0; col < C4_COLS; ++col)
    if (c4_at(s, C4_ROWS - 1, col) == E) return 0;
  return 1;
}
/* connect_four.cc:93-106: gravity drop, win tested before draw. */
static int c4_apply(orc_state* s, int a) {
  if (a < 0 || a >= C4_COLS || c4_at(s, C4_ROWS - 1, a) != E) return -1;
  int row = 0;
  while (c4_at(s, row, a) != E) ++row;
  s->cells[row * C4_COLS + a] = ttt_mark(s->cur_player);
  if (c4_has_line(s, s->cur_player)) s->outcome = s->cur_player;
  else if (c4_is_full(s)) s->outcome = 2;
  s->cur_player = 1 - s->cur_player;
  s->num_moves += 1;
  return 0;
}

/* ---------------------------------------------------------------- breakthrough (8x8) */
#define BT_N 8
/* breakthrough.cc:35-40: direction offsets for black (0..2) then white (3..5). */
static const int kBtDRow[6] = {1, 1, 1, -1, -1, -1};
static const int kBtDCol[6] = {-1, 0, 1, -1, 0, 1};
static int bt_in(int r, int c) { return r >= 0 && r < BT_N && c >= 0 && c < BT_N; }
static int bt_mine(int player) { return player == 0 ? BT_B : BT_W; }
static int bt_theirs(int player) { return player == 0 ? BT_W : BT_B; }

/* breakthrough.cc:217-256 with the action id of spiel_utils.cc:45-58 over bases {8,8,6,2}:
 * id = ((r*8 + c)*6 + dir)*2 + capture. */
static int bt_legal(const orc_state* s, int* out) {
  int n = 0;
  int p = s->cur_player;
  for (int r = 0; r < BT_N; ++r)
    for (int c = 0; c < BT_N; ++c) {
      if (s->cells[r * BT_N + c] != bt_mine(p)) continue;
      for (int o = 0; o < 3; ++o) {
        int dir = p * 3 + o;
        int rp = r + kBtDRow[dir], cp = c + kBtDCol[dir];
        if (!bt_in(rp, cp)) continue;
        int target = s->cells[rp * BT_N + cp];
        if (target == BT_E) out[n++] = ((r * BT_N + c) * 6 + dir) * 2 + 0;
        else if ((o == 0 || o == 2) && target == bt_theirs(p))
          out[n++] = ((r * BT_N + c) * 6 + dir) * 2 + 1;
      }
    }
  return n;
}

/* breakthrough.cc:152-192.  The reference moves whatever the id decodes to and only checks
 * bounds and colours; the oracle additionally rejects ids that are not in the legal list. */
static int bt_apply(orc_state* s, int a) {
  int legal[128];
  int n = bt_legal(s, legal), ok = 0;
  for (int i = 0; i < n; ++i) ok |= (legal[i] == a);
  if (!ok) return -1;
  int cap = a % 2, dir = (a / 2) % 6, c1 = (a / 12) % BT_N, r1 = a / (12 * BT_N);
  int r2 = r1 + kBtDRow[dir], c2 = c1 + kBtDCol[dir];
  (void)cap;
  if (s->cells[r2 * BT_N + c2] == BT_W) s->pieces[1]--;
  else if (s->cells[r2 * BT_N + c2] == BT_B) s->pieces[0]--;
  s->cells[r2 * BT_N + c2] = s->cells[r1 * BT_N + c1];
  s->cells[r1 * BT_N + c1] = BT_E;
  if (s->cur_player == 0 && r2 == BT_N - 1) s->outcome = 0;
  else if (s->cur_player == 1 && r2 == 0) s->outcome = 1;
  s->cur_player = 1 - s->cur_player;
  s->num_moves += 1;
  return 0;
}

/* ---------------------------------------------------------------- hex */
/* hex.cc:206-222: six neighbours with bounds and column-wrap guards. */
static int hex_adjacent(int cell, int n, int* out) {
  int cand[6] = {cell - n, cell - n + 1, cell - 1, cell + 1, cell + n - 1, cell + n};
  int k = 0;
  for (int i = 0; i < 6; ++i) {
    int v = cand[i];
    if (v < 0 || v >= n * n) continue;
    if (v % n == 0 && cell % n == n - 1) continue;
    if (v % n == n - 1 && cell % n == 0) continue;
    out[k++] = v;
  }
  return k;
}

/* hex.cc:55-121: label of a new stone from board edges and neighbours' labels. */
static int hex_new_label(const orc_state* s, int player, int move) {
  int n = s->board_size, nb[6];
  int k = hex_adjacent(move, n, nb);
  if (player == 0) {
    int north = 0, south = 0;
    if (move < n) north = 1;
    else if (move >= n * (n - 1)) south = 1;
    for (int i = 0; i < k; ++i) {
      if (s->cells[nb[i]] == HX_BNORTH) north = 1;
      else if (s->cells[nb[i]] == HX_BSOUTH) south = 1;
    }
    if (north && south) return HX_BWIN;
    if (north) return HX_BNORTH;
    if (south) return HX_BSOUTH;
    return HX_BLACK;
  } else {
    int west = 0, east = 0;
    if (move % n == 0) west = 1;
    else if (move % n == n - 1) east = 1;
    for (int i = 0; i < k; ++i) {
      if (s->cells[nb[i]] == HX_WWEST) west = 1;
      else if (s->cells[nb[i]] == HX_WEAST) east = 1;
    }
    if (west && east) return HX_WWIN;
    if (west) return HX_WWEST;
    if (east) return HX_WEAST;
    return HX_WHITE;
  }
}

/* hex.cc:149-184: place, then flood the edge label through plain stones of the mover. */
static int hex_apply(orc_state* s, int a) {
  int n = s->board_size;
  if (a < 0 || a >= n * n || s->cells[a] != HX_EMPTY) return -1;
  int label = hex_new_label(s, s->cur_player, a);
  s->cells[a] = label;
  if (label == HX_BWIN) s->outcome = 0;
  else if (label == HX_WWIN) s->outcome = 1;
  else if (label != HX_BLACK && label != HX_WHITE) {
    int plain = s->cur_player == 0 ? HX_BLACK : HX_WHITE;
    int stack[ORC_MAX_CELLS], top = 0, nb[6];
    stack[top++] = a;
    while (top > 0) {
      int cell = stack[--top];
      int k = hex_adjacent(cell, n, nb);
      for (int i = 0; i < k; ++i)
        if (s->cells[nb[i]] == plain) {
          s->cells[nb[i]] = label;
          stack[top++] = nb[i];
        }
    }
  }
  s->cur_player = 1 - s->cur_player;
  s->num_moves += 1;
  return 0;
}

/* ---------------------------------------------------------------- pentago (6x6) */
#define PG_N 6
/* pentago.cc:52-75: action = ((y*6 + x) * 8) + r; quadrant = r >> 1 numbered clockwise from
 * the top-left (0 TL, 1 TR, 2 BR, 3 BL: the bit order table pentago.cc:78-85), direction =
 * r & 1 (0 counter-clockwise, 1 clockwise: pentago.cc:132-141 move ring bit i to i-2 / i+2 and
 * every quadrant's ring is numbered clockwise). */
static void pg_rotate(int* cells, int quadrant, int clockwise) {
  int ox = (quadrant == 1 || quadrant == 2) ? 3 : 0, oy = quadrant >= 2 ? 3 : 0;
  int old[3][3];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) old[r][c] = cells[(oy + r) * PG_N + ox + c];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      /* clockwise (y grows downwards): the cell at row r, column c moves to row c, column 2-r */
      int nr = clockwise ? c : 2 - c, nc = clockwise ? 2 - r : r;
      cells[(oy + nr) * PG_N + ox + nc] = old[r][c];
    }
}
/* pentago.cc:104-125 win_mask: every horizontal, vertical and diagonal run of five. */
static int pg_has_five(const int* cells, int code) {
  static const int D[4][2] = {{1, 0}, {0, 1}, {1, 1}, {1, -1}};
  for (int y = 0; y < PG_N; ++y)
    for (int x = 0; x < PG_N; ++x)
      for (int d = 0; d < 4; ++d) {
        int ex = x + 4 * D[d][0], ey = y + 4 * D[d][1], i;
        if (ex < 0 || ex >= PG_N || ey < 0 || ey >= PG_N) continue;
        for (i = 0; i < 5; ++i)
          if (cells[(y + i * D[d][1]) * PG_N + x + i * D[d][0]] != code) break;
        if (i == 5) return 1;
      }
  return 0;
}
/* pentago.cc:147-160 */
static int pg_legal(const orc_state* s, int* out) {
  int n = 0;
  for (int y = 0; y < PG_N; ++y)
    for (int x = 0; x < PG_N; ++x)
      if (s->cells[y * PG_N + x] == 0)
        for (int r = 0; r < 8; ++r) out[n++] = (y * PG_N + x) * 8 + r;
  return n;
}
/* pentago.cc:274-312: place, rotate both colours, then both players' lines: a rotation can
 * complete the opponent's line (they win) or both (draw); a full board is a draw. */
static int pg_apply(orc_state* s, int a) {
  if (a < 0 || a >= 288 || s->cells[a / 8] != 0) return -1;
  s->cells[a / 8] = s->cur_player + 1;
  pg_rotate(s->cells, (a % 8) >> 1, (a % 8) & 1);
  s->num_moves += 1;
  int w0 = pg_has_five(s->cells, 1), w1 = pg_has_five(s->cells, 2);
  if (w0 && w1) s->outcome = 2;
  else if (w0) s->outcome = 0;
  else if (w1) s->outcome = 1;
  else if (s->num_moves == 36) s->outcome = 2;
  s->cur_player = 1 - s->cur_player;
  return 0;
}

/* ---------------------------------------------------------------- othello (8x8) */
#define OT_N 8
#define OT_PASS 64
/* othello.h:46-57 + othello.cc:97-120: the eight directions as (drow, dcol). */
static const int kOtDir[8][2] = {{-1, 0}, {1, 0}, {0, -1}, {0, 1}, {-1, -1}, {-1, 1}, {1, -1}, {1, 1}};
static int ot_on(int r, int c) { return r >= 0 && r < OT_N && c >= 0 && c < OT_N; }
/* othello.cc:126-146: opponent discs from `move` in direction d that end on a disc of `player`. */
static int ot_count_steps(const orc_state* s, int player, int move, int d) {
  int r = move / OT_N + kOtDir[d][0], c = move % OT_N + kOtDir[d][1], count = 0;
  int mine = player + 1;
  while (ot_on(r, c)) {
    int v = s->cells[r * OT_N + c];
    if (v == mine) return count;
    if (v == 0) return 0;
    ++count;
    r += kOtDir[d][0];
    c += kOtDir[d][1];
  }
  return 0;
}
/* othello.cc:148-158, 196-198 */
static int ot_valid(const orc_state* s, int player, int move) {
  if (s->cells[move] != 0) return 0;
  for (int d = 0; d < 8; ++d)
    if (ot_count_steps(s, player, move, d) != 0) return 1;
  return 0;
}
static int ot_regular(const orc_state* s, int player, int* out) {   /* othello.cc:231-239 */
  int n = 0;
  for (int cell = 0; cell < OT_N * OT_N; ++cell)
    if (ot_valid(s, player, cell)) { if (out) out[n] = cell; ++n; }
  return n;
}
/* othello.cc:241-246: a player without a regular move must pass. */
static int ot_legal(const orc_state* s, int* out) {
  int n = ot_regular(s, s->cur_player, out);
  if (n == 0) out[n++] = OT_PASS;
  return n;
}
/* othello.cc:200-229: pass flips the turn; a placement captures in every direction, then the
 * game ends when neither side has a regular move (disc majority wins, equal = tie). */
static int ot_apply(orc_state* s, int a) {
  if (a == OT_PASS) {
    if (ot_regular(s, s->cur_player, 0) != 0) return -1;   /* the oracle also rejects illegal passes */
    s->cur_player = 1 - s->cur_player;
    s->num_moves += 1;
    return 0;
  }
  if (a < 0 || a >= OT_N * OT_N || !ot_valid(s, s->cur_player, a)) return -1;
  int mine = s->cur_player + 1;
  int steps[8];
  for (int d = 0; d < 8; ++d) steps[d] = ot_count_steps(s, s->cur_player, a, d);
  s->cells[a] = mine;
  for (int d = 0; d < 8; ++d) {
    int r = a / OT_N, c = a % OT_N;
    for (int k = 0; k < steps[d]; ++k) {
      r += kOtDir[d][0];
      c += kOtDir[d][1];
      s->cells[r * OT_N + c] = mine;
    }
  }
  s->num_moves += 1;
  if (ot_regular(s, 0, 0) == 0 && ot_regular(s, 1, 0) == 0) {
    int c0 = 0, c1 = 0;
    for (int i = 0; i < OT_N * OT_N; ++i) { c0 += s->cells[i] == 1; c1 += s->cells[i] == 2; }
    s->outcome = c0 > c1 ? 0 : (c0 < c1 ? 1 : 2);
  }
  s->cur_player = 1 - s->cur_player;
  return 0;
}

/* ---------------------------------------------------------------- y (N x N grid, the triangle x + y < N) */
enum { Y_P0 = 0, Y_P1 = 1, Y_NONE = 2, Y_INVALID = 3 };   /* y.h:43-48 */
/* y.h:56-63 */
static int y_xy(int x, int y, int n) {
  return (x >= 0 && y >= 0 && x < n && y < n && x + y < n) ? x + y * n : -1;
}
/* y.cc:106-111: bit 0 the x = 0 edge, bit 1 the y = 0 edge, bit 2 the x + y = N - 1 edge */
static int y_edge(int x, int y, int n) {
  if (y_xy(x, y, n) < 0) return 0;
  return (x == 0 ? 1 : 0) | (y == 0 ? 2 : 0) | (x + y == n - 1 ? 4 : 0);
}
/* y.cc:67-71: neighbour offsets, clockwise */
static const int kYNb[6][2] = {{0, -1}, {1, -1}, {1, 0}, {0, 1}, {-1, 1}, {-1, 0}};
/* y.cc:120-132 */
static void y_init(orc_state* s) {
  int n = s->board_size;
  for (int i = 0; i < n * n; ++i) {
    int x = i % n, y = i / n;
    s->cells[i] = y_xy(x, y, n) >= 0 ? Y_NONE : Y_INVALID;
    s->parent[i] = (int16_t)i;
    s->size[i] = 1;
    s->edge[i] = (uint8_t)y_edge(x, y, n);
  }
}
/* y.cc:293-304: follow the parents to the leader, compress this cell's path only */
static int y_find(orc_state* s, int cell) {
  int parent = s->parent[cell];
  if (parent != cell) {
    do {
      parent = s->parent[parent];
    } while (parent != s->parent[parent]);
    s->parent[cell] = (int16_t)parent;
  }
  return parent;
}
/* y.cc:306-325: the smaller group joins the larger one (ties: b joins a) */
static void y_join(orc_state* s, int a, int b) {
  int la = y_find(s, a), lb = y_find(s, b);
  if (la == lb) return;
  if (s->size[la] < s->size[lb]) { int t = la; la = lb; lb = t; }
  s->parent[lb] = (int16_t)la;
  s->size[la] = (int16_t)(s->size[la] + s->size[lb]);
  s->edge[la] |= s->edge[lb];
}
/* y.cc:269-291 */
static int y_apply(orc_state* s, int a) {
  int n = s->board_size;
  if (a < 0 || a >= n * n || s->cells[a] != Y_NONE) return -1;
  int x = a % n, y = a / n;
  s->cells[a] = s->cur_player;
  s->num_moves += 1;
  for (int d = 0; d < 6; ++d) {
    int m = y_xy(x + kYNb[d][0], y + kYNb[d][1], n);
    if (m >= 0 && s->cells[m] == s->cur_player) y_join(s, a, m);
  }
  if (s->edge[y_find(s, a)] == 7) s->outcome = s->cur_player;
  s->cur_player = 1 - s->cur_player;
  return 0;
}

/* ---------------------------------------------------------------- havannah (radius S, grid (2S-1)^2) */
enum { HV_P0 = 0, HV_P1 = 1, HV_NONE = 2, HV_INVALID = 4 };   /* havannah.h:44-50 */
static int hv_diam(const orc_state* s) { return 2 * s->board_size - 1; }
/* havannah.h:60-69 */
static int hv_xy(int x, int y, int size) {
  int d = 2 * size - 1;
  return (x >= 0 && y >= 0 && x < d && y < d && (y - x < size) && (x - y < size)) ? x + y * d : -1;
}
/* havannah.cc:112-126 */
static int hv_corner(int x, int y, int size) {
  int m = size - 1, e = m * 2;
  if (hv_xy(x, y, size) < 0) return 0;
  if (x == 0 && y == 0) return 1 << 0;
  if (x == m && y == 0) return 1 << 1;
  if (x == e && y == m) return 1 << 2;
  if (x == e && y == e) return 1 << 3;
  if (x == m && y == e) return 1 << 4;
  if (x == 0 && y == m) return 1 << 5;
  return 0;
}
/* havannah.cc:128-143 */
static int hv_edge(int x, int y, int size) {
  int m = size - 1, e = m * 2;
  if (hv_xy(x, y, size) < 0) return 0;
  if (y == 0 && x != 0 && x != m) return 1 << 0;
  if (x - y == m && x != m && x != e) return 1 << 1;
  if (x == e && y != m && y != e) return 1 << 2;
  if (y == e && x != e && x != m) return 1 << 3;
  if (y - x == m && x != m && x != 0) return 1 << 4;
  if (x == 0 && y != m && y != 0) return 1 << 5;
  return 0;
}
/* havannah.cc:61-75: neighbour offsets, clockwise from the upper left */
static const int kHvNb[6][2] = {{-1, -1}, {0, -1}, {1, 0}, {1, 1}, {0, 1}, {-1, 0}};
static int hv_neighbour(const orc_state* s, int cell, int dir) {
  int d = hv_diam(s);
  return hv_xy(cell % d + kHvNb[dir][0], cell / d + kHvNb[dir][1], s->board_size);
}
static int hv_bits(int v) { int n = 0; while (v) { n += v & 1; v >>= 1; } return n; }
/* havannah.cc:159-179 */
static void hv_init(orc_state* s) {
  int d = hv_diam(s);
  for (int i = 0; i < d * d; ++i) {
    int x = i % d, y = i / d;
    s->cells[i] = hv_xy(x, y, s->board_size) >= 0 ? HV_NONE : HV_INVALID;
    s->parent[i] = (int16_t)i;
    s->size[i] = 1;
    s->corner[i] = (uint8_t)hv_corner(x, y, s->board_size);
    s->edge[i] = (uint8_t)hv_edge(x, y, s->board_size);
    s->mark[i] = 0;
  }
}
/* havannah.cc:351-362 */
static int hv_find(orc_state* s, int cell) {
  int parent = s->parent[cell];
  if (parent != cell) {
    do {
      parent = s->parent[parent];
    } while (parent != s->parent[parent]);
    s->parent[cell] = (int16_t)parent;
  }
  return parent;
}
/* havannah.cc:364-383: returns 1 when the cells were in one group already */
static int hv_join(orc_state* s, int a, int b) {
  int la = hv_find(s, a), lb = hv_find(s, b);
  if (la == lb) return 1;
  if (s->size[la] < s->size[lb]) { int t = la; la = lb; lb = t; }
  s->parent[lb] = (int16_t)la;
  s->size[la] = (int16_t)(s->size[la] + s->size[lb]);
  s->corner[la] |= s->corner[lb];
  s->edge[la] |= s->edge[lb];
  return 0;
}
/* havannah.cc:385-399 */
static int hv_ring_dfs(orc_state* s, int cell, int left, int right) {
  if (cell < 0) return 0;
  if (s->cells[cell] != s->cur_player) return 0;
  if (s->mark[cell]) return 1;
  s->mark[cell] = 1;
  int success = 0;
  for (int i = left; !success && i <= right; ++i) {
    int dir = (i + 6) % 6;
    success = hv_ring_dfs(s, hv_neighbour(s, cell, dir), dir - 1, dir + 1);
  }
  s->mark[cell] = 0;
  return success;
}
/* havannah.cc:314-349 */
static int hv_apply(orc_state* s, int a) {
  int d = hv_diam(s);
  if (a < 0 || a >= d * d || s->cells[a] != HV_NONE) return -1;
  s->cells[a] = s->cur_player;
  s->num_moves += 1;
  int already = 0, skip = 0;
  for (int dir = 0; dir < 6; ++dir) {
    if (skip) { skip = 0; continue; }
    int m = hv_neighbour(s, a, dir);
    if (m >= 0 && s->cells[m] == s->cur_player) {
      already |= hv_join(s, a, m);
      skip = 1;
    }
  }
  int g = hv_find(s, a);
  if (hv_bits(s->edge[g]) >= 3 || hv_bits(s->corner[g]) >= 2 || (already && hv_ring_dfs(s, a, 0, 3)))
    s->outcome = s->cur_player;
  else if (s->num_moves == orc_max_game_length(ORC_HAVANNAH, s->board_size))
    s->outcome = 2;
  s->cur_player = 1 - s->cur_player;
  return 0;
}

/* ---------------------------------------------------------------- go (N x N, N <= 9) */
enum { GO_BLACK = 0, GO_WHITE = 1, GO_EMPTY = 2 };   /* go_board.h: GoColor */
/* go_board.cc:28-34: the neighbours in the reference's order (row + 1, col + 1, col - 1, row - 1) */
static int go_nb(const orc_state* s, int cell, int d) {
  int n = s->board_size, r = cell / n, c = cell % n;
  int r2 = r + (d == 0 ? 1 : (d == 3 ? -1 : 0)), c2 = c + (d == 1 ? 1 : (d == 2 ? -1 : 0));
  return (r2 >= 0 && r2 < n && c2 >= 0 && c2 < n) ? r2 * n + c2 : -1;
}
/* Walks the chain of `cell`, marking its stones in seen[] (1) and its distinct liberties (2);
 * returns the number of distinct liberties.  (The reference tracks pseudo-liberty sums per
 * chain, go_board.h:173-192; in_atari() there means exactly one distinct liberty.) */
static int go_chain(const orc_state* s, int cell, uint8_t* seen) {
  int colour = s->cells[cell], libs = 0;
  seen[cell] = 1;
  for (int d = 0; d < 4; ++d) {
    int q = go_nb(s, cell, d);
    if (q < 0 || seen[q]) continue;
    if (s->cells[q] == GO_EMPTY) { seen[q] = 2; ++libs; }
    else if (s->cells[q] == colour) libs += go_chain(s, q, seen);
  }
  return libs;
}
static int go_liberties(const orc_state* s, int cell, uint8_t* seen) {
  memset(seen, 0, 81);
  return go_chain(s, cell, seen);
}
/* go_board.cc:416-441 */
static int go_is_legal(const orc_state* s, int p, int c) {
  uint8_t seen[81];
  if (s->cells[p] != GO_EMPTY || p == s->ko_point) return 0;
  for (int d = 0; d < 4; ++d) {           /* chain(p).num_pseudo_liberties > 0: an empty neighbour */
    int q = go_nb(s, p, d);
    if (q >= 0 && s->cells[q] == GO_EMPTY) return 1;
  }
  for (int d = 0; d < 4; ++d) {           /* connects to a friendly chain that is not in atari */
    int q = go_nb(s, p, d);
    if (q >= 0 && s->cells[q] == c && go_liberties(s, q, seen) != 1) return 1;
  }
  for (int d = 0; d < 4; ++d) {           /* kills an enemy chain that is in atari */
    int q = go_nb(s, p, d);
    if (q >= 0 && s->cells[q] == 1 - c && go_liberties(s, q, seen) == 1) return 1;
  }
  return 0;
}
static int go_legal(const orc_state* s, int* out) {      /* go.cc:127-137 */
  int n = 0, cells = s->board_size * s->board_size;
  for (int p = 0; p < cells; ++p) if (go_is_legal(s, p, s->cur_player)) out[n++] = p;
  out[n++] = cells;                                      /* the pass, always */
  return n;
}
/* go_board.cc:525-553 */
static int go_surrounded(const orc_state* s, int p, uint8_t* marked, int* rb, int* rw) {
  if (marked[p] || (*rb && *rw)) return 0;
  marked[p] = 1;
  int n = 1;
  for (int d = 0; d < 4; ++d) {
    int q = go_nb(s, p, d);
    if (q < 0) continue;
    if (s->cells[q] == GO_BLACK) *rb = 1;
    else if (s->cells[q] == GO_WHITE) *rw = 1;
    else n += go_surrounded(s, q, marked, rb, rw);
  }
  return n;
}
/* go_board.cc:555-598, times two (komi2 = 2 * komi) */
static int go_score2(const orc_state* s) {
  uint8_t marked[81];
  int delta = 0, cells = s->board_size * s->board_size;
  memset(marked, 0, sizeof(marked));
  for (int p = 0; p < cells; ++p) {
    if (s->cells[p] == GO_BLACK) ++delta;
    else if (s->cells[p] == GO_WHITE) --delta;
    else if (!marked[p]) {
      int rb = 0, rw = 0;
      int n = go_surrounded(s, p, marked, &rb, &rw);
      if (rb && !rw) delta += n;
      else if (!rb && rw) delta -= n;
    }
  }
  return 2 * delta - s->komi2;
}
static int go_terminal(const orc_state* s) {             /* go.cc:166-172 */
  int cells = s->board_size * s->board_size;
  if (s->num_moves < 2) return 0;
  return s->num_moves >= 2 * cells || s->superko || (s->last_action == cells && s->prev_action == cells);
}
static void go_remove_chain(orc_state* s, int cell, int colour, int* captured, int* where) {
  s->cells[cell] = GO_EMPTY;
  *captured += 1;
  *where = cell;
  for (int d = 0; d < 4; ++d) {
    int q = go_nb(s, cell, d);
    if (q >= 0 && s->cells[q] == colour) go_remove_chain(s, q, colour, captured, where);
  }
}
/* go.cc:201-211 + go_board.cc:257-289 */
static int go_apply(orc_state* s, int a) {
  int cells = s->board_size * s->board_size, c = s->cur_player;
  uint8_t seen[81];
  if (a < 0 || a > cells) return -1;
  if (a == cells) {
    s->ko_point = -1;
  } else {
    if (!go_is_legal(s, a, c)) return -1;
    int eye = 1;                                         /* played in an enemy eye? */
    for (int d = 0; d < 4; ++d) {
      int q = go_nb(s, a, d);
      if (q >= 0 && (s->cells[q] == c || s->cells[q] == GO_EMPTY)) eye = 0;
    }
    s->cells[a] = c;
    int captured = 0, where = -1;
    for (int d = 0; d < 4; ++d) {                        /* CaptureDeadChains */
      int q = go_nb(s, a, d);
      if (q >= 0 && s->cells[q] == 1 - c && go_liberties(s, q, seen) == 0)
        go_remove_chain(s, q, 1 - c, &captured, &where);
    }
    s->ko_point = (eye && captured == 1) ? where : -1;
  }
  s->prev_action = s->last_action;
  s->last_action = a;
  s->cur_player = 1 - c;
  /* repetitions_: every position so far; a non-pass move that recreates one is a superko */
  int k = s->num_moves + 1;
  for (int p = 0; p < cells; ++p) s->past[k][p] = (uint8_t)s->cells[p];
  if (a != cells)
    for (int j = 0; j < k && !s->superko; ++j)
      if (memcmp(s->past[j], s->past[k], (size_t)cells) == 0) s->superko = 1;
  s->num_moves = k;
  if (go_terminal(s)) {                                  /* go.cc:174-199 */
    if (s->superko) s->outcome = 2;
    else { int sc = go_score2(s); s->outcome = sc > 0 ? 0 : (sc < 0 ? 1 : 2); }
  }
  return 0;
}

/* ---------------------------------------------------------------- generic API */
void orc_init(orc_state* s, int game, int board_size) {
  memset(s, 0, sizeof(*s));
  s->game = game;
  s->outcome = -1;
  s->board_size = game == ORC_HEX ? (board_size > 0 ? board_size : 11) : 0;
  if (game == ORC_Y) {
    /* y.h:38: 19 by default; the reference's playthrough uses 9.  A hex-style default of 11
     * passed by a caller that does not care about the size means "the default". */
    s->board_size = (board_size > 0 && board_size != 11) ? board_size : 19;
    y_init(s);
  }
  if (game == ORC_GO) {
    /* go.h:34-36: komi 7.5 by default; the device game is 9 x 9; the reference's playthrough is
     * go(board_size=7,komi=4.5) */
    s->board_size = (board_size == 7) ? 7 : 9;
    s->komi2 = s->board_size == 7 ? 9 : 15;
    s->ko_point = -1; s->last_action = -1; s->prev_action = -1;
    for (int p = 0; p < s->board_size * s->board_size; ++p) { s->cells[p] = GO_EMPTY; s->past[0][p] = GO_EMPTY; }
  }
  if (game == ORC_HAVANNAH) {
    /* havannah.h:39: radius 8 by default; the reference's playthrough uses 4 (a hex-style
     * default of 11 passed by a caller that does not care means "the default") */
    s->board_size = (board_size > 0 && board_size != 11) ? board_size : 8;
    hv_init(s);
  }
  if (game == ORC_OTHELLO) {
    /* othello.cc:263-269: white on d4/e5, black on e4/d5; black (player 0) moves first */
    s->cells[3 * OT_N + 3] = 2; s->cells[3 * OT_N + 4] = 1;
    s->cells[4 * OT_N + 3] = 1; s->cells[4 * OT_N + 4] = 2;
  }
  if (game == ORC_BT) {
    /* breakthrough.cc:119-142: black on rows 0-1, white on rows 6-7, black (player 0) first. */
    for (int r = 0; r < BT_N; ++r)
      for (int c = 0; c < BT_N; ++c)
        s->cells[r * BT_N + c] = (r <= 1) ? BT_B : (r >= BT_N - 2 ? BT_W : BT_E);
    s->pieces[0] = s->pieces[1] = 2 * BT_N;
  }
}

int orc_is_terminal(const orc_state* s) {
  switch (s->game) {
    case ORC_TTT: return s->outcome != -1 || s->num_moves == 9;     /* tic_tac_toe.cc:139-141 */
    case ORC_C4: return s->outcome != -1;                           /* connect_four.cc:176-178 */
    case ORC_BT: return s->outcome >= 0 || s->pieces[0] == 0 || s->pieces[1] == 0; /* :306-308 */
    case ORC_PENTAGO: return s->outcome != -1;                      /* pentago.h:62 */
    case ORC_OTHELLO: return s->outcome != -1;                      /* othello.cc:284-286 */
    case ORC_Y: return s->outcome != -1;                            /* y.h:129 */
    case ORC_HAVANNAH: return s->outcome != -1;                     /* havannah.h:146 */
    case ORC_GO: return go_terminal(s);
    default: return s->outcome != -1;                               /* hex.cc:248 */
  }
}

int orc_current_player(const orc_state* s) {
  return orc_is_terminal(s) ? ORC_TERMINAL_PLAYER : s->cur_player;
}

int orc_legal_actions(const orc_state* s, int* out) {
  if (orc_is_terminal(s)) return 0;
  int n = 0;
  switch (s->game) {
    case ORC_TTT: /* tic_tac_toe.cc:90-100 */
      for (int c = 0; c < 9; ++c) if (s->cells[c] == E) out[n++] = c;
      return n;
    case ORC_C4: /* connect_four.cc:108-116 */
      for (int c = 0; c < C4_COLS; ++c) if (c4_at(s, C4_ROWS - 1, c) == E) out[n++] = c;
      return n;
    case ORC_BT:
      return bt_legal(s, out);
    case ORC_PENTAGO:
      return pg_legal(s, out);
    case ORC_OTHELLO:
      return ot_legal(s, out);
    case ORC_Y: /* y.cc:138-149 */
      for (int c = 0; c < s->board_size * s->board_size; ++c)
        if (s->cells[c] == Y_NONE) out[n++] = c;
      return n;
    case ORC_GO:
      return go_legal(s, out);
    case ORC_HAVANNAH: /* havannah.cc:182-194 */
      for (int c = 0; c < hv_diam(s) * hv_diam(s); ++c)
        if (s->cells[c] == HV_NONE) out[n++] = c;
      return n;
    default: /* hex.cc:186-196 */
      for (int c = 0; c < s->board_size * s->board_size; ++c)
        if (s->cells[c] == HX_EMPTY) out[n++] = c;
      return n;
  }
}

int orc_apply(orc_state* s, int a) {
  if (orc_is_terminal(s)) return -1;
  switch (s->game) {
    case ORC_TTT: return ttt_apply(s, a);
    case ORC_C4: return c4_apply(s, a);
    case ORC_BT: return bt_apply(s, a);
    case ORC_PENTAGO: return pg_apply(s, a);
    case ORC_OTHELLO: return ot_apply(s, a);
    case ORC_Y: return y_apply(s, a);
    case ORC_HAVANNAH: return hv_apply(s, a);
    case ORC_GO: return go_apply(s, a);
    default: return hex_apply(s, a);
  }
}

void orc_returns(const orc_state* s, int ret[2]) {
  int w = -1;
  if (s->game == ORC_BT) {
    /* breakthrough.cc:310-318 */
    if (s->outcome == 0 || s->pieces[1] == 0) w = 0;
    else if (s->outcome == 1 || s->pieces[0] == 0) w = 1;
  } else if (s->outcome == 0 || s->outcome == 1) {
    w = s->outcome;
  }
  ret[0] = w == 0 ? 1 : (w == 1 ? -1 : 0);
  ret[1] = -ret[0];
}

void orc_observation(const orc_state* s, int player, uint8_t* out) {
  int ncell, plane;
  switch (s->game) {
    case ORC_TTT: /* tic_tac_toe.cc:165-175: plane = raw cell code, not player-relative */
      memset(out, 0, 27);
      for (int c = 0; c < 9; ++c) out[s->cells[c] * 9 + c] = 1;
      return;
    case ORC_C4: /* connect_four.cc:198-221: nought->(p==0?0:1), cross->(p==1?0:1), empty->2 */
      memset(out, 0, 126);
      for (int c = 0; c < 42; ++c) {
        int v = s->cells[c];
        plane = v == O ? (player == 0 ? 0 : 1) : (v == X ? (player == 1 ? 0 : 1) : 2);
        out[plane * 42 + c] = 1;
      }
      return;
    case ORC_BT: /* breakthrough.cc:284-304,326-340: black 0, white 1, empty 2 */
      memset(out, 0, 192);
      for (int c = 0; c < 64; ++c) {
        int v = s->cells[c];
        plane = v == BT_B ? 0 : (v == BT_W ? 1 : 2);
        out[plane * 64 + c] = 1;
      }
      return;
    case ORC_OTHELLO: /* othello.cc:310-329: plane 0 empty, 1 the observer's discs, 2 the opponent's */
      memset(out, 0, 192);
      for (int c = 0; c < 64; ++c) {
        int v = s->cells[c];
        plane = v == 0 ? 0 : ((v - 1) == player ? 1 : 2);
        out[plane * 64 + c] = 1;
      }
      return;
    case ORC_Y: /* y.cc:239-267: plane 0 the observer's stones, 1 the other's, 2 empty; cells
                 * off the board stay zero in every plane */
      ncell = s->board_size * s->board_size;
      memset(out, 0, 3 * ncell);
      for (int c = 0; c < ncell; ++c) {
        int v = s->cells[c];
        if (v == Y_INVALID) continue;
        plane = v == Y_NONE ? 2 : (v == player ? 0 : 1);
        out[plane * ncell + c] = 1;
      }
      return;
    case ORC_GO: /* go.cc:104-125: black, white, empty, then ones when white is to play */
      ncell = s->board_size * s->board_size;
      memset(out, 0, 4 * ncell);
      for (int c = 0; c < ncell; ++c) out[s->cells[c] * ncell + c] = 1;
      if (s->cur_player == 1) memset(out + 3 * ncell, 1, ncell);
      return;
    case ORC_HAVANNAH: /* havannah.cc:284-312: plane 0 the observer's stones, 1 the other's, 2
                        * empty; the cut-off corners stay zero in every plane */
      ncell = hv_diam(s) * hv_diam(s);
      memset(out, 0, 3 * ncell);
      for (int c = 0; c < ncell; ++c) {
        int v = s->cells[c];
        if (v == HV_INVALID) continue;
        plane = v == HV_NONE ? 2 : (v == player ? 0 : 1);
        out[plane * ncell + c] = 1;
      }
      return;
    case ORC_PENTAGO: /* pentago.cc:246-272: plane 0 the observer's stones, 1 the other's, 2 empty */
      memset(out, 0, 108);
      for (int c = 0; c < 36; ++c) {
        int v = s->cells[c];
        plane = v == 0 ? 2 : ((v - 1) == player ? 0 : 1);
        out[plane * 36 + c] = 1;
      }
      return;
    default: /* hex.cc:266-277: plane = state - kMinValueCellState(-4) */
      ncell = s->board_size * s->board_size;
      memset(out, 0, 9 * ncell);
      for (int c = 0; c < ncell; ++c) out[(s->cells[c] + 4) * ncell + c] = 1;
      return;
  }
}

static int orc_from_history(orc_state* s, int game, int board_size, const int64_t* hist, int n) {
  orc_init(s, game, board_size);
  for (int i = 0; i < n; ++i)
    if (orc_apply(s, (int)hist[i]) != 0) return -1;
  return 0;
}

static int orc_replay_from(orc_state* s, const uint32_t* r_table, uint64_t seed, uint64_t playout,
                           int max_steps, int64_t* actions_out, int* len_out, int* terminal_out,
                           int* ret0_out) {
  int legal[ORC_MAX_LEGAL], t = 0, ret[2];
  while (!orc_is_terminal(s) && t < max_steps) {
    int n = orc_legal_actions(s, legal);
    if (n == 0) return -1;
    uint32_t r = r_table ? r_table[t] : oracle_draw(seed, playout, (uint32_t)t);
    int a = legal[r % (uint32_t)n];
    if (actions_out) actions_out[t] = a;
    orc_apply(s, a);
    ++t;
  }
  orc_returns(s, ret);
  *len_out = t;
  *terminal_out = orc_is_terminal(s);
  *ret0_out = ret[0];
  return 0;
}

int orc_replay(int game, int board_size, const int64_t* hist, int n_hist, const uint32_t* r_table,
               uint64_t seed, uint64_t playout, int max_steps, int64_t* actions_out, int* len_out,
               int* terminal_out, int* ret0_out) {
  orc_state s;
  if (orc_from_history(&s, game, board_size, hist, n_hist) != 0) return -2;
  return orc_replay_from(&s, r_table, seed, playout, max_steps, actions_out, len_out,
                         terminal_out, ret0_out);
}

int orc_replay_batch(int game, int board_size, const int64_t* hist_flat, const int64_t* hist_off,
                     int64_t n_roots, int rollouts, uint64_t seed, uint64_t first_playout,
                     int max_steps, int32_t* len_out, int8_t* ret0_out, uint8_t* terminal_out,
                     int16_t* actions_out, int actions_stride) {
  int64_t acts[256];
  for (int64_t i = 0; i < n_roots; ++i) {
    orc_state root;
    if (orc_from_history(&root, game, board_size, hist_flat + hist_off[i],
                         (int)(hist_off[i + 1] - hist_off[i])) != 0)
      return -2;
    for (int j = 0; j < rollouts; ++j) {
      int64_t p = i * rollouts + j;
      orc_state s = root;
      int len, term, ret0;
      if (orc_replay_from(&s, 0, seed, first_playout + (uint64_t)p, max_steps, acts, &len, &term,
                          &ret0) != 0)
        return -1;
      len_out[p] = len;
      ret0_out[p] = (int8_t)ret0;
      terminal_out[p] = (uint8_t)term;
      if (actions_out)
        for (int t = 0; t < len && t < actions_stride; ++t)
          actions_out[p * actions_stride + t] = (int16_t)acts[t];
    }
  }
  return 0;
}

int orc_state_dump(int game, int board_size, const int64_t* hist, int n_hist, int player,
                   uint8_t* obs_out, int64_t* legal_out, int* n_legal, int* cur_player,
                   int* terminal, int* returns_out) {
  orc_state s;
  int legal[ORC_MAX_LEGAL];
  if (orc_from_history(&s, game, board_size, hist, n_hist) != 0) return -2;
  *n_legal = orc_legal_actions(&s, legal);
  if (legal_out) for (int i = 0; i < *n_legal; ++i) legal_out[i] = legal[i];
  *cur_player = orc_current_player(&s);
  *terminal = orc_is_terminal(&s);
  orc_returns(&s, returns_out);
  if (obs_out) orc_observation(&s, player, obs_out);
  return 0;
}

/* Scalar timing loop for bench.py's cpu_baseline (kind "port"): playout p starts from root
 * p mod n_roots and follows the replay contract with Philox draws. */
int64_t orc_random_playouts(int game, int board_size, const int64_t* hist_flat,
                            const int64_t* hist_off, int64_t n_roots, int64_t n_playouts,
                            uint64_t seed, int64_t* sum_ret0_out) {
  int64_t steps = 0, sum = 0;
  for (int64_t p = 0; p < n_playouts; ++p) {
    int64_t i = p % n_roots;
    orc_state s;
    int len, term, ret0;
    if (orc_from_history(&s, game, board_size, hist_flat + hist_off[i],
                         (int)(hist_off[i + 1] - hist_off[i])) != 0)
      return -2;
    if (orc_replay_from(&s, 0, seed, (uint64_t)p, 1 << 30, 0, &len, &term, &ret0) != 0) return -1;
    steps += len;
    sum += ret0;
  }
  *sum_ret0_out = sum;
  return steps;
}
