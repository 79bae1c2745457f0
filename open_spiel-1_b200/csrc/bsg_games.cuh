// Per-game bitboard state machines of the batched playout engine.
//
// Each game is a struct with the same static interface (used by the templated kernels in
// bsg_kernels.cu and by the host helpers of the C-ABI):
//   State                      registers-only state
//   Tables / fill_tables       small lookup tables (shared memory on the device)
//   init, load, store          initial position; packed SoA HBM layout <-> State
//   terminal, outcome, mover   status; won(s) = the game ended with a win of the LAST mover (in all
//                              four games a win is always the mover's: the kernels record only
//                              this flag and the length, the winner follows from the parity)
//   n = step(s, r, tab, action) one playout step: a = legal[r mod n] in ascending action-id
//                              order, apply it, update terminal/outcome.  Branch-free enough
//                              to be run on finished lanes as a dead step (never traps).
//   legal(s, out)              ascending legal action ids (host side: prior, pack, checks)
//   apply(s, action)           apply one action id (host side; false if illegal)
//   obs_bit(s, player, idx)    element idx of the flat CHW observation (reference order)
//   obs_words(s, player, out)  the whole flat observation as a bit string, kObsWords u32
//                              (bit e of the string = element e); what the batched writer expands
//   mask_words(s, out)         State::LegalActionsMask as a bit string, kMaskWords u32
//
// Semantics follow the reference rules cited per function; the representation does not.
#pragma once
#include "bsg_common.cuh"
#ifndef BSG_TTT_LAZY_DRAW
#define BSG_TTT_LAZY_DRAW 1
#endif
#ifndef BSG_Y_ROLLED
#define BSG_Y_ROLLED 1
#endif
#ifndef BSG_OTHELLO_ROLLED
#define BSG_OTHELLO_ROLLED 0
#endif

namespace bsg {

// =========================================================================================
// tic_tac_toe — reference: open_spiel/games/tic_tac_toe.cc
//   x = player 0 (moves first, tic_tac_toe.cc:55-65), o = player 1.  Two 9-bit masks, kept
//   relative to the mover; move selection, cell and line test come from shared-memory tables.
//   packed: one u32 = x | o << 9 | terminal << 18 | outcome << 19.
// =========================================================================================
struct TicTacToe {
  static constexpr int kId = kGameTicTacToe;
  static constexpr int kNumActions = 9, kMaxLen = 9, kMaxDepth = kMaxLen, kObsSize = 27, kStateBytes = 4;
  static constexpr int kObsWords = 1, kMaskWords = 1;
  static constexpr int kPlanes = 1;   // packed SoA planes and their element sizes
  BSG_HD static int plane_bytes(int p) { (void)p; return 4; }
  static constexpr bool kCanStick = false;  // a non-terminal state always has a legal action
  static constexpr bool kAsync = false;     // a step completes within the step (see HexT)
  static constexpr bool kMoverCanLose = false;  // a move never ends the game in the mover's disfavour (see Pentago)
  static constexpr bool kLazyEnd = false;       // the lean kernel sees a terminal state in the step that reaches it (see Othello)
  // A full board is inert for the lean kernel's bookkeeping (see ConnectFour::kInertWhenFull): the
  // row of the empty mask 0 names "cell 9", a bit outside the board, so a dead step on a full
  // board places nothing on it and — the board being a draw — neither side's marks hold a line;
  // the line table covers that tenth bit.  A playout that has made as many moves as its root had
  // empty cells is over: one comparison per group instead of a terminal test per step.
  static constexpr bool kInertWhenFull = BSG_TTT_LAZY_DRAW != 0;
  struct State {
    u32 cur, opp;   // marks of the player to move / of the other player (9-bit masks)
    u32 left;       // empty cells; parity = player to move (x = player 0 first, 9 cells)
    u32 win;        // the last move completed a line
  };
  // One 24-byte row per 9-bit mask e of empty cells (tic_tac_toe.cc:90-100):
  //   word 0 = floor(2^32 / n), word 1 = -n (n = popc e), bytes 8..16 = the k-th empty cell,
  //   ascending.  line[b] = 1 when the 9-bit mask b contains one of the eight lines
  //   (tic_tac_toe.cc:108-118).  A step is then three shared-memory loads and ~8 ALU
  //   instructions (no popcount, no rank-select, no shifts over the eight lines).
  struct Tables {
    u32 row[512][6];
    unsigned char line[1024];       // indexed by a 9-bit mask, or one with the dead step's tenth bit
  };
  static void fill_tables(Tables& t) {
    for (u32 e = 0; e < 512; ++e) {
      unsigned char* cells = (unsigned char*)&t.row[e][2];
      for (int i = 0; i < 16; ++i) cells[i] = 0;
      u32 n = 0;
      for (u32 c = 0; c < 9; ++c) if ((e >> c) & 1) cells[n++] = (unsigned char)c;
      u32 nn = n ? n : 1;   // e = 0 only on finished lanes running dead steps
      if (!n) cells[0] = 9; // ... which then mark a bit outside the board
      t.row[e][0] = modM_entry(nn);
      t.row[e][1] = 0u - nn;
      t.line[e] = t.line[e | 512u] = has_line(e) ? 1 : 0;
    }
  }

  BSG_HD static void init(State& s) { s.cur = 0; s.opp = 0; s.left = 9; s.win = 0; }
  BSG_HD static u32 moves_left(const State& s) { return s.left; }
  BSG_HD static void load(State& s, const void* base, int64_t n, int64_t i) {
    (void)n;
    u32 w = ((const u32*)base)[i];
    u32 x = w & 0x1FF, o = (w >> 9) & 0x1FF;
    u32 stones = (u32)(popc32(x) + popc32(o));
    u32 mv = stones & 1u;
    s.cur = mv ? o : x;
    s.opp = mv ? x : o;
    s.left = 9u - stones;
    s.win = (((w >> 18) & 1u) && ((w >> 19) & 3u) < 2u) ? 1u : 0u;
  }
  BSG_HD static void store(const State& s, void* base, int64_t n, int64_t i) {
    (void)n;
    u32 mv = (u32)mover(s);
    u32 x = mv ? s.opp : s.cur, o = mv ? s.cur : s.opp;
    ((u32*)base)[i] = x | (o << 9) | ((terminal(s) ? 1u : 0u) << 18) | ((u32)outcome(s) << 19);
  }
  BSG_HD static bool terminal(const State& s) { return (s.win != 0) || (s.left == 0); }
  BSG_HD static bool won(const State& s) { return s.win != 0; }  // the last mover won
  BSG_HD static int mover(const State& s) { return (int)((s.left & 1u) ^ 1u); }  // 9 is odd
  BSG_HD static int outcome(const State& s) {
    return s.win ? (mover(s) ^ 1) : (s.left == 0 ? kOutcomeDraw : kOutcomeNone);
  }

  // tic_tac_toe.cc:108-118, the eight lines, as shifts over one 9-bit mask (table builder and
  // host helpers; the kernels read Tables::line).
  BSG_HD static bool has_line(u32 b) {
    u32 rows = b & (b >> 1) & (b >> 2) & 0x49u;
    u32 cols = b & (b >> 3) & (b >> 6) & 0x7u;
    u32 diag = b & (b >> 4) & (b >> 8) & 0x1u;
    u32 anti = (b >> 2) & (b >> 4) & (b >> 6) & 0x1u;
    return (rows | cols | diag | anti) != 0;
  }
  // tic_tac_toe.cc:80-88 + :139-141 (win for the mover only; full board ends the game).
  BSG_HD static void place(State& s, u32 bit, u32 win) {
    u32 b = s.cur | bit;
    s.win = win;
    s.cur = s.opp;
    s.opp = b;
    s.left -= 1u;
  }
  BSG_HD static u32 step(State& s, u32 r, const Tables& tab, int& action) {
    const u32 e = ~(s.cur | s.opp) & 0x1FFu;  // tic_tac_toe.cc:90-100, ascending cell order
#if defined(__CUDA_ARCH__)
    const u32 base = (u32)__cvta_generic_to_shared(&tab);
    const u32 off = base + e * 24u;
    u32 M, negn, cell, win;
    asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(M), "=r"(negn) : "r"(off));
    u32 rem = umulhi32(r, M) * negn + r;
    u32 k = umin32(rem, rem + negn);                 // r mod n
    asm("ld.shared.u8 %0, [%1+8];" : "=r"(cell) : "r"(off + k));
    const u32 bit = 1u << cell;
    asm("ld.shared.u8 %0, [%1+%2];" : "=r"(win) : "r"(base + (s.cur | bit)), "n"(512 * 24));
#else
    const u32* row = tab.row[e];
    const u32 negn = row[1];
    u32 rem = umulhi32(r, row[0]) * negn + r;
    u32 k = umin32(rem, rem + negn);
    const u32 cell = ((const unsigned char*)&row[2])[k];
    const u32 bit = 1u << cell;
    const u32 win = tab.line[s.cur | bit];
#endif
    action = (int)cell;
    place(s, bit, win);
    return 0u - negn;
  }
  BSG_HD static int legal(const State& s, int* out) {
    if (terminal(s)) return 0;
    int n = 0;
    for (int c = 0; c < 9; ++c) if (!(((s.cur | s.opp) >> c) & 1)) out[n++] = c;
    return n;
  }
  BSG_HD static bool apply(State& s, int a) {
    if (terminal(s) || a < 0 || a >= 9 || (((s.cur | s.opp) >> a) & 1)) return false;
    place(s, 1u << a, has_line(s.cur | (1u << a)) ? 1u : 0u);
    return true;
  }
  BSG_HD static u32 legal_bit(const State& s, int a) {  // State::LegalActionsMask, spiel.h:244-252
    return (!terminal(s) && !(((s.cur | s.opp) >> a) & 1u)) ? 1u : 0u;
  }
  // tic_tac_toe.cc:165-175: plane = raw CellState (0 empty, 1 nought = player 1, 2 cross).
  BSG_HD static u32 obs_bit(const State& s, int player, int idx) {
    (void)player;
    int plane = idx / 9, cell = idx - plane * 9;
    u32 x = mover(s) ? s.opp : s.cur, o = mover(s) ? s.cur : s.opp;
    u32 m = plane == 0 ? ~(x | o) : (plane == 1 ? o : x);
    return (m >> cell) & 1u;
  }
  BSG_HD static void obs_words(const State& s, int player, u32* out) {
    (void)player;
    u32 x = mover(s) ? s.opp : s.cur, o = mover(s) ? s.cur : s.opp;
    out[0] = (~(x | o) & 0x1FFu) | (o << 9) | (x << 18);
  }
  BSG_HD static void mask_words(const State& s, u32* out) {
    out[0] = terminal(s) ? 0u : (~(s.cur | s.opp) & 0x1FFu);
  }
};

// =========================================================================================
// connect_four — reference: open_spiel/games/connect_four.cc
//   6 rows x 7 cols, row 0 at the bottom (connect_four.cc:77-83).  Bit = row * 8 + col: one
//   guard column per row keeps horizontal/diagonal shifts from wrapping, and the top row
//   (bits 40..46) gives the 7-bit "column not full" mask with one shift.  Rows 0-3 live in the
//   low 32-bit word, rows 4-5 in the low half of the high word.
//   packed: ONE u64 per state (8 B — the host->device copy of the roots is the end-to-end
//   bound of this path, so the root format is as small as a position gets):
//     bits 0..54   stones of the player to move | one marker bit per column at (height, col)
//                  (everything below a marker is occupied, so the whole position decodes with
//                  three shift-ors: smear the markers down, shift one row down = occupancy);
//     bit  57      the last move made four in a row (terminal by win);
//     bits 58..63  empty cells (42..0; parity = player to move, x = player 0 first).
// =========================================================================================
struct ConnectFour {
  static constexpr int kId = kGameConnectFour;
  static constexpr int kNumActions = 7, kMaxLen = 42, kMaxDepth = kMaxLen, kObsSize = 126, kStateBytes = 8;
  static constexpr int kObsWords = 4, kMaskWords = 1;
  static constexpr int kPlanes = 1;
  BSG_HD static int plane_bytes(int p) { (void)p; return 8; }
  static constexpr bool kCanStick = false;
  static constexpr bool kAsync = false;
  static constexpr bool kMoverCanLose = false;
  static constexpr bool kLazyEnd = false;       // the lean kernel sees a terminal state in the step that reaches it (see Othello)
  struct State {
    u64 cur, opp;  // stones of the player to move / of the other player
    u32 left;      // empty cells; 42 - left stones on the board, parity = player to move
    u32 win;       // the last move completed a line
  };
  // One 72-byte row (nine u64) per 7-bit mask L of non-full columns (connect_four.cc:108-116):
  //   [0] = floor(2^32 / n) | (u64)(-n) << 32   (n = popc L: the fast-modulo constants),
  //   [1 + k] = for k < n the 64-bit cell mask of the k-th open column, ascending.
  // Nine words per row: rows of different L start in different shared-memory banks.
#ifndef BSG_C4_GROUP
#define BSG_C4_GROUP 4
#endif
  static constexpr int kGroupSteps = BSG_C4_GROUP;   // steps between two refills of the lean kernel
#ifndef BSG_C4_MINBLOCKS
#define BSG_C4_MINBLOCKS 1
#endif
  static constexpr int kMinBlocks = BSG_C4_MINBLOCKS;   // A/B switch: register cap of the lean kernel
#ifndef BSG_C4_LAZY_DRAW
#define BSG_C4_LAZY_DRAW 1
#endif
  // A full board has no open column: table row 0 hands step() an empty column mask, nothing is
  // placed, and neither side can own a line (a line ends the game when it is made) — so the lean
  // kernel finds draws from the move count once per group (playout_kernel, kLazyDraw).
  static constexpr bool kInertWhenFull = BSG_C4_LAZY_DRAW != 0;
  BSG_HD static u32 moves_left(const State& s) { return s.left; }
  static constexpr int kRowWords = 9;
  struct Tables { u64 row[128][kRowWords]; };
  static constexpr u64 kColumn = 0x0000010101010101ull;  // the six cells of column 0
  static void fill_tables(Tables& t) {
    for (u32 L = 0; L < 128; ++L) {
      u32 n = 0;
      for (int w = 0; w < kRowWords; ++w) t.row[L][w] = 0;
      for (u32 c = 0; c < 7; ++c)
        if ((L >> c) & 1) t.row[L][1 + n++] = kColumn << c;
      // L = 0 only happens on finished lanes running dead steps: n = 1 keeps k = 0 and the
      // empty column mask places nothing.
      u32 nn = n ? n : 1;
      t.row[L][0] = (u64)modM_entry(nn) | ((u64)(0u - nn) << 32);
    }
  }
  BSG_HD static void init(State& s) { s.cur = 0; s.opp = 0; s.left = 42; s.win = 0; }
  BSG_HD static void load(State& s, const void* base, int64_t n, int64_t i) {
    (void)n;
    u64 w = ((const u64*)base)[i];
    u32 top = (u32)(w >> 32);
    u64 key = w & 0x00FFFFFFFFFFFFFFull;
    u64 sm = key;             // smear every marker (and stone) down its column
    sm |= sm >> 8;
    sm |= sm >> 16;
    sm |= sm >> 32;
    u64 occ = sm >> 8;        // strictly below the markers
    s.cur = key & occ;
    s.opp = occ & ~key;
    s.left = top >> 26;
    s.win = (top >> 25) & 1u;
  }
  // Compact root format for host -> device transfers (bsg_playouts_host_compact): a position is
  // the mover's stones | one marker per column, seven rows of seven columns = 49 bits.  The packed
  // u64 spends a byte per row; the compact format drops every row's unused eighth bit and stores
  // the 49-bit strings back to back (root i at bit 49 * i, little endian).  The meta byte is
  // derived: empty cells from the occupancy, "the last move made four in a row" from the line
  // test on the side that moved last.  The copy of the roots over PCIe is the end-to-end bound
  // of the path, so 49 bits instead of 64 is a quarter off the time.
  static constexpr int kCompactBits = 49;
  BSG_HD static void load_compact(State& s, const void* base, int64_t i) {
    const int64_t bit = 49 * i;
    const u32* p = (const u32*)base + (bit >> 5);          // (the buffer has >= 8 bytes of slack behind it)
    const int sh = (int)(bit & 31);
    const u32 w0 = p[0], w1 = p[1], w2 = p[2];
    const u32 v0 = fshr(w0, w1, sh), v1 = fshr(w1, w2, sh);      // bits 0..31 and 32..63 of the string
    // rows 0..3 (28 bits) and rows 4..6 (21 bits), then every row from its 7-bit slot to its byte
    u32 lo = v0 & 0x0FFFFFFFu, hi = fshr(v0, v1, 28) & 0x001FFFFFu;
    lo = (lo & 0x00003FFFu) | ((lo & 0x0FFFC000u) << 2);
    hi = (hi & 0x00003FFFu) | ((hi & 0x0FFFC000u) << 2);
    lo = (lo & 0x007F007Fu) | ((lo & 0x3F803F80u) << 1);
    hi = (hi & 0x007F007Fu) | ((hi & 0x3F803F80u) << 1);
    const u64 key = (u64)lo | ((u64)hi << 32);
    u64 sm = key;
    sm |= sm >> 8;
    sm |= sm >> 16;
    sm |= sm >> 32;
    const u64 occ = sm >> 8;
    s.cur = key & occ;
    s.opp = occ & ~key;
    s.left = 42u - (u32)popc64(occ);
    s.win = has_line(s.opp) ? 1u : 0u;
  }
  // (host side; the buffer starts zeroed and the roots are stored in ascending order)
  BSG_HD static void store_compact(const State& s, void* base, int64_t i) {
    u64 tmp = 0;
    store(s, &tmp, 1, 0);
    u64 str = 0;
    for (int r = 0; r < 7; ++r) str |= ((tmp >> (8 * r)) & 0x7Full) << (7 * r);
    unsigned char* b = (unsigned char*)base + ((49 * i) >> 3);
    const int sh = (int)((49 * i) & 7);
    for (int k = 0; k < 8; ++k) {                         // 49 + 7 bits span at most 7 bytes
      const unsigned v = (unsigned)((k == 0 ? (str << sh) : (str >> (8 * k - sh))) & 0xFFu);
      if (v) b[k] = (unsigned char)(b[k] | v);
    }
  }
  BSG_HD static void store(const State& s, void* base, int64_t n, int64_t i) {
    (void)n;
    u64 occ = s.cur | s.opp, markers = 0;
    for (int c = 0; c < 7; ++c) {
      u64 e = ~occ & (0x0001010101010101ull << c);   // 7 rows: a full column marks row 6
      markers |= e & (0ull - e);
    }
    u64 meta = (u64)((s.win << 1) | (s.left << 2));
    ((u64*)base)[i] = s.cur | markers | (meta << 56);
  }
  BSG_HD static bool terminal(const State& s) { return (s.win != 0) || (s.left == 0); }
  BSG_HD static bool won(const State& s) { return s.win != 0; }
  BSG_HD static int mover(const State& s) { return (int)(s.left & 1u); }  // 42 is even
  BSG_HD static int outcome(const State& s) {
    return s.win ? (mover(s) ^ 1) : (s.left == 0 ? kOutcomeDraw : kOutcomeNone);
  }

  // Four-in-a-row anywhere on one player's board.  The reference scans every cell in four
  // directions (connect_four.cc:123-152); bit-parallel that is b & b>>d & b>>2d & b>>3d.
  // Rows: d = 1, per 32-bit word (rows do not straddle words), with left shifts (they can run
  // on the multiplier pipe).  Columns and diagonals (d = 8, 7, 9): the lowest cell of such a
  // line is in rows 0..2, so only the LOW word of the product is needed — three funnel
  // shifts of (hi:lo) and two ANDs per direction.  (Measured and rejected: running some of
  // the funnel shifts on the multiplier pipe as IMAD.HI + IMAD; IMAD.HI issues every 4 cycles
  // and every variant was slower than the plain SHF form.)
  BSG_HD static bool has_line(u64 b) {
    u32 lo = (u32)b, hi = (u32)(b >> 32);
    u32 ml = lo & (lo << 1), mh = hi & (hi << 1);
    u32 acc = (ml & (ml << 2)) | (mh & (mh << 2));
    acc |= lo & fshr(lo, hi, 8) & fshr(lo, hi, 16) & fshr(lo, hi, 24);
    acc |= lo & fshr(lo, hi, 7) & fshr(lo, hi, 14) & fshr(lo, hi, 21);
    acc |= lo & fshr(lo, hi, 9) & fshr(lo, hi, 18) & fshr(lo, hi, 27);
    return acc != 0;
  }
  // connect_four.cc:93-106: the stone lands on the lowest empty cell of the column (lowest
  // set bit of the column's empties), win tested before draw, then the turn passes.
  BSG_HD static void drop_mask(State& s, u64 column) {
    u64 e = column & ~(s.cur | s.opp);
    u64 b = s.cur | (e & (0ull - e));
    s.win = has_line(b) ? 1u : 0u;
    s.cur = s.opp;
    s.opp = b;
    s.left -= 1u;
  }
  BSG_HD static u32 open_columns(const State& s) {
    return ~(u32)((s.cur | s.opp) >> 40) & 0x7Fu;  // connect_four.cc:108-116
  }
  BSG_HD static u32 step(State& s, u32 r, const Tables& tab, int& action) {
    // row L of the table straight from the top-row word (L << 8), without isolating L first
    const u32 t = ~(u32)((s.cur | s.opp) >> 32) & 0x7F00u;
#if defined(__CUDA_ARCH__)
    // 32-bit shared-window addressing (generic pointers would cost 64-bit address arithmetic)
    const u32 off = (u32)__cvta_generic_to_shared(&tab) + ((t * 9u) >> 5);   // + L * 72
    u32 M, negn, c_lo, c_hi;
    asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(M), "=r"(negn) : "r"(off));
    u32 rem = umulhi32(r, M) * negn + r;            // r - floor-ish(r / n) * n, in [0, 2n)
    u32 k = umin32(rem, rem + negn);                // r mod n
    asm("ld.shared.v2.u32 {%0, %1}, [%2+8];" : "=r"(c_lo), "=r"(c_hi) : "r"(off + k * 8u));
    u64 column = (u64)c_lo | ((u64)c_hi << 32);
#else
    const u64* row = tab.row[t >> 8];
    const u64 mc = row[0];
    const u32 negn = (u32)(mc >> 32);
    u32 rem = umulhi32(r, (u32)mc) * negn + r;
    u32 k = umin32(rem, rem + negn);
    u64 column = row[1 + k];
#endif
    action = ctz64(column);                          // column index = lowest bit (row 0 cell)
    drop_mask(s, column);
    return 0u - negn;
  }
  BSG_HD static int legal(const State& s, int* out) {
    if (terminal(s)) return 0;
    u32 L = open_columns(s);
    int n = 0;
    for (int c = 0; c < 7; ++c) if ((L >> c) & 1) out[n++] = c;
    return n;
  }
  BSG_HD static bool apply(State& s, int a) {
    if (terminal(s) || a < 0 || a >= 7 || !((open_columns(s) >> a) & 1)) return false;
    drop_mask(s, kColumn << a);
    return true;
  }
  BSG_HD static u32 legal_bit(const State& s, int a) {
    return (!terminal(s) && ((open_columns(s) >> a) & 1u)) ? 1u : 0u;
  }
  // connect_four.cc:198-221: plane 0 = observer's opponent, 1 = observer's own, 2 = empty;
  // element order plane * 42 + row * 7 + col.
  BSG_HD static u32 obs_bit(const State& s, int player, int idx) {
    int plane = idx / 42, cell = idx - plane * 42;
    int row = cell / 7, col = cell - row * 7;
    u64 mine = (player == mover(s)) ? s.cur : s.opp;
    u64 theirs = (player == mover(s)) ? s.opp : s.cur;
    u64 m = plane == 0 ? theirs : (plane == 1 ? mine : ~(mine | theirs));
    return (u32)(m >> (row * 8 + col)) & 1u;
  }
  // Board with 8-bit rows -> the reference's 42-cell order (7-bit rows, cell = row * 7 + col).
  BSG_HD static u64 squeeze(u64 b) {
    u64 r = 0;
#pragma unroll
    for (int row = 0; row < 6; ++row) r |= ((b >> (8 * row)) & 0x7Full) << (7 * row);
    return r;
  }
  BSG_HD static void obs_words(const State& s, int player, u32* out) {
    u64 mine = squeeze((player == mover(s)) ? s.cur : s.opp);
    u64 theirs = squeeze((player == mover(s)) ? s.opp : s.cur);
    u64 empty = ~(mine | theirs) & 0x3FFFFFFFFFFull;
    u64 lo = theirs | (mine << 42);               // plane 0 at bit 0, plane 1 at bit 42
    u64 hi = (mine >> 22) | (empty << 20);        // plane 2 at bit 84
    out[0] = (u32)lo; out[1] = (u32)(lo >> 32); out[2] = (u32)hi; out[3] = (u32)(hi >> 32);
  }
  BSG_HD static void mask_words(const State& s, u32* out) {
    out[0] = terminal(s) ? 0u : open_columns(s);
  }
};

// =========================================================================================
// breakthrough 8x8 — reference: open_spiel/games/breakthrough.cc
//   cell = row * 8 + col; black = player 0 starts on rows 0-1 and moves to row + 1, white on
//   rows 6-7 moves to row - 1 (breakthrough.cc:36-40,119-142).  Two u64 boards kept relative
//   to the mover (own / opp) in ABSOLUTE orientation; the colour only changes shift amounts,
//   expressed as 64-bit rotates so both colours run the same instructions.
//   Action id = ((row*8 + col)*6 + dir)*2 + capture (spiel_utils.cc:45-58, bases {8,8,6,2}),
//   dir = 3*colour + o, o = 0,1,2 <-> dcol = -1,0,+1.  Ascending id order = ascending source
//   cell, then o: the k-th move is found by a weighted rank-select over two bit planes
//   (low = L^F^R, high = maj(L,F,R) of the per-source legality masks L, F, R).
//   packed: SoA plane 0 = 16 B {black, white}; plane 1 = u32 meta
//   (mover | terminal << 1 | outcome << 2).
// =========================================================================================
struct Breakthrough {
  static constexpr int kId = kGameBreakthrough;
  static constexpr int kNumActions = 768, kMaxLen = 209, kMaxDepth = kMaxLen, kObsSize = 192, kStateBytes = 20;
  static constexpr int kObsWords = 6, kMaskWords = 24;
  static constexpr int kPlanes = 2;
  BSG_HD static int plane_bytes(int p) { return p == 0 ? 16 : 4; }
  // A non-terminal state without a legal action cannot arise in play: take
  // the mover's most advanced piece; it is not on the goal row (the game would be over), no own
  // piece stands on the row ahead of it, and at least one of its two diagonal targets is on the
  // board — empty (a move) or enemy (a capture).  Moves keep that invariant, so only a ROOT
  // handed in from outside can be stuck; the kernel tests roots once (has_move) and the
  // per-step test is gone.
  static constexpr bool kCanStick = true;
  static constexpr bool kAsync = false;
  static constexpr bool kMoverCanLose = false;
  static constexpr bool kLazyEnd = false;       // the lean kernel sees a terminal state in the step that reaches it (see Othello)
  static constexpr u64 kCol0 = 0x0101010101010101ull, kCol7 = 0x8080808080808080ull;
  struct State { u64 own, opp; u32 mv, term, outc; };
  // sel4: the last two levels of the weighted rank-select as one look-up.  Index = h4 << 8 |
  // l4 << 4 | k, where cell i of a group of four has l_i + 2 h_i moves and k < 12 is the rank
  // inside the group; value = cell | (rank inside the cell) << 3.
  // selo: direction of a piece's rank-th move; index = rank << 3 | R << 2 | F << 1 | L.
  struct Tables { u32 modM[64]; unsigned char sel4[4096]; unsigned char selo[32]; };
  static void fill_tables(Tables& t) {
    for (u32 n = 0; n < 64; ++n) t.modM[n] = modM_entry(n);
    for (u32 h = 0; h < 16; ++h)
      for (u32 l = 0; l < 16; ++l)
        for (u32 k = 0; k < 16; ++k) {
          u32 rest = k, cell = 0, v = 0;
          for (; cell < 4; ++cell) {
            u32 c = ((l >> cell) & 1u) + 2u * ((h >> cell) & 1u);
            if (rest < c) { v = cell | (rest << 3); break; }
            rest -= c;
          }
          t.sel4[(h << 8) | (l << 4) | k] = (unsigned char)v;
        }
    for (u32 i = 0; i < 32; ++i) {
      u32 d = i & 7u, rank = i >> 3, o = 0;
      for (u32 b = 0, seen = 0; b < 3; ++b)
        if ((d >> b) & 1u) { if (seen == rank) o = b; ++seen; }
      t.selo[i] = (unsigned char)o;
    }
  }

  BSG_HD static void init(State& s) {
    s.own = 0x000000000000FFFFull; s.opp = 0xFFFF000000000000ull;
    s.mv = 0; s.term = 0; s.outc = kOutcomeNone;
  }
  BSG_HD static void load(State& s, const void* base, int64_t n, int64_t i) {
#if defined(__CUDA_ARCH__)
    ulonglong2 w = ((const ulonglong2*)base)[i];
    u64 b = w.x, wh = w.y;
#else
    u64 b = ((const u64*)base)[2 * i], wh = ((const u64*)base)[2 * i + 1];
#endif
    u32 meta = ((const u32*)((const char*)base + 16 * n))[i];
    s.mv = meta & 1; s.term = (meta >> 1) & 1; s.outc = (meta >> 2) & 3;
    s.own = s.mv ? wh : b;
    s.opp = s.mv ? b : wh;
  }
  BSG_HD static void store(const State& s, void* base, int64_t n, int64_t i) {
    ((u64*)base)[2 * i] = s.mv ? s.opp : s.own;
    ((u64*)base)[2 * i + 1] = s.mv ? s.own : s.opp;
    ((u32*)((char*)base + 16 * n))[i] = s.mv | (s.term << 1) | (s.outc << 2);
  }
  BSG_HD static bool terminal(const State& s) { return s.term != 0; }
  BSG_HD static int outcome(const State& s) { return (int)s.outc; }
  BSG_HD static bool won(const State& s) { return s.outc < 2u; }  // the last mover won
  BSG_HD static int mover(const State& s) { return (int)s.mv; }

  BSG_HD static u64 rotr64(u64 x, int sh) { return (x >> sh) | (x << ((64 - sh) & 63)); }

  // Per-source legality masks (breakthrough.cc:217-256): forward needs an empty target,
  // diagonals need a target that is not the mover's own piece (empty = move, enemy = capture).
  // Target of source i is i + 8 - 1/0/+1 for black, i - 8 - 1/0/+1 for white; reading the
  // target's occupancy at the source position is a rotate right by 7/8/9 resp. 55/56/57.
  BSG_HD static void movegen(const State& s, u64& L, u64& F, u64& R) {
    u64 occ = s.own | s.opp;
#if defined(__CUDA_ARCH__)
    // A variable 64-bit rotate costs the compiler four shifts and two merges.  By hand: white's
    // 55/56/57 = 32 + 23/24/25, so swap the words for white (one select each) and rotate by
    // 7/8/9 + 16 * mover < 32 with two funnel shifts.
    const bool w = s.mv != 0;
    const u32 ol = (u32)s.own, oh = (u32)(s.own >> 32), cl = (u32)occ, ch = (u32)(occ >> 32);
    const u32 al = w ? oh : ol, ah = w ? ol : oh, bl = w ? ch : cl, bh = w ? cl : ch;
    const u32 sh = s.mv * 16u + 7u;
    const u32 r7l = __funnelshift_r(al, ah, sh), r7h = __funnelshift_r(ah, al, sh);
    const u32 r8l = __funnelshift_r(bl, bh, sh + 1u), r8h = __funnelshift_r(bh, bl, sh + 1u);
    const u32 r9l = __funnelshift_r(al, ah, sh + 2u), r9h = __funnelshift_r(ah, al, sh + 2u);
    F = (u64)(ol & ~r8l) | ((u64)(oh & ~r8h) << 32);
    L = (u64)(ol & 0xFEFEFEFEu & ~r7l) | ((u64)(oh & 0xFEFEFEFEu & ~r7h) << 32);
    R = (u64)(ol & 0x7F7F7F7Fu & ~r9l) | ((u64)(oh & 0x7F7F7F7Fu & ~r9h) << 32);
#else
    int base = s.mv ? 48 : 0;
    F = s.own & ~rotr64(occ, base + 8);
    L = s.own & ~kCol0 & ~rotr64(s.own, base + 7);
    R = s.own & ~kCol7 & ~rotr64(s.own, base + 9);
#endif
  }
  BSG_HD static bool has_move(const State& s) {
    u64 L, F, R;
    movegen(s, L, F, R);
    return (L | F | R) != 0;
  }
  // breakthrough.cc:152-192: move the piece, remove a captured piece, win on reaching the far
  // row; :306-318: the game also ends when a side has no pieces left.
  BSG_HD static int move(State& s, int src, int o) {
    int dst = src + (s.mv ? -8 : 8) + (o - 1);
    u64 sb = 1ull << src, db = 1ull << (dst & 63);
    int cap = (int)((s.opp >> (dst & 63)) & 1ull);
    u64 own = (s.own & ~sb) | db;
    u64 opp = s.opp & ~db;
    bool reach = s.mv ? (dst < 8) : (dst >= 56);
    bool win = reach || opp == 0;
    int action = ((src * 6) + (int)(3 * s.mv) + o) * 2 + cap;
    s.term = win ? 1u : 0u;
    s.outc = win ? s.mv : (u32)kOutcomeNone;
    s.own = opp;
    s.opp = own;
    s.mv ^= 1u;
    return action;
  }
  BSG_HD static u32 step(State& s, u32 r, const Tables& tab, int& action) {
    u64 L, F, R;
    movegen(s, L, F, R);
    u64 lo = L ^ F ^ R, hi = (L & F) | (L & R) | (F & R);
    u32 n = (u32)(popc64(lo) + 2 * popc64(hi));
    int k = (int)fastmod(r, n, tab.modM[n & 63]);
    // weighted rank-select: first source cell i whose inclusive prefix count exceeds k
    u32 l32 = (u32)lo, h32 = (u32)hi;
    int pos = 0;
    int c = popc32(l32) + 2 * popc32(h32);
    if (k >= c) { k -= c; pos = 32; l32 = (u32)(lo >> 32); h32 = (u32)(hi >> 32); }
#pragma unroll
    for (int w = 16; w >= 4; w >>= 1) {
      // low w bits counted through a left shift (a multiply-pipe instruction) instead of a mask
      c = popc32(l32 << (32 - w)) + 2 * popc32(h32 << (32 - w));
      if (k >= c) { k -= c; pos += w; l32 >>= w; h32 >>= w; }
    }
    const u32 v = tab.sel4[((h32 << 8) & 0xF00u) | ((l32 << 4) & 0xF0u) | ((u32)k & 0xFu)];
    pos = (pos + (int)(v & 3u)) & 63;
    // the piece's moves in o = 0,1,2 order; v >> 3 is the rank among them
    const u32 d = (u32)((L >> pos) & 1ull) | ((u32)((F >> pos) & 1ull) << 1) | ((u32)((R >> pos) & 1ull) << 2);
    const int o = (int)tab.selo[(v & 0x18u) | d];
    action = move(s, pos, o);
    return n;
  }
  BSG_HD static int legal(const State& s, int* out) {
    if (terminal(s)) return 0;
    u64 L, F, R;
    movegen(s, L, F, R);
    int n = 0;
    for (int i = 0; i < 64; ++i) {
      for (int o = 0; o < 3; ++o) {
        u64 m = o == 0 ? L : (o == 1 ? F : R);
        if (!((m >> i) & 1ull)) continue;
        int dst = i + (s.mv ? -8 : 8) + (o - 1);
        int cap = (int)((s.opp >> dst) & 1ull);
        out[n++] = ((i * 6) + (int)(3 * s.mv) + o) * 2 + cap;
      }
    }
    return n;
  }
  BSG_HD static bool apply(State& s, int a) {
    int list[64];
    int n = legal(s, list);
    bool ok = false;
    for (int i = 0; i < n; ++i) ok |= (list[i] == a);
    if (!ok) return false;
    int dir = (a / 2) % 6, src = a / 12;
    move(s, src, dir - 3 * (int)s.mv);
    return true;
  }
  BSG_HD static u32 legal_bit(const State& s, int a) {
    int cap = a & 1, dir = (a >> 1) % 6, src = a / 12;
    int o = dir - 3 * (int)s.mv;
    if (s.term || o < 0 || o > 2) return 0u;
    u64 L, F, R;
    movegen(s, L, F, R);
    u64 m = o == 0 ? L : (o == 1 ? F : R);
    int dst = (src + (s.mv ? -8 : 8) + (o - 1)) & 63;
    return (((m >> src) & 1ull) && (int)((s.opp >> dst) & 1ull) == cap) ? 1u : 0u;
  }
  // breakthrough.cc:284-304,326-340: plane 0 black, 1 white, 2 empty, same for both observers.
  BSG_HD static u32 obs_bit(const State& s, int player, int idx) {
    (void)player;
    int plane = idx >> 6, cell = idx & 63;
    u64 black = s.mv ? s.opp : s.own, white = s.mv ? s.own : s.opp;
    u64 m = plane == 0 ? black : (plane == 1 ? white : ~(black | white));
    return (u32)(m >> cell) & 1u;
  }
  BSG_HD static void obs_words(const State& s, int player, u32* out) {
    (void)player;
    u64 black = s.mv ? s.opp : s.own, white = s.mv ? s.own : s.opp, empty = ~(black | white);
    out[0] = (u32)black; out[1] = (u32)(black >> 32);
    out[2] = (u32)white; out[3] = (u32)(white >> 32);
    out[4] = (u32)empty; out[5] = (u32)(empty >> 32);
  }
  // 12 action bits per source cell: bit (3*mover + o)*2 + capture.  Built for mover 0 (six
  // bits per cell at 12*cell), then the whole string moves up by 6 bits for mover 1.
  BSG_HD static void mask_words(const State& s, u32* out) {
    u64 L, F, R;
    movegen(s, L, F, R);
    if (s.term) { L = 0; F = 0; R = 0; }
    int base = s.mv ? 48 : 0;
    u64 CL = rotr64(s.opp, base + 7), CR = rotr64(s.opp, base + 9);  // enemy on the diagonal target
    u64 a0 = L & ~CL, a1 = L & CL, a4 = R & ~CR, a5 = R & CR;
#pragma unroll
    for (int w = 0; w < kMaskWords; ++w) out[w] = 0;
#pragma unroll
    for (int i = 0; i < 64; ++i) {
      u32 pat = (u32)((a0 >> i) & 1ull) | ((u32)((a1 >> i) & 1ull) << 1) |
                ((u32)((F >> i) & 1ull) << 2) | ((u32)((a4 >> i) & 1ull) << 4) |
                ((u32)((a5 >> i) & 1ull) << 5);
      const int pos = 12 * i, w = pos >> 5, sh = pos & 31;
      out[w] |= pat << sh;
      if (sh > 26) out[w + 1] |= pat >> (32 - sh);
    }
    if (s.mv) {
#pragma unroll
      for (int w = kMaskWords - 1; w >= 1; --w) out[w] = (out[w] << 6) | (out[w - 1] >> 26);
      out[0] <<= 6;
    }
  }
};

// =========================================================================================
// hex N x N (N = 11 in the product; N = 5 only for the reference's golden playthrough) —
// reference: open_spiel/games/hex.cc
//   cell = row * N + col in a 128-bit mask (4 x u32).  Black = player 0 moves first and
//   connects row 0 ("north") with row N-1; white connects col 0 ("west") with col N-1
//   (hex.cc:55-121).  The reference keeps, per stone, whether its group touches one of its
//   colour's edges (9 cell states, hex.h:51-61) and floods that label through adjacent PLAIN
//   stones of the mover when an edge-connected stone is placed (hex.cc:149-184).  Here the same
//   sets are bit masks per side — O (all stones), A (north / west connected), B (south / east
//   connected); plain = O & ~A & ~B — and the flood is a bit-parallel dilation.
//
//   The flood is ASYNCHRONOUS on the device (kAsync): placing a stone only records the first
//   frontier (the plain neighbours of the new stone, one table look-up); every later kernel
//   iteration runs ONE dilation ring for every lane (straight-line code) and a lane makes its
//   next move only when its frontier is empty.  Flood depth has a fat tail in random play
//   (11 % of the moves flood, 1.4 % need >= 3 rings, some 15+); a per-move flood loop made the
//   whole warp wait for its deepest lane (measured: 58 % of the kernel's instructions ran in
//   that loop with 2.3 of 32 lanes active).  Host helpers flood synchronously (`finish`).
//   A winning stone is a member of both A and B; no flood runs on the winning move.
//   packed: SoA, 4 planes of 16 B: black, white, edgeA (north|west), edgeB (south|east).
// =========================================================================================
template <int N>
struct HexT {
  static constexpr int kId = kGameHex;
  static constexpr int kCells = N * N;
  static constexpr int kNumActions = N * N, kMaxLen = N * N, kMaxDepth = kMaxLen, kObsSize = 9 * N * N;
  static constexpr int kStateBytes = 64;
  static constexpr int kObsWords = (9 * N * N + 31) / 32, kMaskWords = 4;
  static constexpr int kPlanes = 4;
  BSG_HD static int plane_bytes(int p) { (void)p; return 16; }
  static constexpr bool kCanStick = false;
  static constexpr bool kAsync = true;
#ifndef BSG_HEX_MINBLOCKS
#define BSG_HEX_MINBLOCKS 1   // A/B switch: resident CTAs per SM the lean kernel is register-capped for
#endif
  static constexpr int kMinBlocks = BSG_HEX_MINBLOCKS;
  static constexpr bool kMoverCanLose = false;
  static constexpr bool kLazyEnd = false;       // the lean kernel sees a terminal state in the step that reaches it (see Othello)
  struct State {
    u32 cO[4], cA[4], cB[4];   // side to move: all stones / edge-A connected / edge-B connected
    u32 oO[4], oA[4], oB[4];   // the side that moved last
    u32 F[4];                  // pending flood frontier: plain stones of the o side that are
                               // adjacent to newly labelled ones and not labelled yet
    u32 fA;                    // ~0: the pending flood spreads label A; 0: label B
    u32 left;                  // empty cells
    u32 mv;                    // player to move (0 black, 1 white)
    u32 win;                   // the last move connected both edges
  };
  // nb[cell]: neighbour mask; word 3 bits 28..31 carry the edge flags of the cell:
  // bit 28 row 0, 29 row N-1, 30 col 0, 31 col N-1 (never board cells since N*N <= 124).
  // one[cell]: one-hot mask.  sel8[b][k]: position of the k-th set bit of byte b.
  struct Tables {
    u32 nb[128][4];
    u32 one[128][4];
    u32 modM[128];
    unsigned char sel8[256][8];
  };
  static void fill_tables(Tables& t) {
    for (u32 n = 0; n < 128; ++n) t.modM[n] = modM_entry(n);
    for (int c = 0; c < 128; ++c)
      for (int w = 0; w < 4; ++w) { t.nb[c][w] = 0; t.one[c][w] = (w == (c >> 5)) ? 1u << (c & 31) : 0u; }
    for (int b = 0; b < 256; ++b) {
      int k = 0;
      for (int i = 0; i < 8; ++i) t.sel8[b][i] = 0;
      for (int i = 0; i < 8; ++i) if ((b >> i) & 1) t.sel8[b][k++] = (unsigned char)i;
    }
    for (int cell = 0; cell < kCells; ++cell) {
      int row = cell / N, col = cell % N;
      const int dr[6] = {-1, -1, 0, 0, 1, 1}, dc[6] = {0, 1, -1, 1, -1, 0};  // hex.cc:206-222
      for (int k = 0; k < 6; ++k) {
        int r2 = row + dr[k], c2 = col + dc[k];
        if (r2 < 0 || r2 >= N || c2 < 0 || c2 >= N) continue;
        int v = r2 * N + c2;
        t.nb[cell][v >> 5] |= 1u << (v & 31);
      }
      if (row == 0) t.nb[cell][3] |= 1u << 28;
      else if (row == N - 1) t.nb[cell][3] |= 1u << 29;   // "else if" as in hex.cc:70-74
      if (col == 0) t.nb[cell][3] |= 1u << 30;
      else if (col == N - 1) t.nb[cell][3] |= 1u << 31;   // hex.cc:95-99
    }
  }
  // Compile-time masks of word w: board cells, cells not in column 0, cells not in column N-1.
  BSG_HDC static u32 word_mask(int w, int kind) {
    u32 m = 0;
    for (int b = 0; b < 32; ++b) {
      int cell = 32 * w + b;
      if (cell >= kCells) break;
      int col = cell % N;
      bool in = kind == 0 ? true : (kind == 1 ? col != 0 : col != N - 1);
      if (in) m |= 1u << b;
    }
    return m;
  }
  BSG_HD static u32 board_word(int w) {
    constexpr u32 m[4] = {word_mask(0, 0), word_mask(1, 0), word_mask(2, 0), word_mask(3, 0)};
    return m[w];
  }
  BSG_HD static u32 not_col0(int w) {
    constexpr u32 m[4] = {word_mask(0, 1), word_mask(1, 1), word_mask(2, 1), word_mask(3, 1)};
    return m[w];
  }
  BSG_HD static u32 not_col_last(int w) {
    constexpr u32 m[4] = {word_mask(0, 2), word_mask(1, 2), word_mask(2, 2), word_mask(3, 2)};
    return m[w];
  }

  BSG_HD static void init(State& s) {
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      s.cO[w] = s.cA[w] = s.cB[w] = s.oO[w] = s.oA[w] = s.oB[w] = 0;
      s.F[w] = 0;
    }
    s.fA = 0; s.left = kCells; s.mv = 0; s.win = 0;
  }
  BSG_HD static void load(State& s, const void* base, int64_t n, int64_t i) {
    u32 bl[4], wh[4], ea[4], eb[4];
#if defined(__CUDA_ARCH__)
    const uint4* p = (const uint4*)base;
    uint4 q0 = p[i], q1 = p[n + i], q2 = p[2 * n + i], q3 = p[3 * n + i];
    bl[0] = q0.x; bl[1] = q0.y; bl[2] = q0.z; bl[3] = q0.w;
    wh[0] = q1.x; wh[1] = q1.y; wh[2] = q1.z; wh[3] = q1.w;
    ea[0] = q2.x; ea[1] = q2.y; ea[2] = q2.z; ea[3] = q2.w;
    eb[0] = q3.x; eb[1] = q3.y; eb[2] = q3.z; eb[3] = q3.w;
#else
    const u32* p = (const u32*)base;
    for (int w = 0; w < 4; ++w) {
      bl[w] = p[4 * i + w]; wh[w] = p[4 * (n + i) + w];
      ea[w] = p[4 * (2 * n + i) + w]; eb[w] = p[4 * (3 * n + i) + w];
    }
#endif
    int nb = 0, nw = 0;
    u32 won = 0;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      nb += popc32(bl[w]); nw += popc32(wh[w]);
      won |= (bl[w] | wh[w]) & ea[w] & eb[w];
    }
    s.mv = nb > nw ? 1u : 0u;  // black moves first and turns alternate
    s.win = won ? 1u : 0u;
    s.left = (u32)(kCells - nb - nw);
    s.fA = 0;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      u32 mine = s.mv ? wh[w] : bl[w], theirs = s.mv ? bl[w] : wh[w];
      s.cO[w] = mine; s.cA[w] = mine & ea[w]; s.cB[w] = mine & eb[w];
      s.oO[w] = theirs; s.oA[w] = theirs & ea[w]; s.oB[w] = theirs & eb[w];
      s.F[w] = 0;
    }
  }
  BSG_HD static void store(const State& s0, void* base, int64_t n, int64_t i) {
    State s = s0;
    finish(s);   // a packed state always carries complete edge-connected sets
    u32* p = (u32*)base;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      p[4 * i + w] = s.mv ? s.oO[w] : s.cO[w];
      p[4 * (n + i) + w] = s.mv ? s.cO[w] : s.oO[w];
      p[4 * (2 * n + i) + w] = s.cA[w] | s.oA[w];
      p[4 * (3 * n + i) + w] = s.cB[w] | s.oB[w];
    }
  }
  BSG_HD static bool terminal(const State& s) { return s.win != 0; }
  BSG_HD static bool won(const State& s) { return s.win != 0; }
  BSG_HD static int mover(const State& s) { return (int)s.mv; }
  BSG_HD static int outcome(const State& s) { return s.win ? (int)(s.mv ^ 1u) : kOutcomeNone; }

  // One ring of neighbours of every cell in f (hex.cc:206-222 for a whole set at once),
  // restricted to `within`.  East/west neighbours are 128-bit shifts by one with the column
  // wrap masked; the two northern neighbours (r-1,c), (r-1,c+1) are (f | east) shifted one row
  // down in index, the two southern ones (f | west) shifted one row up.
  BSG_HD static void dilate_within(const u32 f[4], const u32 within[4], u32 out[4]) {
    u32 E[4], W[4], fe[4], fw[4], nn[4], ss[4];
    shl_words<1, 4>(f, E);
    shr_words<1, 4>(f, W);
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      fe[w] = f[w] | (E[w] & not_col0(w));          // a stone in the last column has no east neighbour
      fw[w] = f[w] | (W[w] & not_col_last(w));      // a stone in column 0 has no west neighbour
    }
    shr_words<N, 4>(fe, nn);
    shl_words<N, 4>(fw, ss);
    // (f itself rides along in fe | fw: every caller's `within` excludes the frontier)
#pragma unroll
    for (int w = 0; w < 4; ++w) out[w] = (fe[w] | fw[w] | nn[w] | ss[w]) & within[w];
  }
  // The pending flood (hex.cc:158-182, one ring per call): label the frontier, then the next
  // frontier is its ring of neighbours among the stones that are still plain.
  BSG_HD static void ring(State& s) {
    u32 P[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      s.oA[w] |= s.F[w] & s.fA;
      s.oB[w] |= s.F[w] & ~s.fA;
      P[w] = s.oO[w] & ~(s.oA[w] | s.oB[w]);
    }
    u32 d[4];
    dilate_within(s.F, P, d);
#pragma unroll
    for (int w = 0; w < 4; ++w) s.F[w] = d[w];
  }
  BSG_HD static bool pending(const State& s) { return (s.F[0] | s.F[1] | s.F[2] | s.F[3]) != 0; }
  BSG_HD static void finish(State& s) { while (pending(s)) ring(s); }

  // hex.cc:149-184 for the mover: label from edges + neighbours, win, or start the flood.
  BSG_HD static void place(State& s, int cell, const Tables& t) {
    u32 nb[4], m[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) { nb[w] = t.nb[cell][w]; m[w] = t.one[cell][w]; }
    u32 flags = nb[3] >> (28 + 2 * s.mv);  // bit 0: on edge A, bit 1: on edge B
    u32 hitA = 0, hitB = 0;
#pragma unroll
    for (int w = 0; w < 4; ++w) { hitA |= nb[w] & s.cA[w]; hitB |= nb[w] & s.cB[w]; }
    bool tA = (hitA != 0) || (flags & 1u), tB = (hitB != 0) || (flags & 2u);
    u32 toA = tA ? 0xFFFFFFFFu : 0u, toB = tB ? 0xFFFFFFFFu : 0u;
    u32 go = (tA != tB) ? 0xFFFFFFFFu : 0u;   // exactly one label: it spreads through plain stones
#if defined(__CUDA_ARCH__) && BSG_HEX_OPAQUE_MASKS
    // keep the three masks in registers: a | (m & mask) is then ONE 3-input LOP3 per word instead
    // of a select on the predicate followed by an OR
    asm("" : "+r"(toA));
    asm("" : "+r"(toB));
    asm("" : "+r"(go));
#endif
    u32 nO[4], nA[4], nB[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      nO[w] = s.cO[w] | m[w];
      nA[w] = s.cA[w] | (m[w] & toA);
      nB[w] = s.cB[w] | (m[w] & toB);
      u32 plain = nO[w] & ~(nA[w] | nB[w]);
      u32 f = nb[w] & plain & go;
      // pass the turn: the mover's sets become the o side
      s.cO[w] = s.oO[w]; s.cA[w] = s.oA[w]; s.cB[w] = s.oB[w];
      s.oO[w] = nO[w]; s.oA[w] = nA[w]; s.oB[w] = nB[w];
      s.F[w] = f;
    }
    s.fA = toA;
    s.win = (tA && tB) ? 1u : 0u;
    s.mv ^= 1u;
    s.left -= 1u;
  }
  BSG_HD static u32 empty_word(const State& s, int w) { return ~(s.cO[w] | s.oO[w]) & board_word(w); }
#ifndef BSG_HEX_OPAQUE_MASKS
#define BSG_HEX_OPAQUE_MASKS 1
#endif
  // k-th (ascending) empty cell, k < left.  Word by prefix popcounts, then halves of the word,
  // then a byte table.
  BSG_HD static int select_empty(const State& s, u32 k, const Tables& tab) {
    u32 e0 = empty_word(s, 0), e1 = empty_word(s, 1), e2 = empty_word(s, 2), e3 = empty_word(s, 3);
    u32 c0 = (u32)popc32(e0), c1 = c0 + (u32)popc32(e1), c2 = c1 + (u32)popc32(e2);
    bool g0 = k >= c0, g1 = k >= c1, g2 = k >= c2;
    u32 x = g2 ? e3 : (g1 ? e2 : (g0 ? e1 : e0));
    u32 below = g2 ? c2 : (g1 ? c1 : (g0 ? c0 : 0u));
    u32 pos = g2 ? 96u : (g1 ? 64u : (g0 ? 32u : 0u));
    u32 kk = k - below;
    u32 h = (u32)popc32(x & 0xFFFFu);
    if (kk >= h) { kk -= h; x >>= 16; pos += 16; }
    h = (u32)popc32(x & 0xFFu);
    if (kk >= h) { kk -= h; x >>= 8; pos += 8; }
    return (int)((pos + tab.sel8[x & 0xFFu][kk & 7u]) & 127u);
  }
  BSG_HD static u32 step(State& s, u32 r, const Tables& tab, int& action) {
    u32 n = s.left;                                        // hex.cc:186-196: every empty cell
    u32 k = fastmod(r, n, tab.modM[n & 127]);
    int cell = select_empty(s, k, tab);
    action = cell;
    place(s, cell, tab);
    return n;
  }
  BSG_HD static int legal(const State& s, int* out) {
    if (terminal(s)) return 0;
    int n = 0;
    for (int c = 0; c < kCells; ++c) if ((empty_word(s, c >> 5) >> (c & 31)) & 1u) out[n++] = c;
    return n;
  }
  BSG_HD static bool apply(State& s, int a, const Tables& t) {
    if (terminal(s) || a < 0 || a >= kCells || !((empty_word(s, a >> 5) >> (a & 31)) & 1u)) return false;
    place(s, a, t);
    finish(s);      // host helpers flood synchronously
    return true;
  }
  BSG_HD static u32 legal_bit(const State& s, int a) {
    return (!s.win && ((empty_word(s, a >> 5) >> (a & 31)) & 1u)) ? 1u : 0u;
  }
  // Per-word absolute sets for the observation: black/white x {A only, B only, plain, both}.
  BSG_HD static void abs_sets(const State& s, int w, u32& bA, u32& bB, u32& bO, u32& wA, u32& wB, u32& wO) {
    bA = s.mv ? s.oA[w] : s.cA[w]; bB = s.mv ? s.oB[w] : s.cB[w]; bO = s.mv ? s.oO[w] : s.cO[w];
    wA = s.mv ? s.cA[w] : s.oA[w]; wB = s.mv ? s.cB[w] : s.oB[w]; wO = s.mv ? s.cO[w] : s.oO[w];
  }
  // hex.cc:266-277: plane = CellState + 4: 0 WhiteWin, 1 WhiteWest, 2 WhiteEast, 3 White,
  // 4 Empty, 5 Black, 6 BlackSouth, 7 BlackNorth, 8 BlackWin; same for both observers.
  BSG_HD static u32 plane_word(const State& s, int plane, int w) {
    u32 bA, bB, bO, wA, wB, wO;
    abs_sets(s, w, bA, bB, bO, wA, wB, wO);
    switch (plane) {
      case 0: return wA & wB;
      case 1: return wA & ~wB;
      case 2: return wB & ~wA;
      case 3: return wO & ~(wA | wB);
      case 4: return empty_word(s, w);
      case 5: return bO & ~(bA | bB);
      case 6: return bB & ~bA;
      case 7: return bA & ~bB;
      default: return bA & bB;
    }
  }
  BSG_HD static u32 obs_bit(const State& s, int player, int idx) {
    (void)player;
    int plane = idx / kCells, cell = idx - plane * kCells;
    return (plane_word(s, plane, cell >> 5) >> (cell & 31)) & 1u;
  }
  BSG_HD static void obs_words(const State& s, int player, u32* out) {
    (void)player;
#pragma unroll
    for (int w = 0; w < kObsWords; ++w) out[w] = 0;
#pragma unroll
    for (int p = 0; p < 9; ++p) {
      const int off = p * kCells, wi = off >> 5, sh = off & 31;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        u32 m = plane_word(s, p, j);
        if (wi + j < kObsWords) out[wi + j] |= m << sh;
        if (sh != 0 && wi + j + 1 < kObsWords) out[wi + j + 1] |= m >> ((32 - sh) & 31);
      }
    }
  }
  BSG_HD static void mask_words(const State& s, u32* out) {
#pragma unroll
    for (int w = 0; w < 4; ++w) out[w] = s.win ? 0u : empty_word(s, w);
  }

  // ---- value-only playouts (bsg_playout_io.value_only).  Stones never leave a hex board and
  // a full board has exactly one winner, so the side that connects first is the side that is
  // connected when the same move sequence is played on until the board is full: the playout
  // needs no labels and no flood per move, only one flood of the full board afterwards
  // (fill_winner, run by a second kernel with every lane busy).  Same draws, same moves, same
  // winner as the exact playout; only the move count differs.
  static constexpr bool kFillable = true;
  static constexpr int kFillWords = 4;
  BSG_HD static u32 step_fill(State& s, u32 r, const Tables& tab, int& action) {
    const u32 n = s.left;
    const u32 k = fastmod(r, n, tab.modM[n & 127]);
    const int cell = select_empty(s, k, tab);
    action = cell;
    // a lane whose board is full runs dead steps until its warp's group ends: they must not
    // touch the board, which IS the result here
    const u32 live = n ? 0xFFFFFFFFu : 0u;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const u32 mine = s.cO[w] | (tab.one[cell][w] & live);
      s.cO[w] = s.oO[w];
      s.oO[w] = mine;
    }
    s.mv ^= 1u;
    s.left -= live & 1u;
    return n;
  }
  BSG_HD static bool fill_done(const State& s) { return s.left == 0 || s.win != 0; }
  // what the second kernel needs: black's stones (on a full board the rest is white's)
  BSG_HD static void fill_store(const State& s, u32* out) {
#pragma unroll
    for (int w = 0; w < 4; ++w) out[w] = s.mv ? s.oO[w] : s.cO[w];
  }
  // 0: black connects row 0 to row N-1 through black stones; 1: it does not (white wins a
  // full or finished board)
  BSG_HD static int fill_winner(const u32* black) {
    u32 reach[4], f[4], rest[4];
    constexpr int last = N * (N - 1);
    u32 any = 0;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      u32 row0 = 0;
      if (32 * w < N) row0 = (N - 32 * w >= 32) ? 0xFFFFFFFFu : ((1u << (N - 32 * w)) - 1u);
      f[w] = black[w] & row0;
      reach[w] = f[w];
      any |= f[w];
    }
    while (any) {
      u32 d[4];
#pragma unroll
      for (int w = 0; w < 4; ++w) rest[w] = black[w] & ~reach[w];
      dilate_within(f, rest, d);
      any = 0;
#pragma unroll
      for (int w = 0; w < 4; ++w) { f[w] = d[w]; reach[w] |= d[w]; any |= d[w]; }
    }
    u32 hit = 0;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      u32 m = 0;   // cells last .. N*N-1 of word w
      for (int b = 0; b < 32; ++b) { const int c = 32 * w + b; if (c >= last && c < N * N) m |= 1u << b; }
      hit |= reach[w] & m;
    }
    return hit ? 0 : 1;
  }
};
typedef HexT<11> Hex;

// =========================================================================================
// othello — reference: open_spiel/games/othello.cc   
//   8 x 8, cell = row * 8 + col, black = player 0 moves first (othello.cc:263-269).  A move
//   places a disc that brackets opponent discs in at least one of eight directions and flips
//   them (othello.cc:126-158,200-229); a player without such a move must PASS (action 64,
//   othello.cc:241-246) and the game ends right after a placement that leaves neither side a
//   move — disc majority wins, equal counts tie (othello.cc:214-226).  The winner is decided
//   by counting, so the last mover can win or lose (kMoverCanLose).
//   Bit = cell in one u64 per colour.  Legal moves: the parallel-prefix bracket scan in four
//   axis pairs; flips: eight directional fills from the new disc.  The mover's move set is
//   carried in the state (the terminal test needs the next mover's moves anyway), so a step
//   runs one move generation, or a second one in the rare case that the next mover has none
//   (the lean kernel defers that second one to an iteration of its own: step_lazy).
//   The reference reports MaxGameLength() = 64 (othello.h:142); with passes a game can take up
//   to 60 placements + 59 passes, which is what kMaxDepth bounds.
//   packed: plane 0: 16 B {black, white}; plane 1: u32 = mover | terminal << 1 | outcome << 2.
// =========================================================================================
struct Othello {
  static constexpr int kId = kGameOthello;
  static constexpr bool kRolledGroup = BSG_OTHELLO_ROLLED;   // playout kernel: step group not unrolled (A/B)
  static constexpr int kNumActions = 65, kMaxLen = 64, kMaxDepth = 120, kObsSize = 192, kStateBytes = 20;
  static constexpr int kObsWords = 6, kMaskWords = 3;
  static constexpr int kPlanes = 2;
  BSG_HD static int plane_bytes(int p) { return p == 0 ? 16 : 4; }
  static constexpr bool kCanStick = false;    // a player without a placement passes
  static constexpr bool kAsync = false;
  static constexpr bool kMoverCanLose = true;
  // The lean kernel learns that a playout is over one iteration late (step_lazy below): the
  // "neither side can move" test after a placement needs a second move generation that only
  // ~1 step in 40 uses but, as a branch, most warp-steps would pay for.
  static constexpr bool kLazyEnd = true;
  static constexpr int kPass = 64;
  static constexpr u64 kInner = 0x7E7E7E7E7E7E7E7Eull;   // columns 1..6: no wrap for +-1, +-7, +-9
  struct State {
    u64 own, opp;   // discs of the player to move / of the other player
    u64 moves;      // placements open to the player to move (0: must pass)
    u32 mv;         // player to move
    u32 term;       // neither side can move
    u32 res;        // at the end: 1 = the last mover has more discs, 2 = fewer, 0 = equal
  };
  struct Tables { u32 modM[64]; };
  static void fill_tables(Tables& t) { for (u32 n = 0; n < 64; ++n) t.modM[n] = modM_entry(n); }

  // Cells where `p` may place: empty cells at the far end of a run of `o` discs that starts
  // next to a `p` disc (othello.cc:126-158 for every cell at once).  One axis = two directions.
  BSG_HD static u64 axis_moves(u64 p, u64 mask, int d) {
    u64 fl = mask & (p << d), fr = mask & (p >> d);
    fl |= mask & (fl << d);   fr |= mask & (fr >> d);
    const u64 ml = mask & (mask << d), mr = mask & (mask >> d);
    fl |= ml & (fl << (2 * d)); fr |= mr & (fr >> (2 * d));
    fl |= ml & (fl << (2 * d)); fr |= mr & (fr >> (2 * d));
    return (fl << d) | (fr >> d);
  }
  BSG_HD static u64 get_moves(u64 p, u64 o) {
    const u64 inner = o & kInner;
    const u64 m = axis_moves(p, inner, 1) | axis_moves(p, o, 8) | axis_moves(p, inner, 7) | axis_moves(p, inner, 9);
    return m & ~(p | o);
  }
  // Discs flipped by a disc of `p` on bit x: per direction the run of `o` discs from x, kept
  // when a `p` disc ends it (othello.cc:126-146,160-177).
  BSG_HD static u64 get_flips(u64 x, u64 p, u64 o) {
    const u64 inner = o & kInner;
    u64 flips = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int d = i == 0 ? 1 : (i == 1 ? 8 : (i == 2 ? 7 : 9));
      const u64 mask = i == 1 ? o : inner;
      u64 fl = mask & (x << d), fr = mask & (x >> d);
#pragma unroll
      for (int k = 0; k < 5; ++k) { fl |= mask & (fl << d); fr |= mask & (fr >> d); }
      if (p & (fl << d)) flips |= fl;
      if (p & (fr >> d)) flips |= fr;
    }
    return flips;
  }

  BSG_HD static void init(State& s) {
    s.own = (1ull << 28) | (1ull << 35);   // black e4, d5
    s.opp = (1ull << 27) | (1ull << 36);   // white d4, e5
    s.mv = 0; s.term = 0; s.res = 0;
    s.moves = get_moves(s.own, s.opp);
  }
  BSG_HD static void load(State& s, const void* base, int64_t n, int64_t i) {
#if defined(__CUDA_ARCH__)
    ulonglong2 w = ((const ulonglong2*)base)[i];
    u64 b = w.x, wh = w.y;
#else
    u64 b = ((const u64*)base)[2 * i], wh = ((const u64*)base)[2 * i + 1];
#endif
    u32 meta = ((const u32*)((const char*)base + 16 * n))[i];
    s.mv = meta & 1u; s.term = (meta >> 1) & 1u;
    const u32 outc = (meta >> 2) & 3u;
    s.own = s.mv ? wh : b;
    s.opp = s.mv ? b : wh;
    s.moves = s.term ? 0ull : get_moves(s.own, s.opp);
    s.res = (!s.term || outc >= 2u) ? 0u : (outc == (s.mv ^ 1u) ? 1u : 2u);
  }
  BSG_HD static void store(const State& s, void* base, int64_t n, int64_t i) {
    ((u64*)base)[2 * i] = s.mv ? s.opp : s.own;
    ((u64*)base)[2 * i + 1] = s.mv ? s.own : s.opp;
    ((u32*)((char*)base + 16 * n))[i] = s.mv | (s.term << 1) | ((u32)outcome(s) << 2);
  }
  BSG_HD static bool same_state(const State& a, const State& b) {
    return a.own == b.own && a.opp == b.opp && a.moves == b.moves && a.mv == b.mv && a.term == b.term &&
           (!a.term || a.res == b.res);
  }
  BSG_HD static bool terminal(const State& s) { return s.term != 0; }
  BSG_HD static bool won(const State& s) { return s.term && s.res == 1u; }
  BSG_HD static bool lost(const State& s) { return s.term && s.res == 2u; }
  BSG_HD static int mover(const State& s) { return (int)s.mv; }
  BSG_HD static int outcome(const State& s) {
    if (!s.term) return kOutcomeNone;
    return s.res == 1u ? (int)(s.mv ^ 1u) : (s.res == 2u ? (int)s.mv : kOutcomeDraw);
  }

  // othello.cc:200-229.  pos = 64 is the pass.
  BSG_HD static void play(State& s, int pos) {
    if (pos == kPass) {
      const u64 t = s.own; s.own = s.opp; s.opp = t;
      s.mv ^= 1u;
      s.moves = get_moves(s.own, s.opp);
      return;
    }
    const u64 x = 1ull << pos;
    const u64 flips = get_flips(x, s.own, s.opp);
    const u64 mine = s.own | x | flips, theirs = s.opp & ~flips;
    u64 next = get_moves(theirs, mine);           // the other player's placements
    if (next == 0 && get_moves(mine, theirs) == 0) {
      const int a = popc64(mine), b = popc64(theirs);
      s.term = 1u;
      s.res = a > b ? 1u : (a < b ? 2u : 0u);
    }
    s.own = theirs;
    s.opp = mine;
    s.moves = next;
    s.mv ^= 1u;
  }
  // Flips for the device step: the same eight directional runs, filled by doubling (a run is
  // at most six discs: 1 + 1 + 2 + 2) like the move generator, three fill steps instead of six.
  BSG_HD static u64 get_flips_fast(u64 x, u64 p, u64 o) {
    const u64 inner = o & kInner;
    u64 flips = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int d = i == 0 ? 1 : (i == 1 ? 8 : (i == 2 ? 7 : 9));
      const u64 mask = i == 1 ? o : inner;
      u64 fl = mask & (x << d), fr = mask & (x >> d);
      fl |= mask & (fl << d);   fr |= mask & (fr >> d);
      const u64 ml = mask & (mask << d), mr = mask & (mask >> d);
      fl |= ml & (fl << (2 * d)); fr |= mr & (fr >> (2 * d));
      fl |= ml & (fl << (2 * d)); fr |= mr & (fr >> (2 * d));
      flips |= (p & (fl << d)) ? fl : 0ull;
      flips |= (p & (fr >> d)) ? fr : 0ull;
    }
    return flips;
  }
  // One iteration of the lean kernel, straight-line for every lane.  A state with an empty
  // move set that is not marked terminal is UNRESOLVED: either the player to move must pass or
  // the game is over (othello.cc:214-229,241-246).  With g = the moves of the other side after
  // this iteration's placement (or, without a placement, of the other side as it stands):
  //   moves != 0           place a disc; the other side is to move with g (g may be empty)
  //   moves == 0, g != 0   the pass: the other side is to move with g
  //   moves == 0, g == 0   nobody can move: the game ended with the previous placement; no
  //                        action is made (made = 0) and the state becomes terminal
  // Every path is the same flips + one move generation; the result equals play()'s.
  BSG_HD static u32 step_lazy(State& s, u32 r, const Tables& tab, int& action, u32& made) {
    const u64 m = s.moves;
    const u32 n = (u32)popc64(m);
    u32 k = fastmod(r, n, tab.modM[n & 63u]);
    u32 w = (u32)m;
    int pos = 0;
    const u32 c = (u32)popc32(w);
    if (k >= c) { k -= c; pos = 32; w = (u32)(m >> 32); }
    pos += select32(w, (int)k);
    const u64 x = n ? (1ull << (pos & 63)) : 0ull;
    const u64 flips = get_flips_fast(x, s.own, s.opp);
    const u64 mine = s.own | x | flips, theirs = s.opp & ~flips;
    const u64 g = get_moves(theirs, mine);
    const bool over = (n == 0) & (g == 0);
    if (over) {   // the last mover's discs are s.opp
      const int a = popc64(s.opp), b = popc64(s.own);
      s.term = 1u;
      s.res = a > b ? 1u : (a < b ? 2u : 0u);
    } else {
      s.own = theirs;
      s.opp = mine;
      s.mv ^= 1u;
    }
    s.moves = g;
    made = over ? 0u : 1u;
    action = n ? pos : kPass;
    return n ? n : 1u;
  }
  BSG_HD static u32 step(State& s, u32 r, const Tables& tab, int& action) {
    const u64 m = s.moves;
    const u32 n = (u32)popc64(m);
    int pos = kPass;
    if (n != 0) {
      u32 k = fastmod(r, n, tab.modM[n & 63u]);
      u32 w = (u32)m;
      pos = 0;
      const u32 c = (u32)popc32(w);
      if (k >= c) { k -= c; pos = 32; w = (u32)(m >> 32); }
      pos += select32(w, (int)k);
    }
    action = pos;
    play(s, pos);
    return n ? n : 1u;
  }
  BSG_HD static int legal(const State& s, int* out) {
    if (s.term) return 0;
    int n = 0;
    for (int c = 0; c < 64; ++c) if ((s.moves >> c) & 1ull) out[n++] = c;
    if (n == 0) out[n++] = kPass;
    return n;
  }
  BSG_HD static bool apply(State& s, int a) {
    if (s.term || a < 0 || a > kPass) return false;
    if (a == kPass) { if (s.moves != 0) return false; }   // (the reference lets any pass through)
    else if (!((s.moves >> a) & 1ull)) return false;
    play(s, a);
    return true;
  }
  BSG_HD static u32 legal_bit(const State& s, int a) {
    if (s.term) return 0u;
    return a == kPass ? (s.moves == 0 ? 1u : 0u) : (u32)((s.moves >> a) & 1ull);
  }
  // othello.cc:310-329: plane 0 empty, 1 the observer's discs, 2 the opponent's.
  BSG_HD static u32 obs_bit(const State& s, int player, int idx) {
    const int plane = idx >> 6, cell = idx & 63;
    const u64 mine = ((int)s.mv == player) ? s.own : s.opp, theirs = ((int)s.mv == player) ? s.opp : s.own;
    const u64 m = plane == 0 ? ~(mine | theirs) : (plane == 1 ? mine : theirs);
    return (u32)(m >> cell) & 1u;
  }
  BSG_HD static void obs_words(const State& s, int player, u32* out) {
    const u64 mine = ((int)s.mv == player) ? s.own : s.opp, theirs = ((int)s.mv == player) ? s.opp : s.own;
    const u64 e = ~(mine | theirs);
    out[0] = (u32)e; out[1] = (u32)(e >> 32);
    out[2] = (u32)mine; out[3] = (u32)(mine >> 32);
    out[4] = (u32)theirs; out[5] = (u32)(theirs >> 32);
  }
  BSG_HD static void mask_words(const State& s, u32* out) {
    const u64 m = s.term ? 0ull : s.moves;
    out[0] = (u32)m; out[1] = (u32)(m >> 32);
    out[2] = (!s.term && m == 0) ? 1u : 0u;
  }
};


// =========================================================================================
// pentago — reference: open_spiel/games/pentago.cc   (one more
//   perfect-information board game on the same kernel framework)
//   6 x 6 board, a move places a stone on an empty cell and then turns one 3 x 3 quadrant a
//   quarter turn: action = (y * 6 + x) * 8 + r, quadrant = r >> 1 numbered clockwise from the
//   top-left (0 TL, 1 TR, 2 BR, 3 BL), r & 1 = 0 counter-clockwise / 1 clockwise
//   (pentago.cc:52-75,78-85,132-141).  Five in a row wins — and because the turn comes after
//   the stone, a move can complete the OPPONENT's line (they win) or both (draw): the one game
//   here where the mover can lose by moving (kMoverCanLose; pentago.cc:290-308).
//   Bit = y * 8 + x in one u64 per colour (two guard columns keep shifts from wrapping), which
//   is also the reference's ascending action order, so the k-th legal action is the (k >> 3)-th
//   empty cell with rotation k & 7.  The quadrant turn is a 512-entry look-up on the nine
//   cells; the 32 win lines are four bit-parallel run-of-five tests.
//   packed: one 16-byte element = {player 0's stones | outcome code << 48, player 1's stones}.
// =========================================================================================
struct Pentago {
  static constexpr int kId = kGamePentago;
  static constexpr int kNumActions = 288, kMaxLen = 36, kMaxDepth = kMaxLen, kObsSize = 108, kStateBytes = 16;
  static constexpr int kObsWords = 4, kMaskWords = 9;
  static constexpr int kPlanes = 1;
  BSG_HD static int plane_bytes(int p) { (void)p; return 16; }
  static constexpr bool kCanStick = false;
  static constexpr bool kAsync = false;
  static constexpr bool kMoverCanLose = true;
  static constexpr bool kLazyEnd = false;       // the lean kernel sees a terminal state in the step that reaches it (see Othello)
  static constexpr u64 kBoard = 0x00003F3F3F3F3F3Full;
  struct State {
    u64 cur, opp;   // stones of the player to move / of the player who moved last
    u32 left;       // empty cells (36..0); parity = player to move (player 0 first, 36 is even)
    u32 res;        // bit 0: the last mover has five in a row; bit 1: the player to move has
  };
  // rot[d][m]: the nine cells m (bit 3 * row + col) of a quadrant after a quarter turn, d = 0
  // counter-clockwise / 1 clockwise, already spread to board rows (bits 0-2, 8-10, 16-18).
  // sel6[row << 3 | k]: column of the k-th empty cell of a board row (6-bit mask).
  struct Tables { u32 modM[64]; u32 rot[2][512]; unsigned char sel6[512]; };
  static void fill_tables(Tables& t) {
    for (u32 e = 0; e < 64; ++e) t.modM[e] = modM_entry(e * 8u);   // n = 8 * empty cells
    for (u32 row = 0; row < 64; ++row)
      for (u32 k = 0; k < 8; ++k) {
        u32 seen = 0, col = 0;
        for (u32 c = 0; c < 6; ++c)
          if ((row >> c) & 1u) { if (seen == k) col = c; ++seen; }
        t.sel6[(row << 3) | k] = (unsigned char)col;
      }
    for (u32 d = 0; d < 2; ++d)
      for (u32 m = 0; m < 512; ++m) {
        u32 out = 0;
        for (u32 r = 0; r < 3; ++r)
          for (u32 c = 0; c < 3; ++c)
            if ((m >> (3 * r + c)) & 1u) {
              // clockwise, y growing downwards: (row r, col c) -> (row c, col 2 - r)
              u32 nr = d ? c : 2 - c, nc = d ? 2 - r : r;
              out |= 1u << (8 * nr + nc);
            }
        t.rot[d][m] = out;
      }
  }
  BSG_HD static void init(State& s) { s.cur = 0; s.opp = 0; s.left = 36; s.res = 0; }
  BSG_HD static int mover(const State& s) { return (int)(s.left & 1u); }
  BSG_HD static bool terminal(const State& s) { return s.res != 0 || s.left == 0; }
  BSG_HD static bool won(const State& s) { return s.res == 1u; }    // the last mover won
  BSG_HD static bool lost(const State& s) { return s.res == 2u; }   // ... lost by moving
  BSG_HD static int outcome(const State& s) {
    if (s.res == 1u) return mover(s) ^ 1;
    if (s.res == 2u) return mover(s);
    return (s.res == 3u || s.left == 0) ? kOutcomeDraw : kOutcomeNone;
  }
  BSG_HD static void load(State& s, const void* base, int64_t n, int64_t i) {
    (void)n;
#if defined(__CUDA_ARCH__)
    ulonglong2 w = ((const ulonglong2*)base)[i];
    u64 w0 = w.x, p1 = w.y;
#else
    u64 w0 = ((const u64*)base)[2 * i], p1 = ((const u64*)base)[2 * i + 1];
#endif
    u64 p0 = w0 & kBoard;
    u32 outc = (u32)(w0 >> 48) & 3u;
    s.left = 36u - (u32)(popc64(p0) + popc64(p1));
    u32 mv = s.left & 1u;
    s.cur = mv ? p1 : p0;
    s.opp = mv ? p0 : p1;
    // outcome code -> who has a line, relative to the last mover (1 - mv)
    s.res = outc == (u32)kOutcomeNone ? 0u : (outc == (u32)kOutcomeDraw ? 3u : (outc == (mv ^ 1u) ? 1u : 2u));
  }
  BSG_HD static void store(const State& s, void* base, int64_t n, int64_t i) {
    (void)n;
    u32 mv = (u32)mover(s);
    u64 p0 = mv ? s.opp : s.cur, p1 = mv ? s.cur : s.opp;
    ((u64*)base)[2 * i] = p0 | ((u64)(u32)outcome(s) << 48);
    ((u64*)base)[2 * i + 1] = p1;
  }
  // pentago.cc:104-125: the 32 runs of five, as four bit-parallel directions, on the two
  // 32-bit words (rows 0-3 / rows 4-5).  Rows: inside each word, with left shifts (they can
  // run on the multiplier pipe).  Columns and diagonals: the top cell of such a run is in row
  // 0 or 1, so only the LOW word of the product matters — pairs, then pairs of pairs (both
  // inside rows 0-3 of the low word), then the fifth cell four rows further down, which is
  // the high word shifted into place.
  BSG_HD static bool has_five(u64 b) {
    const u32 lo = (u32)b, hi = (u32)(b >> 32);
    u32 tl = lo & (lo << 1), th = hi & (hi << 1);
    u32 acc = (tl & (tl << 2) & (lo << 4)) | (th & (th << 2) & (hi << 4));
    u32 t = lo & fshr(lo, hi, 8);                 // columns: rows r, r+1
    acc |= t & (t >> 16) & hi;                    //   r..r+3, then row r+4 = bit + 32
    t = lo & fshr(lo, hi, 9);                     // diagonals down-right
    acc |= t & (t >> 18) & (hi >> 4);             //   fifth cell at bit + 36
    t = lo & fshr(lo, hi, 7);                     // diagonals down-left
    acc |= t & (t >> 14) & fshr(lo, hi, 28);      //   fifth cell at bit + 28
    return acc != 0;
  }
  // Bit offset of the quadrant's top-left cell: TL 0, TR 3, BR 27, BL 24.
  BSG_HD static u32 quadrant_shift(u32 q) { return (0x181B0300u >> (8u * q)) & 0xFFu; }
  // Quarter turn: the nine cells of the quadrant through the look-up.  (A word-level variant
  // — funnel the half to one word, bit-select it back — compiled to more instructions.)
  BSG_HD static u64 turn(u64 b, u32 sh, u32 dir, const Tables& tab) {
    u32 w = (u32)(b >> sh);
    u32 m = (w & 7u) | ((w >> 5) & 0x38u) | ((w >> 10) & 0x1C0u);
    return (b & ~(0x070707ull << sh)) | ((u64)tab.rot[dir][m] << sh);
  }
  // pentago.cc:274-312 from the placed stone on: turn both colours, test both players' lines.
  BSG_HD static void play(State& s, int pos, u32 r, const Tables& tab) {
    const u32 sh = quadrant_shift(r >> 1), dir = r & 1u;
    u64 mine = turn(s.cur | (1ull << pos), sh, dir, tab);
    u64 theirs = turn(s.opp, sh, dir, tab);
    s.res = (has_five(mine) ? 1u : 0u) | (has_five(theirs) ? 2u : 0u);
    s.cur = theirs;
    s.opp = mine;
    s.left -= 1u;
  }
  BSG_HD static int cell_of(int pos) { return (pos >> 3) * 6 + (pos & 7); }
  BSG_HD static int pos_of(int cell) { return (cell / 6) * 8 + cell % 6; }
  BSG_HD static u32 step(State& s, u32 r, const Tables& tab, int& action) {
    const u64 e = ~(s.cur | s.opp) & kBoard;               // pentago.cc:147-160
    const u32 n = s.left * 8u;
    const u32 k = fastmod(r, n, tab.modM[s.left & 63u]);   // dead lanes: left wraps, index stays in range
    u32 j = k >> 3;                                        // the j-th empty cell, ascending
    u32 w = (u32)e;
    int pos = 0;
    u32 c = (u32)popc32(w);
    if (j >= c) { j -= c; pos = 32; w = (u32)(e >> 32); }
    c = (u32)popc32(w << 16); if (j >= c) { j -= c; pos += 16; w >>= 16; }   // two rows
    c = (u32)popc32(w << 24); if (j >= c) { j -= c; pos += 8; w >>= 8; }     // the row
    pos += (int)tab.sel6[((w << 3) & 0x1F8u) | (j & 7u)];                    // the column
    pos &= 63;
    action = cell_of(pos) * 8 + (int)(k & 7u);
    play(s, pos, k & 7u, tab);
    return n;
  }
  BSG_HD static int legal(const State& s, int* out) {
    if (terminal(s)) return 0;
    int n = 0;
    for (int cell = 0; cell < 36; ++cell)
      if (!(((s.cur | s.opp) >> pos_of(cell)) & 1ull))
        for (int r = 0; r < 8; ++r) out[n++] = cell * 8 + r;
    return n;
  }
  BSG_HD static bool apply(State& s, int a, const Tables& tab) {
    if (terminal(s) || a < 0 || a >= kNumActions) return false;
    const int pos = pos_of(a >> 3);
    if (((s.cur | s.opp) >> pos) & 1ull) return false;
    play(s, pos, (u32)a & 7u, tab);
    return true;
  }
  BSG_HD static u32 legal_bit(const State& s, int a) {
    return (!terminal(s) && !(((s.cur | s.opp) >> pos_of(a >> 3)) & 1ull)) ? 1u : 0u;
  }
  // Rows of 8 bits -> rows of 6 bits (cell order of the observation planes).
  BSG_HD static u64 compact(u64 b) {
    u64 out = 0;
#pragma unroll
    for (int y = 0; y < 6; ++y) out |= ((b >> (8 * y)) & 0x3Full) << (6 * y);
    return out;
  }
  // pentago.cc:246-272: plane 0 the observer's stones, 1 the other player's, 2 empty.
  BSG_HD static u32 obs_bit(const State& s, int player, int idx) {
    int plane = idx / 36, pos = pos_of(idx - plane * 36);
    u64 own = (mover(s) == player) ? s.cur : s.opp, other = (mover(s) == player) ? s.opp : s.cur;
    u64 m = plane == 0 ? own : (plane == 1 ? other : ~(own | other));
    return (u32)(m >> pos) & 1u;
  }
  BSG_HD static void obs_words(const State& s, int player, u32* out) {
    u64 own = (mover(s) == player) ? s.cur : s.opp, other = (mover(s) == player) ? s.opp : s.cur;
    u64 a = compact(own), b = compact(other), e = compact(~(own | other) & kBoard);
    out[0] = (u32)a;                                         // bits 0..35 own, 36..71 other, 72..107 empty
    out[1] = (u32)(a >> 32) | (u32)(b << 4);
    out[2] = (u32)(b >> 28) | (u32)(e << 8);
    out[3] = (u32)(e >> 24);
  }
  // eight action bits per empty cell: byte `cell` of the 288-bit string
  BSG_HD static void mask_words(const State& s, u32* out) {
    u64 e = terminal(s) ? 0ull : compact(~(s.cur | s.opp) & kBoard);
#pragma unroll
    for (int w = 0; w < kMaskWords; ++w) {
      u32 nib = (u32)(e >> (4 * w)) & 0xFu;
      out[w] = ((nib & 1u) ? 0xFFu : 0u) | ((nib & 2u) ? 0xFF00u : 0u) |
               ((nib & 4u) ? 0xFF0000u : 0u) | ((nib & 8u) ? 0xFF000000u : 0u);
    }
  }
};


// =========================================================================================
// y N (N = 19, the reference's default, in the product; N = 9 only for the reference's golden
// playthrough) — reference: open_spiel/games/y.cc   
//   The board is the triangle x + y < N of an N x N grid, cell = x + y * N (y.h:56-63), so
//   NumDistinctActions = N * N with the cut-off corner never legal (y.h:174-178).  Player 0
//   ("kPlayer1", drawn O) moves first.  A player wins when ONE group of their stones touches
//   all three edges x = 0, y = 0, x + y = N - 1 (y.cc:106-111,285-287); corners touch two.  The
//   board cannot fill up without a winner, so there is no draw.  The six neighbours
//   (y.cc:67-71) are the offsets -N, -N+1, +1, +N, +N-1, -1: the hex neighbourhood, so one
//   ring of a set is the same four shifts as in HexT, on W = ceil((N*(N-1)+1)/32) words.
//
//   The reference keeps groups in a union-find with an edge bit-set per leader
//   (y.cc:293-325).  Here every stone carries its group's edge set as three bit planes
//   E0, E1, E2 over the grid (both colours in one plane: a cell has one owner): a new stone
//   takes the union L of its own edges and its neighbours' groups' edge sets, wins when L is
//   all three, and otherwise L floods through the mover's stones that lack a label of L — the
//   union of the merged groups.  The flood is the asynchronous one of HexT (kAsync): placing
//   a stone records the first frontier, every kernel iteration runs one ring for every lane,
//   a lane moves only when its frontier is empty; host helpers flood synchronously.
//   Stones are kept as A (all stones) and P (player 1's), so the mover's stones are one LOP3
//   and nothing is swapped when the turn passes.
//   packed: SoA, 5 planes of 4*PW bytes (PW = W rounded up to 4 words): player 0's stones,
//   player 1's stones, E0, E1, E2; plane 5: u32 = last move + 1 (0: none; y.cc:160-207 draws
//   it in brackets).
// =========================================================================================
template <int N>
struct YT {
  static constexpr int kId = kGameY;
  static constexpr bool kRolledGroup = BSG_Y_ROLLED;   // playout kernel: step group not unrolled (A/B)
  static constexpr int kGrid = N * N;
  static constexpr int kCells = N * (N + 1) / 2;
  static constexpr int kBits = N * (N - 1) + 1;          // highest on-board cell index + 1
  static constexpr int W = (kBits + 31) / 32;            // words that can hold a stone
  static constexpr int PW = (W + 3) / 4 * 4;             // words per packed plane (16-byte multiples)
  static_assert(PW <= 12 && N <= 19, "mask tables below are written out for at most 12 words");
  static constexpr int kNumActions = kGrid, kMaxLen = kCells, kMaxDepth = kMaxLen, kObsSize = 3 * kGrid;
  static constexpr int kStateBytes = 5 * 4 * PW + 4;
  static constexpr int kObsWords = (3 * kGrid + 31) / 32, kMaskWords = (kGrid + 31) / 32;
  static constexpr int kPlanes = 6;
  BSG_HD static int plane_bytes(int p) { return p < 5 ? 4 * PW : 4; }
  static constexpr bool kCanStick = false;
  static constexpr bool kAsync = true;
  static constexpr int kMinBlocks = 2;        // lean kernel: register cap (128) for two CTAs per SM
  static constexpr bool kMoverCanLose = false;
  static constexpr bool kLazyEnd = false;
  struct State {
    u32 A[W], P[W];            // all stones / player 1's stones
    u32 E0[W], E1[W], E2[W];   // stones whose group touches edge x = 0 / y = 0 / x + y = N - 1
    u32 F[W];                  // pending flood frontier: stones of the flood's owner next to
                               // newly labelled ones that still lack a label being spread
    u32 m0, m1, m2;            // ~0: the pending flood spreads that label
    u32 fq;                    // ~0: the flood's owner is player 1
    u32 left;                  // empty cells
    u32 mv;                    // player to move
    u32 win;                   // the last move connected all three edges
    u32 last;                  // last move + 1 (0: none)
  };
  // nb[cell]: neighbour mask (words 0..W-1).  edge[cell]: the cell's own edge set (bit 0 x = 0,
  // bit 1 y = 0, bit 2 x + y = N - 1).  sel8[b][k]: position of the k-th set bit of byte b.
  struct Tables {
    u32 nb[kGrid][PW];
    u32 modM[256];
    unsigned char edge[(kGrid + 3) / 4 * 4];
    unsigned char sel8[256][8];
  };
  BSG_HDC static bool on_board(int x, int y) { return x >= 0 && y >= 0 && x < N && y < N && x + y < N; }
  static void fill_tables(Tables& t) {
    for (u32 n = 0; n < 256; ++n) t.modM[n] = modM_entry(n);
    for (int b = 0; b < 256; ++b) {
      int k = 0;
      for (int i = 0; i < 8; ++i) t.sel8[b][i] = 0;
      for (int i = 0; i < 8; ++i) if ((b >> i) & 1) t.sel8[b][k++] = (unsigned char)i;
    }
    for (int i = 0; i < (int)sizeof(t.edge); ++i) t.edge[i] = 0;
    const int dx[6] = {0, 1, 1, 0, -1, -1}, dy[6] = {-1, -1, 0, 1, 1, 0};   // y.cc:67-71
    for (int cell = 0; cell < kGrid; ++cell) {
      for (int w = 0; w < PW; ++w) t.nb[cell][w] = 0;
      const int x = cell % N, y = cell / N;
      if (!on_board(x, y)) continue;
      for (int k = 0; k < 6; ++k) {
        const int x2 = x + dx[k], y2 = y + dy[k];
        if (!on_board(x2, y2)) continue;
        const int v = x2 + y2 * N;
        t.nb[cell][v >> 5] |= 1u << (v & 31);
      }
      t.edge[cell] = (unsigned char)((x == 0 ? 1 : 0) | (y == 0 ? 2 : 0) | (x + y == N - 1 ? 4 : 0));
    }
  }
  // Compile-time masks of word w: board cells (kind 0), board cells on edge 0 / 1 / 2 (kinds 3 / 4 / 5).
  BSG_HDC static u32 word_mask(int w, int kind) {
    u32 m = 0;
    for (int b = 0; b < 32; ++b) {
      const int cell = 32 * w + b;
      if (cell >= kGrid) break;
      const int x = cell % N, y = cell / N;
      const bool in = on_board(x, y) &&
                      (kind == 0 || (kind == 3 ? x == 0 : (kind == 4 ? y == 0 : x + y == N - 1)));
      if (in) m |= 1u << b;
    }
    return m;
  }
#define BSG_Y_MASKS(kind)                                                                        \
  {word_mask(0, kind), word_mask(1, kind), word_mask(2, kind),  word_mask(3, kind),              \
   word_mask(4, kind), word_mask(5, kind), word_mask(6, kind),  word_mask(7, kind),              \
   word_mask(8, kind), word_mask(9, kind), word_mask(10, kind), word_mask(11, kind)}
  BSG_HD static u32 board_word(int w) { constexpr u32 m[12] = BSG_Y_MASKS(0); return m[w]; }
  BSG_HD static u32 edge_word(int e, int w) {   // board cells on edge e (0: x = 0, 1: y = 0, 2: x + y = N - 1)
    constexpr u32 m0[12] = BSG_Y_MASKS(3);
    constexpr u32 m1[12] = BSG_Y_MASKS(4);
    constexpr u32 m2[12] = BSG_Y_MASKS(5);
    return e == 0 ? m0[w] : (e == 1 ? m1[w] : m2[w]);
  }
#undef BSG_Y_MASKS

  BSG_HD static void init(State& s) {
#pragma unroll
    for (int w = 0; w < W; ++w) s.A[w] = s.P[w] = s.E0[w] = s.E1[w] = s.E2[w] = s.F[w] = 0;
    s.m0 = s.m1 = s.m2 = 0; s.fq = 0;
    s.left = kCells; s.mv = 0; s.win = 0; s.last = 0;
  }
  BSG_HD static void load(State& s, const void* base, int64_t n, int64_t i) {
    u32 pl[5][PW];
#if defined(__CUDA_ARCH__)
#pragma unroll
    for (int p = 0; p < 5; ++p) {
      const uint4* q = (const uint4*)((const char*)base + (int64_t)p * 4 * PW * n) + i * (PW / 4);
#pragma unroll
      for (int j = 0; j < PW / 4; ++j) {
        const uint4 v = q[j];
        pl[p][4 * j] = v.x; pl[p][4 * j + 1] = v.y; pl[p][4 * j + 2] = v.z; pl[p][4 * j + 3] = v.w;
      }
    }
#else
    for (int p = 0; p < 5; ++p) {
      const u32* q = (const u32*)((const char*)base + (int64_t)p * 4 * PW * n) + i * PW;
      for (int j = 0; j < PW; ++j) pl[p][j] = q[j];
    }
#endif
    s.last = ((const u32*)((const char*)base + (int64_t)5 * 4 * PW * n))[i];
    int n0 = 0, n1 = 0;
    u32 won = 0;
#pragma unroll
    for (int w = 0; w < W; ++w) {
      n0 += popc32(pl[0][w]); n1 += popc32(pl[1][w]);
      s.A[w] = pl[0][w] | pl[1][w]; s.P[w] = pl[1][w];
      s.E0[w] = pl[2][w]; s.E1[w] = pl[3][w]; s.E2[w] = pl[4][w];
      s.F[w] = 0;
      won |= pl[2][w] & pl[3][w] & pl[4][w];
    }
    s.mv = n0 > n1 ? 1u : 0u;   // player 0 moves first and turns alternate
    s.win = won ? 1u : 0u;
    s.left = (u32)(kCells - n0 - n1);
    s.m0 = s.m1 = s.m2 = 0; s.fq = 0;
  }
  BSG_HD static void store(const State& s0, void* base, int64_t n, int64_t i) {
    State s = s0;
    finish(s);   // a packed state always carries complete edge sets
#pragma unroll
    for (int p = 0; p < 5; ++p) {
      u32* q = (u32*)((char*)base + (int64_t)p * 4 * PW * n) + i * PW;
#pragma unroll
      for (int w = 0; w < PW; ++w) {
        u32 v = 0;
        if (w < W) v = p == 0 ? (s.A[w] & ~s.P[w]) : (p == 1 ? s.P[w] : (p == 2 ? s.E0[w] : (p == 3 ? s.E1[w] : s.E2[w])));
        q[w] = v;
      }
    }
    ((u32*)((char*)base + (int64_t)5 * 4 * PW * n))[i] = s.last;
  }
  BSG_HD static bool terminal(const State& s) { return s.win != 0; }
  BSG_HD static bool won(const State& s) { return s.win != 0; }
  BSG_HD static int mover(const State& s) { return (int)s.mv; }
  BSG_HD static int outcome(const State& s) { return s.win ? (int)(s.mv ^ 1u) : kOutcomeNone; }

  // One ring of neighbours of every cell of f, restricted to `within` (which lies on the
  // board).  x+1 / x-1 are shifts by one with the row wrap masked; (x, y-1) and (x+1, y-1) are
  // (f | east) shifted down by N, (x, y+1) and (x-1, y+1) are (f | west) shifted up by N.
  BSG_HD static void dilate_within(const u32 f[W], const u32 within[W], u32 out[W]) {
    u32 e[W], x[W], fe[W], fw[W], nn[W], ss[W];
    shl_words<1, W>(f, e);
    shr_words<1, W>(f, x);
    // f lies on the board, so the only shifts by one that wrap onto a board cell are
    // (N-1, 0) -> (0, 1) and back: bits N and N - 1 of word 0; every other stray bit lands
    // off the board and is dropped by `within`
    e[0] &= ~(1u << N);
    x[0] &= ~(1u << (N - 1));
#pragma unroll
    for (int w = 0; w < W; ++w) {
      fe[w] = f[w] | e[w];
      fw[w] = f[w] | x[w];
    }
    static_assert(N < 32, "row shifts are written for N < 32");
    shr_words<N, W>(fe, nn);
    shl_words<N, W>(fw, ss);
#pragma unroll
    for (int w = 0; w < W; ++w) out[w] = (fe[w] | fw[w] | nn[w] | ss[w]) & within[w];
  }
  // The pending flood, one ring per call: the frontier takes the labels being spread, the next
  // frontier is its ring among the owner's stones that still lack one of them.
  BSG_HD static void ring(State& s) {
    u32 within[W], d[W];
#pragma unroll
    for (int w = 0; w < W; ++w) {
      s.E0[w] |= s.F[w] & s.m0;
      s.E1[w] |= s.F[w] & s.m1;
      s.E2[w] |= s.F[w] & s.m2;
      const u32 own = s.A[w] & ~(s.P[w] ^ s.fq);
      const u32 full = (s.E0[w] | ~s.m0) & (s.E1[w] | ~s.m1) & (s.E2[w] | ~s.m2);
      within[w] = own & ~full;
    }
    dilate_within(s.F, within, d);
#pragma unroll
    for (int w = 0; w < W; ++w) s.F[w] = d[w];
  }
  BSG_HD static bool pending(const State& s) {
    u32 any = 0;
#pragma unroll
    for (int w = 0; w < W; ++w) any |= s.F[w];
    return any != 0;
  }
  BSG_HD static void finish(State& s) { while (pending(s)) ring(s); }

  // y.cc:269-291 for the mover.  The new stone's group has the edges of the stone itself and
  // of every neighbouring group of the mover (y.cc:279-283: JoinGroups ORs the edge sets).
  BSG_HD static void place(State& s, int cell, const Tables& t) {
    const u32 q = s.mv ? 0xFFFFFFFFu : 0u;
    const u32 edges = t.edge[cell];
    const u32 cw = (u32)cell >> 5, cb = 1u << (cell & 31);
    u32 nb[W], mine[W];
    u32 h0 = 0, h1 = 0, h2 = 0;
#pragma unroll
    for (int w = 0; w < W; ++w) {
      nb[w] = t.nb[cell][w];
      mine[w] = s.A[w] & ~(s.P[w] ^ q);
      const u32 x = nb[w] & mine[w];
      h0 |= x & s.E0[w]; h1 |= x & s.E1[w]; h2 |= x & s.E2[w];
    }
    const bool t0 = h0 != 0 || (edges & 1u), t1 = h1 != 0 || (edges & 2u), t2 = h2 != 0 || (edges & 4u);
    const u32 l0 = t0 ? 0xFFFFFFFFu : 0u, l1 = t1 ? 0xFFFFFFFFu : 0u, l2 = t2 ? 0xFFFFFFFFu : 0u;
    const bool win = t0 && t1 && t2;
    const u32 go = win ? 0u : 0xFFFFFFFFu;   // the kernel stops at a win; store() completes the labels
#pragma unroll
    for (int w = 0; w < W; ++w) {
      const u32 m = (u32)w == cw ? cb : 0u;
      s.A[w] |= m;
      s.P[w] |= m & q;
      // the new stone rides in the frontier: the next ring (run before any lane moves again,
      // and by store()) gives it the labels L, together with its neighbours that lack one
      const u32 full = (s.E0[w] | ~l0) & (s.E1[w] | ~l1) & (s.E2[w] | ~l2);
      s.F[w] = (nb[w] & mine[w] & ~full & go) | m;
    }
    s.m0 = l0; s.m1 = l1; s.m2 = l2; s.fq = q;
    s.win = win ? 1u : 0u;
    s.mv ^= 1u;
    s.left -= 1u;
    s.last = (u32)cell + 1u;
  }
  BSG_HD static u32 empty_word(const State& s, int w) { return ~s.A[w] & board_word(w); }
  // k-th (ascending) empty cell, k < left: the word by running popcounts, then halves of the
  // word, then a byte table.
  BSG_HD static int select_empty(const State& s, u32 k, const Tables& tab) {
    u32 x = empty_word(s, 0), below = 0, pos = 0, cum = 0;
#pragma unroll
    for (int w = 0; w + 1 < W; ++w) {
      cum += (u32)popc32(empty_word(s, w));
      const bool g = k >= cum;
      x = g ? empty_word(s, w + 1) : x;
      below = g ? cum : below;
      pos = g ? 32u * (u32)(w + 1) : pos;
    }
    u32 kk = k - below;
    u32 h = (u32)popc32(x & 0xFFFFu);
    if (kk >= h) { kk -= h; x >>= 16; pos += 16; }
    h = (u32)popc32(x & 0xFFu);
    if (kk >= h) { kk -= h; x >>= 8; pos += 8; }
    const u32 cell = pos + tab.sel8[x & 0xFFu][kk & 7u];
    return (int)(cell < (u32)kGrid ? cell : 0u);   // a dead lane's garbage stays inside the tables
  }
  BSG_HD static u32 step(State& s, u32 r, const Tables& tab, int& action) {
    const u32 n = s.left;                                  // y.cc:138-149: every empty cell
    const u32 k = fastmod(r, n, tab.modM[n & 255u]);
    const int cell = select_empty(s, k, tab);
    action = cell;
    place(s, cell, tab);
    return n;
  }
  BSG_HD static int legal(const State& s, int* out) {
    if (terminal(s)) return 0;
    int n = 0;
    for (int c = 0; c < kBits; ++c) if ((empty_word(s, c >> 5) >> (c & 31)) & 1u) out[n++] = c;
    return n;
  }
  BSG_HD static u32 legal_bit(const State& s, int a) {
    return (!s.win && a >= 0 && a < kBits && ((empty_word(s, a >> 5) >> (a & 31)) & 1u)) ? 1u : 0u;
  }
  BSG_HD static bool apply(State& s, int a, const Tables& t) {
    if (!legal_bit(s, a)) return false;
    place(s, a, t);
    finish(s);      // host helpers flood synchronously
    return true;
  }
  // y.cc:239-267: plane 0 the observer's stones, 1 the other player's, 2 empty board cells;
  // grid cells off the board are zero in every plane.
  BSG_HD static u32 plane_word(const State& s, int player, int plane, int w) {
    const u32 q = player ? 0xFFFFFFFFu : 0u;
    if (plane == 0) return s.A[w] & ~(s.P[w] ^ q);
    if (plane == 1) return s.A[w] & (s.P[w] ^ q);
    return empty_word(s, w);
  }
  BSG_HD static u32 obs_bit(const State& s, int player, int idx) {
    const int plane = idx / kGrid, cell = idx - plane * kGrid;
    if (cell >= kBits) return 0u;
    return (plane_word(s, player, plane, cell >> 5) >> (cell & 31)) & 1u;
  }
  BSG_HD static void obs_words(const State& s, int player, u32* out) {
#pragma unroll
    for (int w = 0; w < kObsWords; ++w) out[w] = 0;
#pragma unroll
    for (int p = 0; p < 3; ++p) {
      const int off = p * kGrid, wi = off >> 5, sh = off & 31;
#pragma unroll
      for (int j = 0; j < W; ++j) {
        const u32 m = plane_word(s, player, p, j);
        if (wi + j < kObsWords) out[wi + j] |= m << sh;
        if (sh != 0 && wi + j + 1 < kObsWords) out[wi + j + 1] |= m >> ((32 - sh) & 31);
      }
    }
  }
  BSG_HD static void mask_words(const State& s, u32* out) {
#pragma unroll
    for (int w = 0; w < kMaskWords; ++w) out[w] = (w < W && !s.win) ? empty_word(s, w < W ? w : 0) : 0u;
  }

  // ---- value-only playouts (see HexT): a full y board has exactly one winner and stones
  // never leave the board, so the first player to join the three edges is the one whose
  // stones join them on the full board.
  static constexpr bool kFillable = true;
  static constexpr int kFillWords = PW;
  BSG_HD static u32 step_fill(State& s, u32 r, const Tables& tab, int& action) {
    const u32 n = s.left;
    const u32 k = fastmod(r, n, tab.modM[n & 255u]);
    const int cell = select_empty(s, k, tab);
    action = cell;
    const u32 q = s.mv ? 0xFFFFFFFFu : 0u;
    // (n = 0: a dead step of a lane whose board is full must not touch the board)
    const u32 cw = (u32)cell >> 5, cb = n ? 1u << (cell & 31) : 0u;
#pragma unroll
    for (int w = 0; w < W; ++w) {
      const u32 m = (u32)w == cw ? cb : 0u;
      s.A[w] |= m;
      s.P[w] |= m & q;
    }
    s.mv ^= 1u;
    s.left -= n ? 1u : 0u;
    return n;
  }
  BSG_HD static bool fill_done(const State& s) { return s.left == 0 || s.win != 0; }
  BSG_HD static void fill_store(const State& s, u32* out) {   // player 0's stones
#pragma unroll
    for (int w = 0; w < PW; ++w) out[w] = w < W ? (s.A[w] & ~s.P[w]) : 0u;
  }
  // stones of `set` connected to `seed` (seed within set)
  BSG_HD static void flood(const u32* set, u32* seed_reach) {
    u32 f[W], rest[W], d[W];
    u32 any = 0;
#pragma unroll
    for (int w = 0; w < W; ++w) { f[w] = seed_reach[w]; any |= f[w]; }
    while (any) {
#pragma unroll
      for (int w = 0; w < W; ++w) rest[w] = set[w] & ~seed_reach[w];
      dilate_within(f, rest, d);
      any = 0;
#pragma unroll
      for (int w = 0; w < W; ++w) { f[w] = d[w]; seed_reach[w] |= d[w]; any |= d[w]; }
    }
  }
  // 0: a group of player 0 touches all three edges; 1: none does (player 1 wins a full or
  // finished board).  Groups touching edge 0, of those the groups touching edge 1, of those
  // any stone on edge 2.
  BSG_HD static int fill_winner(const u32* p0) {
    u32 r0[W], r01[W];
#pragma unroll
    for (int w = 0; w < W; ++w) r0[w] = p0[w] & edge_word(0, w);
    flood(p0, r0);
#pragma unroll
    for (int w = 0; w < W; ++w) r01[w] = r0[w] & edge_word(1, w);
    flood(r0, r01);
    u32 hit = 0;
#pragma unroll
    for (int w = 0; w < W; ++w) hit |= r01[w] & edge_word(2, w);
    return hit ? 0 : 1;
  }
};
typedef YT<19> Y;


}  // namespace bsg

#include "bsg_havannah.cuh"
#include "bsg_go.cuh"
