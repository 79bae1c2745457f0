// havannah — reference: open_spiel/games/havannah.{h,cc}   
//
//   Hexagonal board of radius S (the reference's default board_size 8: a 15 x 15 grid with two
//   corners cut off, 169 cells; action id = x + y * (2S-1), havannah.h:60-69, 221-225).  Players
//   alternate placing stones (no swap); the mover wins when the group of the new stone touches
//   three EDGES (corners excluded) or two CORNERS, or closes a RING (havannah.cc:314-349); a full
//   board is a draw.
//
//   What has to be reproduced bit for bit is the reference's ring rule, which is defined by its
//   own search, not by topology: a ring is looked for only when one of the joins made for the
//   new stone found the neighbour ALREADY in the stone's group (havannah.cc:326-338 — neighbours
//   are visited clockwise and the direction after a joined one is skipped), and the search is a
//   depth-first walk that never turns sharper than 60 degrees and succeeds when it re-enters a
//   cell of its own path (CheckRingDFS, havannah.cc:385-399, started in directions 0..3).
//   Group identities are therefore kept the way the reference keeps them, as a union-find over
//   the cells — one byte per cell, in the lane's local memory — with the corner / edge sets at
//   the leaders; which cell leads a group has no influence on any result, so a stone that joins
//   a group simply points at that group's leader (flat trees, short finds).  The ring search runs
//   as an explicit-stack loop over the same (cell, first direction, last direction) frames as
//   the reference's recursion — and only when the mover's stones CAN hold a ring: a solid hexagon,
//   or a hole, counted by Euler characteristic from popcounts (ring_possible).
//   Stones are also kept as bit sets (W words per colour) for the empty-cell rank-select, the
//   observation planes and the packed state.
//
//   This is a different kernel shape from the flood / line-test games: ~1.2 KB of per-lane state
//   in local memory, data-dependent loops in the step.  It runs through the same templated
//   kernels; finished lanes leave step() at once instead of running dead steps.
//   packed: SoA, plane 0 / 1: player 0's / player 1's stones (PW words each), plane 2: u32 =
//   mover | outcome << 1 | (last move + 1) << 8.
#pragma once
#include "bsg_common.cuh"
#ifndef BSG_HV_MINBLOCKS
#define BSG_HV_MINBLOCKS 4
#endif
#ifndef BSG_HV_ROLLED
#define BSG_HV_ROLLED 1
#endif

namespace bsg {

template <int S>
struct HavannahT {
  static constexpr int kId = kGameHavannah;
  static constexpr bool kLocalHeavy = true;   // playout kernel: L1 over shared memory (the state is local memory)
  static constexpr int kMinBlocks = BSG_HV_MINBLOCKS;   // lean kernel: resident blocks per SM asked of the compiler
  static constexpr bool kRolledGroup = BSG_HV_ROLLED;   // playout kernel: step group not unrolled (A/B)
  static constexpr int D = 2 * S - 1;
  static constexpr int kGrid = D * D;
  static constexpr int kCells = D * D - S * (S - 1);     // havannah.cc:166-168
  static constexpr int W = (kGrid + 31) / 32;
  static constexpr int PW = (W + 3) / 4 * 4;             // words per packed plane (16-byte multiples)
  static_assert(kGrid <= 255, "cell ids are bytes");
  static constexpr int kNumActions = kGrid, kMaxLen = kCells, kMaxDepth = kCells, kObsSize = 3 * kGrid;
  static constexpr int kStateBytes = 2 * 4 * PW + 4;
  static constexpr int kObsWords = (3 * kGrid + 31) / 32, kMaskWords = W;
  static constexpr int kPlanes = 3;
  BSG_HD static int plane_bytes(int p) { return p < 2 ? 4 * PW : 4; }
  static constexpr bool kCanStick = false;
  static constexpr bool kAsync = false;
  static constexpr bool kMoverCanLose = false;
  static constexpr bool kLazyEnd = false;
  struct State {
    u32 st[2][W];                    // stones of player 0 / player 1
    __align__(4) unsigned char parent[(kGrid + 3) / 4 * 4];   // union-find over the cells (havannah.h:96-131)
    __align__(4) unsigned short ec[(kGrid + 1) / 2 * 2];      // at a leader: corners (bits 0..5) | edges (bits 6..11) of its group
    u32 left;                        // empty cells
    u32 comps;                       // groups of player 0 | groups of player 1 << 16 (for ring_possible)
    u32 mv;                          // player to move
    u32 outc;                        // kOutcomeNone / kOutcomeP0 / kOutcomeP1 / kOutcomeDraw
    u32 last;                        // last move + 1 (0: none)
  };
  // sel8[b][k]: position of the k-th set bit of byte b.  nb[cell][d]: neighbour d of the cell
  // (havannah.cc:61-75), kOff when it is off the board.
  static constexpr int kOff = 255;
  struct Tables {
    u32 modM[256];
    unsigned char sel8[256][8];
    unsigned char nb[(kGrid + 3) / 4 * 4][8];
  };
  BSG_HDC static bool on_board(int x, int y) {           // havannah.h:60-69 CalcXY
    return x >= 0 && y >= 0 && x < D && y < D && (y - x < S) && (x - y < S);
  }
  BSG_HDC static int corner_of(int x, int y) {           // havannah.cc:112-126
    const int m = S - 1, e = 2 * (S - 1);
    if (x == 0 && y == 0) return 1 << 0;
    if (x == m && y == 0) return 1 << 1;
    if (x == e && y == m) return 1 << 2;
    if (x == e && y == e) return 1 << 3;
    if (x == m && y == e) return 1 << 4;
    if (x == 0 && y == m) return 1 << 5;
    return 0;
  }
  BSG_HDC static int edge_of(int x, int y) {             // havannah.cc:128-143
    const int m = S - 1, e = 2 * (S - 1);
    if (y == 0 && x != 0 && x != m) return 1 << 0;
    if (x - y == m && x != m && x != e) return 1 << 1;
    if (x == e && y != m && y != e) return 1 << 2;
    if (y == e && x != e && x != m) return 1 << 3;
    if (y - x == m && x != m && x != 0) return 1 << 4;
    if (x == 0 && y != m && y != 0) return 1 << 5;
    return 0;
  }
  static void fill_tables(Tables& t) {
    for (u32 n = 0; n < 256; ++n) t.modM[n] = modM_entry(n);
    for (int b = 0; b < 256; ++b) {
      int k = 0;
      for (int i = 0; i < 8; ++i) t.sel8[b][i] = 0;
      for (int i = 0; i < 8; ++i) if ((b >> i) & 1) t.sel8[b][k++] = (unsigned char)i;
    }
    for (int c = 0; c < (int)(sizeof(t.nb) / 8); ++c)
      for (int d = 0; d < 8; ++d) {
        const int m = (c < kGrid && d < 6 && on_board(c % D, c / D)) ? neighbour(c, d) : -1;
        t.nb[c][d] = (unsigned char)(m < 0 ? kOff : m);
      }
  }
  // Board cells of word w (compile time).
  BSG_HDC static u32 word_mask(int w) {
    u32 m = 0;
    for (int b = 0; b < 32; ++b) {
      const int cell = 32 * w + b;
      if (cell >= kGrid) break;
      if (on_board(cell % D, cell / D)) m |= 1u << b;
    }
    return m;
  }
  BSG_HD static u32 board_word(int w) {
    constexpr u32 m[8] = {word_mask(0), word_mask(1), word_mask(2), word_mask(3),
                          word_mask(4), word_mask(5), word_mask(6), word_mask(7)};
    static_assert(W <= 8, "mask table is written out for at most 8 words");
    return m[w];
  }
  BSG_HD static bool has(const u32* set, int cell) { return (set[cell >> 5] >> (cell & 31)) & 1u; }

  // Direct neighbours, clockwise (havannah.cc:61-75): direction d of (x, y), or -1 off the board.
  BSG_HD static int neighbour(int cell, int d) {
    const int x = cell % D, y = cell / D;
    const int dx = (d == 2 || d == 3) ? 1 : ((d == 0 || d == 5) ? -1 : 0);
    const int dy = (d == 0 || d == 1) ? -1 : ((d == 3 || d == 4) ? 1 : 0);
    const int x2 = x + dx, y2 = y + dy;
    return on_board(x2, y2) ? x2 + y2 * D : -1;
  }

  // havannah.cc:351-362 (path halving instead of compressing one cell: same groups)
  BSG_HD static int find(State& s, int cell) {
    int c = cell;
    while (s.parent[c] != c) {
      const int p = s.parent[c];
      s.parent[c] = s.parent[p];
      c = s.parent[c];
    }
    return c;
  }
  // havannah.cc:364-383; true when the two cells were in one group already
  // (a's group goes under b's leader: a cell that joins an existing group points straight at
  // its leader, so the trees stay flat and find() is one or two hops)
  BSG_HD static bool join(State& s, int a, int b) {
    const int la = find(s, a), lb = find(s, b);
    if (la == lb) return true;
    s.parent[la] = (unsigned char)lb;
    s.ec[lb] = (unsigned short)(s.ec[lb] | s.ec[la]);
    return false;
  }

  // Corner and edge bits of every cell, two cells to a word (compile time; read through the
  // constant bank).  A group starts as one cell (havannah.cc:170-176).
  struct Ec0 {
    u32 v[(kGrid + 1) / 2];
    constexpr Ec0() : v() {
      for (int c = 0; c < kGrid; ++c) {
        const int x = c % D, y = c / D;
        const u32 e = on_board(x, y) ? (u32)(corner_of(x, y) | (edge_of(x, y) << 6)) : 0u;
        v[c >> 1] |= e << (16 * (c & 1));
      }
    }
  };
  BSG_HD static void reset_groups(State& s) {            // every cell its own group
    constexpr Ec0 ec0;
    u32* par = (u32*)s.parent;
    for (int k = 0; k < (kGrid + 3) / 4; ++k) par[k] = 0x03020100u + 0x04040404u * (u32)k;
    u32* ec = (u32*)s.ec;
    for (int k = 0; k < (kGrid + 1) / 2; ++k) ec[k] = ec0.v[k];
  }
  BSG_HD static void init(State& s) {
    for (int w = 0; w < W; ++w) s.st[0][w] = s.st[1][w] = 0;
    reset_groups(s);
    s.left = kCells; s.mv = 0; s.outc = kOutcomeNone; s.last = 0; s.comps = 0;
  }
  // Cell c, fresh (L < 0) or already under the leader L, meets its friendly neighbour m.
  BSG_HD static int attach(State& s, int c, int m, int L, u32& merges) {
    const int lb = find(s, m);
    merges += (L < 0 || lb != L) ? 1u : 0u;
    if (L < 0) {
      s.parent[c] = (unsigned char)lb;
      s.ec[lb] = (unsigned short)(s.ec[lb] | s.ec[c]);
      return lb;
    }
    if (lb != L) {
      s.parent[lb] = (unsigned char)L;
      s.ec[L] = (unsigned short)(s.ec[L] | s.ec[lb]);
    }
    return L;
  }
  // Groups from the stones: in ascending order every stone meets its lower-index neighbours of
  // its colour (directions 5, 1, 0: cells c - 1, c - D, c - D - 1; a subtraction that leaves the
  // row lands off the board, where no stone is — see nb6).  c - D - 1 touches the other two, so
  // when it is there it is the only one to meet.  Which neighbours are there is read off three
  // shifted copies of the stones in registers.
  BSG_HD static void rebuild_groups(State& s) {
    reset_groups(s);
    s.comps = 0;
    for (int p = 0; p < 2; ++p) {
      u32 merges = 0, stones = 0;
      u32 st[W], f1[W], fd[W], fd1[W];
      for (int w = 0; w < W; ++w) st[w] = s.st[p][w];
      shl_words<1, W>(st, f1);
      shl_words<D, W>(st, fd);
      shl_words<D + 1, W>(st, fd1);
      for (int w = 0; w < W; ++w) {
        u32 m = st[w];
        while (m) {
          const int bit = ctz32(m);
          const int c = 32 * w + bit;
          m &= m - 1u;
          stones += 1u;
          if ((fd1[w] >> bit) & 1u) { attach(s, c, c - D - 1, -1, merges); continue; }
          int L = -1;
          if ((f1[w] >> bit) & 1u) L = attach(s, c, c - 1, L, merges);
          if ((fd[w] >> bit) & 1u) L = attach(s, c, c - D, L, merges);
        }
      }
      s.comps |= (stones - merges) << (16 * p);
    }
  }
  BSG_HD static void load(State& s, const void* base, int64_t n, int64_t i) {
    const u32* p0 = (const u32*)base + i * PW;
    const u32* p1 = (const u32*)base + (int64_t)PW * n + i * PW;
    const u32 meta = ((const u32*)base + (int64_t)2 * PW * n)[i];
    int stones = 0;
    for (int w = 0; w < W; ++w) {
      s.st[0][w] = p0[w]; s.st[1][w] = p1[w];
      stones += popc32(p0[w]) + popc32(p1[w]);
    }
    s.left = (u32)(kCells - stones);
    s.mv = meta & 1u;
    s.outc = (meta >> 1) & 3u;
    s.last = meta >> 8;
    rebuild_groups(s);
  }
  BSG_HD static void store(const State& s, void* base, int64_t n, int64_t i) {
    u32* p0 = (u32*)base + i * PW;
    u32* p1 = (u32*)base + (int64_t)PW * n + i * PW;
    for (int w = 0; w < PW; ++w) { p0[w] = w < W ? s.st[0][w] : 0u; p1[w] = w < W ? s.st[1][w] : 0u; }
    ((u32*)base + (int64_t)2 * PW * n)[i] = s.mv | (s.outc << 1) | (s.last << 8);
  }
  BSG_HD static bool terminal(const State& s) { return s.outc != (u32)kOutcomeNone; }
  BSG_HD static bool won(const State& s) { return s.outc == (u32)kOutcomeP0 || s.outc == (u32)kOutcomeP1; }
  BSG_HD static int mover(const State& s) { return (int)s.mv; }
  BSG_HD static int outcome(const State& s) { return (int)s.outc; }

  // The six neighbours of every cell of a set of board cells, as shifts of the x + y * D grid by
  // 1, D and D + 1.  A shift that leaves a row lands off the board (the board is a hexagon in
  // the grid: (D - 1, y) on the board means (0, y + 1) and (0, y + 2) are not, and the same at
  // the other end), so the board mask is all the clipping there is.
  BSG_HD static void nb6(const u32* f, u32* out) {
    u32 a[W], b[W], c[W], d[W], e[W], g[W];
    shl_words<1, W>(f, a);
    shr_words<1, W>(f, b);
    shl_words<D, W>(f, c);
    shr_words<D, W>(f, d);
    shl_words<D + 1, W>(f, e);
    shr_words<D + 1, W>(f, g);
    for (int w = 0; w < W; ++w) out[w] = (a[w] | b[w] | c[w] | d[w] | e[w] | g[w]) & board_word(w);
  }
  // Can `mine` hold a ring at all?  The reference's search (below) finds a closed walk that turns
  // by at most 60 degrees per step except where it closes; a closed walk on the triangular
  // lattice with at most one sharper corner has a lattice point inside it (the walks without one
  // are width-one strips and the side-two triangle: two or three sharp corners).  That cell is
  // either not the mover's — then the mover's stones have a HOLE — or every cell inside is the
  // mover's and one of them has six neighbours of the mover's.  Holes are counted without a
  // search: the stones with their adjacencies and their filled triangles are a complex in the
  // plane with Euler characteristic V - E + F = groups - holes, the groups are counted as they
  // merge (State::comps), and V, E, F are popcounts of the set and-ed with shifted copies of
  // itself.  All of it is set arithmetic in registers; the search, a backtracking walk over
  // local-memory stacks, runs only when a ring is possible.
  BSG_HD static bool ring_possible(const u32* mine_in, u32 groups) {
    // (one shifted copy at a time besides the c - D - 1 one: few live registers)
    u32 mine[W], e[W], sol[W], t[W];
#pragma unroll
    for (int w = 0; w < W; ++w) mine[w] = mine_in[w];
    int chi = 0;
    shl_words<D + 1, W>(mine, e);      // the stone at c - D - 1 (touches the next two)
    for (int w = 0; w < W; ++w) { sol[w] = mine[w] & e[w]; chi += popc32(mine[w]) - popc32(sol[w]); }
    shl_words<1, W>(mine, t);          // c - 1
    for (int w = 0; w < W; ++w) {
      const u32 m = mine[w] & t[w];
      chi += popc32(m & e[w]) - popc32(m);
      sol[w] &= t[w];
    }
    shl_words<D, W>(mine, t);          // c - D
    for (int w = 0; w < W; ++w) {
      const u32 m = mine[w] & t[w];
      chi += popc32(m & e[w]) - popc32(m);
      sol[w] &= t[w];
    }
    shr_words<1, W>(mine, t);
    for (int w = 0; w < W; ++w) sol[w] &= t[w];
    shr_words<D, W>(mine, t);
    for (int w = 0; w < W; ++w) sol[w] &= t[w];
    shr_words<D + 1, W>(mine, t);
    u32 solid = 0;
    for (int w = 0; w < W; ++w) solid |= sol[w] & t[w];
    return solid != 0u || (int)groups - chi > 0;
  }

  // CheckRingDFS(move, 0, 3) (havannah.cc:385-399) as a loop over the recursion's frames
  // (cell, next direction to try, last direction).  A frame is entered on an unmarked stone of
  // `mine`; meeting a marked one — a cell of the current path — is a ring.
  BSG_HD static bool ring_from(const u32* mine, int start, const Tables& t) {
#if defined(BSG_HV_NORING)
    return false;   // timing experiment only
#endif
    u32 mark[W];
    for (int w = 0; w < W; ++w) mark[w] = 0;
    unsigned char cell[kCells + 1];
    signed char next[kCells + 1], last[kCells + 1];
    int top = 0;
    cell[0] = (unsigned char)start; next[0] = 0; last[0] = 3;
    mark[start >> 5] |= 1u << (start & 31);
    while (top >= 0) {
      if (next[top] > last[top]) {                         // every direction failed: unmark, return
        const int c = cell[top];
        mark[c >> 5] &= ~(1u << (c & 31));
        --top;
        continue;
      }
      const int i = next[top]++;
      const int dir = (i + 6) % 6;
      const int m = t.nb[cell[top]][dir];
      if (m == kOff || !has(mine, m)) continue;
      if (has(mark, m)) return true;
      if (top + 1 > kCells) return false;                  // (cannot happen: a path has distinct cells)
      mark[m >> 5] |= 1u << (m & 31);
      ++top;
      cell[top] = (unsigned char)m; next[top] = (signed char)(dir - 1); last[top] = (signed char)(dir + 1);
    }
    return false;
  }

  // havannah.cc:314-349 for the mover.
  BSG_HD static void place(State& s, int c, const Tables& t) {
    const u32 me = s.mv;
    s.st[me][c >> 5] |= 1u << (c & 31);
    s.left -= 1u;
    s.last = (u32)c + 1u;
    bool already = false, skip = false;
    // c is a fresh one-cell group: it goes under the leader L of the first friendly neighbour, the
    // leaders of further neighbours go under L too, and the group's corner / edge bits are
    // gathered in a register and written once.  `already`: a neighbour that is in c's group
    // before it is joined (havannah.cc:364-383 returns that).
    int L = -1;
    u32 g = s.ec[c], merges = 0;
    for (int d = 0; d < 6; ++d) {
      if (skip) { skip = false; continue; }
      const int m = t.nb[c][d];
      if (m != kOff && has(s.st[me], m)) {
        const int lb = find(s, m);
        if (L < 0) { L = lb; s.parent[c] = (unsigned char)lb; g |= s.ec[lb]; ++merges; }
        else if (lb == L) already = true;
        else { s.parent[lb] = (unsigned char)L; g |= s.ec[lb]; ++merges; }
        skip = true;     // the next neighbour touches this one: same group or a sharp corner
      }
    }
    if (L >= 0) s.ec[L] = (unsigned short)g;
    s.comps += (1u - merges) << (16u * me);              // (two's complement: the other half is not touched
                                                         //  as long as this half stays >= 0, which it does)
    const bool win = popc32(g >> 6) >= 3 || popc32(g & 63u) >= 2 ||
                     (already && ring_possible(s.st[me], (s.comps >> (16u * me)) & 0xFFFFu) && ring_from(s.st[me], c, t));
    if (win) s.outc = me;
    else if (s.left == 0) s.outc = (u32)kOutcomeDraw;
    s.mv = me ^ 1u;
  }

  BSG_HD static u32 empty_word(const State& s, int w) { return ~(s.st[0][w] | s.st[1][w]) & board_word(w); }
  // k-th (ascending) empty cell, k < left
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
    return (int)(cell < (u32)kGrid ? cell : 0u);
  }
  BSG_HD static u32 step(State& s, u32 r, const Tables& tab, int& action) {
    // a finished lane leaves at once: the group arrays and the ring search are not made for
    // dead steps on a dead state
    if (terminal(s) || s.left == 0) { action = 0; return 1u; }
    const u32 n = s.left;                                  // havannah.cc:182-194: every empty cell
    const u32 k = fastmod(r, n, tab.modM[n & 255u]);
    const int c = select_empty(s, k, tab);
    action = c;
    place(s, c, tab);
    return n;
  }
  BSG_HD static int legal(const State& s, int* out) {
    if (terminal(s)) return 0;
    int n = 0;
    for (int c = 0; c < kGrid; ++c) if ((empty_word(s, c >> 5) >> (c & 31)) & 1u) out[n++] = c;
    return n;
  }
  BSG_HD static bool apply(State& s, int a, const Tables& t) {
    if (terminal(s) || a < 0 || a >= kGrid || !((empty_word(s, a >> 5) >> (a & 31)) & 1u)) return false;
    place(s, a, t);
    return true;
  }
  BSG_HD static u32 legal_bit(const State& s, int a) {
    return (!terminal(s) && ((empty_word(s, a >> 5) >> (a & 31)) & 1u)) ? 1u : 0u;
  }
  // havannah.cc:284-312: plane 0 the observer's stones, 1 the other player's, 2 empty; the
  // cut-off corners stay zero in every plane.
  BSG_HD static u32 plane_word(const State& s, int player, int plane, int w) {
    return plane == 0 ? s.st[player][w] : (plane == 1 ? s.st[player ^ 1][w] : empty_word(s, w));
  }
  BSG_HD static u32 obs_bit(const State& s, int player, int idx) {
    const int plane = idx / kGrid, cell = idx - plane * kGrid;
    return (plane_word(s, player, plane, cell >> 5) >> (cell & 31)) & 1u;
  }
  BSG_HD static void obs_words(const State& s, int player, u32* out) {
    for (int w = 0; w < kObsWords; ++w) out[w] = 0;
    for (int p = 0; p < 3; ++p) {
      const int off = p * kGrid, wi = off >> 5, sh = off & 31;
      for (int j = 0; j < W; ++j) {
        const u32 m = plane_word(s, player, p, j);
        if (wi + j < kObsWords) out[wi + j] |= m << sh;
        if (sh != 0 && wi + j + 1 < kObsWords) out[wi + j + 1] |= m >> ((32 - sh) & 31);
      }
    }
  }
  BSG_HD static void mask_words(const State& s, u32* out) {
    for (int w = 0; w < W; ++w) out[w] = terminal(s) ? 0u : empty_word(s, w);
  }
};
typedef HavannahT<8> Havannah;

}  // namespace bsg
