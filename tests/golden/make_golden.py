#!/usr/bin/env python
"""Generates tests/golden/*.json from the reference tree (run in the build container only).

Sources (all under /root/reference, read in place, never copied verbatim):
  * open_spiel/integration_tests/playthroughs/{tic_tac_toe,connect_four,breakthrough,
    hex(board_size=5),pentago,othello,y(board_size=9),havannah(board_size=4)}.txt — the reference's own golden playthroughs; per state we keep the
    action history, LegalActions(), CurrentPlayer(), IsTerminal(), Returns() and both
    players' ObservationTensor planes (as flat 0/1 lists in CHW order).
  * open_spiel/games/connect_four_test.cc:33-80 — FastLoss / DeserializeDraw known answers.
  * open_spiel/integration_tests/example_connect_four_state.txt — a mid-game board string.
  * outputs of the unmodified reference compiled here (oracle/_ref/libspiel_ref.so) for what
    no reference file pins: hex 11x11 and mid-game replay vectors under the Philox contract.

Usage: python tests/golden/make_golden.py
"""
import ctypes
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/open_spiel"
PT = os.path.join(REF, "integration_tests", "playthroughs")

GAMES = {
    "tic_tac_toe": ("tic_tac_toe.txt", (3, 3, 3)),
    "connect_four": ("connect_four.txt", (3, 6, 7)),
    "breakthrough": ("breakthrough.txt", (3, 8, 8)),
    "hex5": ("hex(board_size=5).txt", (9, 5, 5)),
    "pentago": ("pentago.txt", (3, 6, 6)),
    "othello": ("othello.txt", (3, 8, 8)),
    "y9": ("y(board_size=9).txt", (3, 9, 9)),
    "havannah4": ("havannah(board_size=4).txt", (3, 7, 7)),
    "go7": ("go.txt", (4, 7, 7)),
}


def parse_planes(lines, shape):
  """lines: H text rows, each with C groups of W glyphs separated by two spaces."""
  c, h, w = shape
  assert len(lines) == h, (len(lines), shape)
  out = np.zeros(shape, np.uint8)
  for r, line in enumerate(lines):
    groups = line.split("  ")
    assert len(groups) == c, (line, shape)
    for p, g in enumerate(groups):
      assert len(g) == w
      for x, ch in enumerate(g):
        out[p, r, x] = 1 if ch == "◉" else 0
  return out.ravel().tolist()


def parse_playthrough(path, shape):
  states = []
  cur = None
  actions = []
  with open(path, encoding="utf-8") as f:
    lines = [l.rstrip("\n") for l in f]
  i = 0
  while i < len(lines):
    line = lines[i]
    if line.startswith("# State "):
      cur = {"history": list(actions)}
      states.append(cur)
    elif cur is not None:
      m = re.match(r"IsTerminal\(\) = (\w+)", line)
      if m:
        cur["terminal"] = m.group(1) == "True"
      m = re.match(r"CurrentPlayer\(\) = (-?\d+)", line)
      if m:
        cur["current_player"] = int(m.group(1))
      m = re.match(r"History\(\) = \[(.*)\]", line)
      if m:
        h = [int(x) for x in m.group(1).split(",") if x.strip()]
        assert h == cur["history"], (h, cur["history"])
      m = re.match(r"ObservationTensor\((\d)\):$", line)
      if m:
        cur["obs%s" % m.group(1)] = parse_planes(lines[i + 1:i + 1 + shape[1]], shape)
        i += shape[1]
      m = re.match(r"Returns\(\) = \[(.*)\]", line)
      if m:
        cur["returns"] = [int(float(x)) for x in m.group(1).split(",")]
      m = re.match(r"LegalActions\(\) = \[(.*)\]", line)
      if m:
        cur["legal"] = [int(x) for x in m.group(1).split(",") if x.strip()]
      m = re.match(r"action: (\d+)", line)
      if m:
        actions.append(int(m.group(1)))
    i += 1
  for s in states:
    s.setdefault("legal", [])  # terminal states print no LegalActions()
    assert "obs0" in s and "obs1" in s and "returns" in s
  return states


def board_string_to_history(rows):
  """connect_four board string (top row first) -> some legal alternating history reaching it."""
  grid = [list(r) for r in rows][::-1]  # bottom first
  xs = [(r, c) for c in range(7) for r in range(6) if grid[r][c] == "x"]
  os_ = [(r, c) for c in range(7) for r in range(6) if grid[r][c] == "o"]
  heights = [0] * 7
  hist = []
  placed = set()

  def solve(turn):
    if len(placed) == len(xs) + len(os_):
      return True
    pool = xs if turn == 0 else os_
    for (r, c) in pool:
      if (r, c) not in placed and heights[c] == r:
        placed.add((r, c)); heights[c] += 1; hist.append(c)
        if solve(1 - turn):
          return True
        placed.discard((r, c)); heights[c] -= 1; hist.pop()
    return False

  assert solve(0)
  return list(hist)


def reference_vectors(lib):
  """Vectors only the compiled reference can pin: Philox-contract replays + state dumps."""
  out = {}
  lib.ref_replay.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                             ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, ctypes.c_void_p,
                             ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
  lib.ref_state_dump.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                 ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                 ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
  sizes = {"tic_tac_toe": 27, "connect_four": 126, "breakthrough": 192, "hex": 1089, "pentago": 108, "othello": 192, "y": 1083}
  for game in ["tic_tac_toe", "connect_four", "breakthrough", "hex", "pentago", "othello", "y"]:
    recs = []
    for playout in range(12):
      # root = first `d` moves of another replay (seed 7), then replay with seed 20240611.
      d = [0, 0, 1, 2, 3, 4, 5, 6, 8, 10, 12, 14][playout]
      acts = np.zeros(256, np.int64)
      ln, term, ret = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
      lib.ref_replay(game.encode(), None, 0, None, 7, playout, d, acts.ctypes.data,
                     ctypes.byref(ln), ctypes.byref(term), ctypes.byref(ret))
      root = acts[:ln.value].tolist()
      if term.value:
        root = root[:-1]
      rh = np.array(root, np.int64)
      lib.ref_replay(game.encode(), rh.ctypes.data, len(root), None, 20240611, playout, 1000,
                     acts.ctypes.data, ctypes.byref(ln), ctypes.byref(term), ctypes.byref(ret))
      full = root + acts[:ln.value].tolist()
      rec = {"root": root, "seed": 20240611, "playout": playout,
             "actions": acts[:ln.value].tolist(), "terminal": term.value, "ret0": ret.value}
      # observation / legal dump at the final state and at the midpoint, both observers
      dumps = []
      for cut in sorted({len(full), (len(root) + len(full)) // 2}):
        h = np.array(full[:cut], np.int64)
        for player in (0, 1):
          obs = np.zeros(sizes[game], np.float64)
          legal = np.zeros(1024, np.int64)
          nl, cp, tm = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
          rets = np.zeros(2, np.float64)
          lib.ref_state_dump(game.encode(), h.ctypes.data, cut, player, obs.ctypes.data,
                             legal.ctypes.data, ctypes.byref(nl), ctypes.byref(cp),
                             ctypes.byref(tm), rets.ctypes.data)
          dumps.append({"cut": cut, "player": player,
                        "obs_on": np.nonzero(obs)[0].tolist(),
                        "legal": legal[:nl.value].tolist(), "current_player": cp.value,
                        "terminal": tm.value, "returns": [int(rets[0]), int(rets[1])]})
      rec["dumps"] = dumps
      recs.append(rec)
    out[game] = recs
  return out


def main():
  for name, (fname, shape) in GAMES.items():
    states = parse_playthrough(os.path.join(PT, fname), shape)
    with open(os.path.join(HERE, "playthrough_%s.json" % name), "w") as f:
      json.dump({"source": "open_spiel/integration_tests/playthroughs/" + fname,
                 "obs_shape": list(shape), "states": states}, f, separators=(",", ":"))
    print(name, len(states), "states")

  with open(os.path.join(REF, "integration_tests", "example_connect_four_state.txt")) as f:
    rows = [l.strip() for l in f if l.strip()]
  known = {
      "connect_four_fast_loss": {  # connect_four_test.cc:33-53
          "history": [3, 3, 4, 4, 2, 2, 1], "terminal_before_last": False, "terminal": True,
          "returns": [1, -1],
          "board": [".......", ".......", ".......", ".......", "..ooo..", ".xxxx.."]},
      "connect_four_draw": {  # connect_four_test.cc:62-80
          "board": ["ooxxxoo", "xxoooxx", "ooxxxoo", "xxoooxx", "ooxxxoo", "xxoooxx"],
          "terminal": True, "returns": [0, 0]},
      "connect_four_example_state": {  # integration_tests/example_connect_four_state.txt
          "board": rows, "history": board_string_to_history(rows)},
      "tic_tac_toe_counts": {  # python/algorithms/get_all_states_test.py:29-45, tic_tac_toe.h:42
          "histories": 549946, "boards": 5478},
      "mcts_solver": [  # python/algorithms/mcts_test.py:115-153 (cell = 3*row + col)
          {"history": [4, 0, 8], "root_outcome": 0, "best_actions": [2, 6]},
          {"history": [4, 0, 8, 1, 2], "root_outcome": -1},
          {"history": [1, 8], "root_outcome": 1, "best_actions": [2]}],
  }
  with open(os.path.join(HERE, "known_answers.json"), "w") as f:
    json.dump(known, f, indent=1)

  lib = ctypes.CDLL(os.path.join(REPO, "oracle", "_ref", "libspiel_ref.so"))
  vec = reference_vectors(lib)
  with open(os.path.join(HERE, "reference_vectors.json"), "w") as f:
    json.dump({"source": "unmodified reference via oracle/_ref/libspiel_ref.so "
                         "(oracle/ref_wrapper.cc: ref_replay, ref_state_dump)",
               "games": vec}, f, separators=(",", ":"))
  print("reference vectors written")


if __name__ == "__main__":
  sys.exit(main())
