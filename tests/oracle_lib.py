"""ctypes access to the CPU checkers (TEST INFRASTRUCTURE ONLY; never imported by the product).

  Oracle    — oracle/_ref/libbsg_oracle.so, the plain-C restatement (oracle/bsg_oracle.c)
  Reference — oracle/_ref/libspiel_ref.so, the unmodified reference C++ (oracle/ref_wrapper.cc);
              only present where it was built (this container; it travels to the GPU box).
"""
import ctypes
import os
import subprocess

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(REPO, "oracle")
GAME_IDS = {"tic_tac_toe": 0, "connect_four": 1, "breakthrough": 2, "hex": 3, "pentago": 5, "othello": 6, "y": 7, "havannah": 9, "go": 11, "go(board_size=9)": 11}
OBS_SIZE = {"tic_tac_toe": 27, "connect_four": 126, "breakthrough": 192, "hex": 1089, "pentago": 108, "othello": 192, "y": 1083, "havannah": 675, "go": 324, "go(board_size=9)": 324}
vp = ctypes.c_void_p


def obs_size(game, board_size=11):
  if game == "hex":
    return 9 * board_size * board_size
  if game == "y":
    n = 19 if board_size == 11 else board_size
    return 3 * n * n
  if game == "havannah":
    n = 8 if board_size == 11 else board_size
    return 3 * (2 * n - 1) * (2 * n - 1)
  if game.startswith("go"):
    n = 7 if board_size == 7 else 9
    return 4 * n * n
  return OBS_SIZE[game]


def _ptr(a):
  return None if a is None else a.ctypes.data


def build_oracle():
  so = os.path.join(ORACLE_DIR, "_ref", "libbsg_oracle.so")
  src = os.path.join(ORACLE_DIR, "bsg_oracle.c")
  if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
    subprocess.check_call(["make", "-C", ORACLE_DIR, "_ref/libbsg_oracle.so"],
                          stdout=subprocess.DEVNULL)
  return so


def ragged_from_log(log, lens):
  """Dense action log (int16 [n, stride], lens [n]) -> (flat int64, off int64) without loops."""
  lens = np.asarray(lens, np.int64)
  off = np.zeros(len(lens) + 1, np.int64)
  np.cumsum(lens, out=off[1:])
  keep = np.arange(log.shape[1])[None, :] < lens[:, None]
  flat = log[keep].astype(np.int64)
  if flat.size == 0:
    flat = np.zeros(1, np.int64)
  return flat, off


def ragged(histories):
  if isinstance(histories, tuple):      # already (flat, off)
    return histories
  off = np.zeros(len(histories) + 1, np.int64)
  for i, h in enumerate(histories):
    off[i + 1] = off[i] + len(h)
  flat = np.zeros(max(int(off[-1]), 1), np.int64)
  for i, h in enumerate(histories):
    flat[off[i]:off[i + 1]] = h
  return flat, off


class Oracle:
  """The C restatement.  `game` is one of GAME_IDS; hex takes board_size (default 11)."""

  def __init__(self):
    self.lib = ctypes.CDLL(build_oracle())
    L = self.lib
    L.orc_replay.argtypes = [ctypes.c_int, ctypes.c_int, vp, ctypes.c_int, vp, ctypes.c_uint64,
                             ctypes.c_uint64, ctypes.c_int, vp, vp, vp, vp]
    L.orc_replay_batch.argtypes = [ctypes.c_int, ctypes.c_int, vp, vp, ctypes.c_int64,
                                   ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int,
                                   vp, vp, vp, vp, ctypes.c_int]
    L.orc_state_dump.argtypes = [ctypes.c_int, ctypes.c_int, vp, ctypes.c_int, ctypes.c_int, vp,
                                 vp, vp, vp, vp, vp]
    L.orc_random_playouts.argtypes = [ctypes.c_int, ctypes.c_int, vp, vp, ctypes.c_int64,
                                      ctypes.c_int64, ctypes.c_uint64, vp]
    L.orc_random_playouts.restype = ctypes.c_int64
    L.orc_draw.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint32]
    L.orc_draw.restype = ctypes.c_uint32

  def draw(self, seed, playout, t):
    return self.lib.orc_draw(seed, playout, t)

  def replay(self, game, hist, r_table=None, seed=0, playout=0, max_steps=1 << 20, board_size=11):
    h = np.asarray(hist, np.int64)
    acts = np.zeros(256, np.int64)
    ln, term, ret = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    rt = None if r_table is None else np.asarray(r_table, np.uint32)
    rc = self.lib.orc_replay(GAME_IDS[game], board_size, _ptr(h), len(h), _ptr(rt), seed, playout,
                             max_steps, _ptr(acts), ctypes.byref(ln), ctypes.byref(term),
                             ctypes.byref(ret))
    if rc != 0:
      raise RuntimeError("orc_replay failed rc=%d" % rc)
    return acts[:ln.value].tolist(), term.value, ret.value

  def replay_batch(self, game, histories, rollouts=1, seed=0, first_playout=0,
                   max_steps=1 << 20, actions_stride=0, board_size=11):
    flat, off = ragged(histories)
    n_roots = len(off) - 1
    n = n_roots * rollouts
    lens = np.zeros(n, np.int32)
    ret0 = np.zeros(n, np.int8)
    term = np.zeros(n, np.uint8)
    acts = np.full((n, actions_stride), -1, np.int16) if actions_stride else None
    rc = self.lib.orc_replay_batch(GAME_IDS[game], board_size, _ptr(flat), _ptr(off),
                                   n_roots, rollouts, seed, first_playout, max_steps,
                                   _ptr(lens), _ptr(ret0), _ptr(term), _ptr(acts), actions_stride)
    if rc != 0:
      raise RuntimeError("orc_replay_batch failed rc=%d" % rc)
    return lens, ret0, term, acts

  def state_dump(self, game, hist, player=0, board_size=11):
    h = np.asarray(hist, np.int64)
    size = obs_size(game, board_size)
    obs = np.zeros(size, np.uint8)
    legal = np.zeros(1024, np.int64)
    nl, cp, tm = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    rets = np.zeros(2, np.int32)
    rc = self.lib.orc_state_dump(GAME_IDS[game], board_size, _ptr(h), len(h), player, _ptr(obs),
                                 _ptr(legal), ctypes.byref(nl), ctypes.byref(cp),
                                 ctypes.byref(tm), _ptr(rets))
    if rc != 0:
      raise RuntimeError("orc_state_dump failed rc=%d (illegal history?)" % rc)
    return {"obs": obs, "legal": legal[:nl.value].tolist(), "current_player": cp.value,
            "terminal": bool(tm.value), "returns": rets.tolist()}

  def random_playouts(self, game, histories, n_playouts, seed=0, board_size=11):
    flat, off = ragged(histories)
    s = ctypes.c_int64()
    steps = self.lib.orc_random_playouts(GAME_IDS[game], board_size, _ptr(flat), _ptr(off),
                                         len(histories), n_playouts, seed, ctypes.byref(s))
    if steps < 0:
      raise RuntimeError("orc_random_playouts failed")
    return steps, s.value


def reference_path():
  return os.path.join(ORACLE_DIR, "_ref", "libspiel_ref.so")


class Reference:
  """The unmodified reference C++ behind oracle/ref_wrapper.cc."""

  def __init__(self):
    self.lib = ctypes.CDLL(reference_path())
    L = self.lib
    L.ref_replay.argtypes = [ctypes.c_char_p, vp, ctypes.c_int, vp, ctypes.c_uint64,
                             ctypes.c_uint64, ctypes.c_int, vp, vp, vp, vp]
    L.ref_replay_batch.argtypes = [ctypes.c_char_p, vp, vp, ctypes.c_int64, ctypes.c_int,
                                   ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, ctypes.c_int,
                                   vp, vp, vp, vp, ctypes.c_int]
    L.ref_state_dump.argtypes = [ctypes.c_char_p, vp, ctypes.c_int, ctypes.c_int, vp, vp, vp, vp,
                                 vp, vp]
    L.ref_replay_log.argtypes = [ctypes.c_char_p, vp, ctypes.c_int, vp, ctypes.c_int64, ctypes.c_int,
                                 ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, vp, vp, vp]
    L.ref_state_dump_log.argtypes = [ctypes.c_char_p, vp, ctypes.c_int, vp, ctypes.c_int64, vp,
                                     ctypes.c_int, vp, vp, vp, vp]
    L.ref_game_info.argtypes = [ctypes.c_char_p, vp, vp, vp]
    L.ref_random_playouts.argtypes = [ctypes.c_char_p, vp, vp, ctypes.c_int64, ctypes.c_int64,
                                      ctypes.c_uint64, ctypes.c_int, ctypes.c_int, vp, vp]
    L.ref_random_playouts.restype = ctypes.c_double
    L.ref_mcts_selfplay.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_double, ctypes.c_int, ctypes.c_uint64, ctypes.c_int,
                                    vp, vp, vp]
    L.ref_mcts_selfplay.restype = ctypes.c_double
    L.ref_mcts_step.argtypes = [ctypes.c_char_p, vp, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                ctypes.c_double, ctypes.c_int, ctypes.c_int]
    L.ref_mcts_step.restype = ctypes.c_int64
    L.ref_to_string.argtypes = [ctypes.c_char_p, vp, ctypes.c_int, ctypes.c_int, ctypes.c_char_p,
                                ctypes.c_int]
    L.ref_serialize_game_and_state.argtypes = [ctypes.c_char_p, vp, ctypes.c_int, ctypes.c_char_p,
                                               ctypes.c_int]
    L.ref_deserialize_dump.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int, vp, vp, vp,
                                       vp, vp, vp, ctypes.c_char_p, ctypes.c_int]

  @staticmethod
  def game_string(game, board_size=11):
    if game == "y":   # default 19 (y.h:38); callers that pass the hex default mean "default"
      return ("y(board_size=%d)" % (19 if board_size == 11 else board_size)).encode()
    if game.startswith("go"):   # the device game is 9 x 9 at the default komi; the playthrough is 7 x 7 at komi 4.5
      return b"go(board_size=7,komi=4.5)" if board_size == 7 else b"go(board_size=9)"
    if game == "havannah":   # default 8 (havannah.h:39)
      return ("havannah(board_size=%d)" % (8 if board_size == 11 else board_size)).encode()
    return ("hex(board_size=%d)" % board_size if game == "hex" else game).encode()

  def replay(self, game, hist, r_table=None, seed=0, playout=0, max_steps=1 << 20, board_size=11):
    h = np.asarray(hist, np.int64)
    acts = np.zeros(256, np.int64)
    ln, term, ret = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    rt = None if r_table is None else np.asarray(r_table, np.uint32)
    rc = self.lib.ref_replay(self.game_string(game, board_size), _ptr(h), len(h), _ptr(rt), seed,
                             playout, max_steps, _ptr(acts), ctypes.byref(ln), ctypes.byref(term),
                             ctypes.byref(ret))
    if rc != 0:
      raise RuntimeError("ref_replay failed")
    return acts[:ln.value].tolist(), term.value, ret.value

  def replay_batch(self, game, histories, rollouts=1, seed=0, first_playout=0,
                   max_steps=1 << 20, actions_stride=0, threads=8, board_size=11):
    flat, off = ragged(histories)
    n_roots = len(off) - 1
    n = n_roots * rollouts
    lens = np.zeros(n, np.int32)
    ret0 = np.zeros(n, np.int8)
    term = np.zeros(n, np.uint8)
    acts = np.full((n, actions_stride), -1, np.int16) if actions_stride else None
    rc = self.lib.ref_replay_batch(self.game_string(game, board_size), _ptr(flat), _ptr(off),
                                   n_roots, rollouts, seed, first_playout, max_steps,
                                   threads, _ptr(lens), _ptr(ret0), _ptr(term), _ptr(acts),
                                   actions_stride)
    if rc != 0:
      raise RuntimeError("ref_replay_batch failed")
    return lens, ret0, term, acts

  def replay_log(self, game, log, root_len, rollouts=1, seed=0, first_playout=0, threads=8):
    """ref_replay_log: replay from roots given as a dense int16 action log (roots.make_roots
    want_log).  Returns (lens int32, ret0 int8, terminal uint8) per playout."""
    log = np.ascontiguousarray(log, np.int16)
    root_len = np.ascontiguousarray(root_len, np.int32)
    n = len(root_len) * rollouts
    lens, ret0, term = np.zeros(n, np.int32), np.zeros(n, np.int8), np.zeros(n, np.uint8)
    rc = self.lib.ref_replay_log(self.game_string(game), _ptr(log), log.shape[1], _ptr(root_len),
                                 len(root_len), rollouts, seed, first_playout, threads, _ptr(lens),
                                 _ptr(ret0), _ptr(term))
    if rc != 0:
      raise RuntimeError("ref_replay_log failed")
    return lens, ret0, term

  def state_dump_log(self, game, log, hist_len, players=None, threads=8, want_obs=True, want_mask=True):
    """ref_state_dump_log: ObservationTensor(players[i]) and LegalActionsMask as bytes, current
    player and terminal flag of the states reached by the rows of a dense int16 action log."""
    log = np.ascontiguousarray(log, np.int16)
    hist_len = np.ascontiguousarray(hist_len, np.int32)
    n = len(hist_len)
    na, ml, osz = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    self.lib.ref_game_info(self.game_string(game), ctypes.byref(na), ctypes.byref(ml), ctypes.byref(osz))
    obs = np.zeros((n, osz.value), np.uint8) if want_obs else None
    mask = np.zeros((n, na.value), np.uint8) if want_mask else None
    cur, term = np.zeros(n, np.int32), np.zeros(n, np.uint8)
    pl = None if players is None else np.ascontiguousarray(players, np.uint8)
    rc = self.lib.ref_state_dump_log(self.game_string(game), _ptr(log), log.shape[1], _ptr(hist_len),
                                     n, _ptr(pl), threads, _ptr(obs), _ptr(mask), _ptr(cur), _ptr(term))
    if rc != 0:
      raise RuntimeError("ref_state_dump_log failed")
    return {"obs": obs, "mask": mask, "current_player": cur, "terminal": term}

  def state_dump(self, game, hist, player=0, board_size=11):
    h = np.asarray(hist, np.int64)
    size = obs_size(game, board_size)
    obs = np.zeros(size, np.float64)
    legal = np.zeros(1024, np.int64)
    nl, cp, tm = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    rets = np.zeros(2, np.float64)
    self.lib.ref_state_dump(self.game_string(game, board_size), _ptr(h), len(h), player,
                            _ptr(obs), _ptr(legal), ctypes.byref(nl), ctypes.byref(cp),
                            ctypes.byref(tm), _ptr(rets))
    return {"obs": obs.astype(np.uint8), "legal": legal[:nl.value].tolist(),
            "current_player": cp.value, "terminal": bool(tm.value),
            "returns": [int(rets[0]), int(rets[1])]}

  def random_playouts(self, game, histories, n_playouts, seed=0, threads=1, with_obs=False,
                      board_size=11):
    flat, off = ragged(histories)
    steps, s = ctypes.c_int64(), ctypes.c_int64()
    secs = self.lib.ref_random_playouts(self.game_string(game, board_size), _ptr(flat), _ptr(off),
                                        len(histories), n_playouts, seed, threads, int(with_obs),
                                        ctypes.byref(steps), ctypes.byref(s))
    return secs, steps.value, s.value

  def mcts_selfplay(self, game, n_games, max_sims, n_rollouts, uct_c=2.0, solve=True, seed=1,
                    threads=1):
    mv, sims, s = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_int64()
    secs = self.lib.ref_mcts_selfplay(self.game_string(game), n_games, max_sims, n_rollouts,
                                      uct_c, int(solve), seed, threads, ctypes.byref(mv),
                                      ctypes.byref(sims), ctypes.byref(s))
    return secs, mv.value, sims.value, s.value

  def to_string(self, game, hist, serialize=False, board_size=11):
    h = np.asarray(hist, np.int64)
    buf = ctypes.create_string_buffer(4096)
    n = self.lib.ref_to_string(self.game_string(game, board_size), _ptr(h), len(h),
                               int(serialize), buf, 4096)
    assert n >= 0
    return buf.raw[:n].decode()

  def serialize_game_and_state(self, game, hist, board_size=11):
    h = np.asarray(hist, np.int64)
    buf = ctypes.create_string_buffer(8192)
    n = self.lib.ref_serialize_game_and_state(self.game_string(game, board_size), _ptr(h), len(h),
                                              buf, 8192)
    assert n >= 0
    return buf.raw[:n].decode()

  def deserialize_dump(self, game, text, player=None, board_size=11):
    size = obs_size(game, board_size)
    obs = np.zeros(size, np.float64)
    legal = np.zeros(1024, np.int64)
    nl, cp, tm = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    rets = np.zeros(2, np.float64)
    out = ctypes.create_string_buffer(4096)
    self.lib.ref_deserialize_dump(self.game_string(game, board_size), text.encode(),
                                  -1 if player is None else player, _ptr(obs), _ptr(legal),
                                  ctypes.byref(nl), ctypes.byref(cp), ctypes.byref(tm), _ptr(rets),
                                  out, 4096)
    return {"obs": obs.astype(np.uint8), "legal": legal[:nl.value].tolist(),
            "current_player": cp.value, "terminal": bool(tm.value),
            "returns": [int(rets[0]), int(rets[1])], "text": out.value.decode()}

  def mcts_step(self, game, hist, max_sims, n_rollouts=1, uct_c=2.0, solve=True, seed=1):
    h = np.asarray(hist, np.int64)
    return self.lib.ref_mcts_step(self.game_string(game), _ptr(h), len(h), max_sims, n_rollouts,
                                  uct_c, int(solve), seed)


def reference_python():
  """(pyspiel, mcts) — the reduced `pyspiel` module over the unmodified reference C++
  (oracle/pyspiel_min.cc) and the reference's python/algorithms/mcts.py byte-compiled from
  where it lies, both built by oracle/Makefile into oracle/_ref/.  None when not built."""
  import glob
  import importlib.machinery
  import importlib.util
  import sys
  ref_dir = os.path.dirname(reference_path())
  pyc = os.path.join(ref_dir, "ref_mcts.bytecode")
  if not glob.glob(os.path.join(ref_dir, "pyspiel*.so")) or not os.path.exists(pyc):
    return None
  if "ref_mcts" in sys.modules:
    return sys.modules["pyspiel"], sys.modules["ref_mcts"]
  if ref_dir not in sys.path:
    sys.path.insert(0, ref_dir)
  import pyspiel
  loader = importlib.machinery.SourcelessFileLoader("ref_mcts", pyc)
  spec = importlib.util.spec_from_loader("ref_mcts", loader)
  module = importlib.util.module_from_spec(spec)
  loader.exec_module(module)
  sys.modules["ref_mcts"] = module
  return pyspiel, module
