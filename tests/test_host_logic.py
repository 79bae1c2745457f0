"""CPU tests of the engine's bitboard rules through the C-ABI host helpers (no GPU needed).

The same __host__ __device__ game code runs inside the CUDA kernels, so these tests pin the
rules, the action ordering, the observation planes and — through bsg_host_replay_step — the
exact selection path of the kernels (look-up tables, fast modulo, rank-select) against the
oracle and the reference's golden playthroughs.  Playouts themselves are GPU-only.
"""
import ctypes
import json
import os

import numpy as np
import pytest

import open_spiel_1_b200 as bsg
from open_spiel_1_b200 import _cabi as C
from oracle_lib import Oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GAMES = [("tic_tac_toe", "tic_tac_toe", 11), ("connect_four", "connect_four", 11),
         ("breakthrough", "breakthrough", 11), ("hex", "hex", 11), ("hex(board_size=5)", "hex", 5),
         ("pentago", "pentago", 11), ("othello", "othello", 11),
         ("y", "y", 11), ("y(board_size=9)", "y", 9),
         ("havannah", "havannah", 11), ("havannah(board_size=4)", "havannah", 4),
         ("go(board_size=9)", "go", 9), ("go(board_size=7,komi=4.5)", "go", 7)]   # oracle: 11 = the default size, 19


@pytest.fixture(scope="module")
def oracle():
  return Oracle()


def check_state(oracle, st, oname, bs, hist):
  for player in (0, 1):
    d = oracle.state_dump(oname, hist, player, board_size=bs)
    assert st.legal_actions() == d["legal"], hist
    assert st.current_player() == d["current_player"]
    assert st.is_terminal() == d["terminal"]
    assert [int(x) for x in st.returns()] == d["returns"]
    assert np.array_equal(np.array(st.observation_tensor(player), np.uint8), d["obs"]), (hist, player)
    mask = st.legal_actions_mask()
    assert [i for i, m in enumerate(mask) if m] == d["legal"]


@pytest.mark.parametrize("fixture,name", [("tic_tac_toe", "tic_tac_toe"),
                                          ("connect_four", "connect_four"),
                                          ("breakthrough", "breakthrough"),
                                          ("hex5", "hex(board_size=5)"),
                                          ("pentago", "pentago"), ("othello", "othello"),
                                          ("y9", "y(board_size=9)"),
                                          ("havannah4", "havannah(board_size=4)"),
                                          ("go7", "go(board_size=7,komi=4.5)")])
def test_golden_playthroughs(fixture, name):
  with open(os.path.join(GOLD, "playthrough_%s.json" % fixture)) as f:
    data = json.load(f)
  game = bsg.load_game(name)
  assert game.observation_tensor_shape() == data["obs_shape"]
  for rec in data["states"]:
    st = game.new_initial_state()
    for a in rec["history"]:
      st.apply_action(a)
    assert st.legal_actions() == rec["legal"]
    assert st.current_player() == rec["current_player"]
    assert st.is_terminal() == rec["terminal"]
    assert [int(x) for x in st.returns()] == rec["returns"]
    if not rec["terminal"]:
      for p in (0, 1):
        assert [int(x) for x in st.observation_tensor(p)] == rec["obs%d" % p]
    else:
      for p in (0, 1):
        assert [int(x) for x in st.observation_tensor(p)] == rec["obs%d" % p]


@pytest.mark.parametrize("name,oname,bs", GAMES)
def test_random_games_match_oracle(oracle, name, oname, bs):
  game = bsg.load_game(name)
  rng = np.random.RandomState(7)
  n_games = 12 if oname in ("hex", "breakthrough") and bs == 11 else 40
  for g in range(n_games):
    st = game.new_initial_state()
    hist = []
    check_state(oracle, st, oname, bs, hist)
    while not st.is_terminal():
      la = st.legal_actions()
      a = la[rng.randint(len(la))]
      st.apply_action(a)
      hist.append(a)
      if rng.rand() < 0.35 or st.is_terminal():
        check_state(oracle, st, oname, bs, hist)
    # packing the whole history at once gives the same bytes as stepping
    assert np.array_equal(bsg.pack_histories(name, [hist]), st.packed())


@pytest.mark.parametrize("name,oname,bs", GAMES)
def test_kernel_step_path_matches_oracle_replay(oracle, name, oname, bs):
  """Drives bsg_host_replay_step (the kernels' step()) with the Philox draws and compares the
  whole action sequence, length, terminal flag and return with the oracle's replay."""
  gid = bsg.game_id(name)
  n = 25 if bs == 11 and oname in ("hex", "breakthrough", "y") else 150
  if oname == "go" and bs == 7:
    n = 1200        # enough 7 x 7 games for a few that end by superko (asserted below)
  superko = 0
  for playout in range(n):
    seed = 20240611
    packed = bsg.pack_histories(name, [[]]).copy()
    acts = []
    t = 0
    while True:
      cur, term, ret0 = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
      C.check(C.lib.bsg_host_state_info(gid, packed.ctypes.data, 1, 0, ctypes.byref(cur),
                                        ctypes.byref(term), ctypes.byref(ret0)))
      if term.value:
        break
      a = ctypes.c_int64()
      C.check(C.lib.bsg_host_replay_step(gid, packed.ctypes.data, 1, 0,
                                         bsg.draw(seed, playout, t), ctypes.byref(a)))
      acts.append(a.value)
      t += 1
    o_acts, o_term, o_ret = oracle.replay(oname, [], seed=seed, playout=playout, board_size=bs)
    assert acts == o_acts, (name, playout)
    assert o_term == 1 and ret0.value == o_ret
    if oname == "go" and len(acts) < 2 * bs * bs and not (acts[-1] == bs * bs and acts[-2] == bs * bs):
      superko += 1  # neither two passes nor the length cap: a position came back
  if oname == "go" and bs == 7:
    assert superko >= 1


def test_draw_matches_oracle(oracle):
  for seed, playout, t in [(0, 0, 0), (1, 2, 3), (2**63 + 5, 2**40 + 17, 77), (123, 456, 1000)]:
    assert bsg.draw(seed, playout, t) == oracle.draw(seed, playout, t)


def test_illegal_and_errors():
  g = bsg.load_game("tic_tac_toe")
  s = g.new_initial_state()
  s.apply_action(4)
  with pytest.raises(RuntimeError):
    s.apply_action(4)
  with pytest.raises(RuntimeError):
    s.apply_action(99)
  with pytest.raises(RuntimeError):
    bsg.load_game("chess")
  with pytest.raises(RuntimeError):
    bsg.load_game("hex(board_size=7)")
  with pytest.raises(RuntimeError):
    bsg.pack_histories("connect_four", [[0, 0, 0, 0, 0, 0, 0]])  # column overflow
  with pytest.raises(RuntimeError):
    s.observation_tensor(2)  # invalid player (api_test.py:314 expects an error)
  assert s.legal_actions(0) == [] and len(s.legal_actions(1)) == 8  # empty for the non-mover


def test_connect_four_known_answers():
  with open(os.path.join(GOLD, "known_answers.json")) as f:
    k = json.load(f)
  g = bsg.load_game("connect_four")
  s = g.new_initial_state()
  for a in k["connect_four_fast_loss"]["history"][:-1]:
    s.apply_action(a)
  assert not s.is_terminal()
  s.apply_action(k["connect_four_fast_loss"]["history"][-1])
  assert s.is_terminal() and s.returns() == [1.0, -1.0] and s.current_player() == -4


@pytest.mark.parametrize("name,oname,bs,n", [("hex", "hex", 11, 120), ("hex(board_size=5)", "hex", 5, 300),
                                             ("y", "y", 11, 60), ("y(board_size=9)", "y", 9, 300)])
def test_value_only_fill_gives_the_exact_playouts_winner(oracle, name, oname, bs, n):
  """bsg_playout_io.value_only: the kernels' step_fill + fill_winner (plain stones to a full
  board, then one reading of the winner) against the oracle's exact replay of the same playout
  id from the same mid-game root: same winner; the moves are the empty cells of the root."""
  gid = bsg.game_id(name)
  cells = {"hex": bs * bs, "y": (19 if bs == 11 else bs) * ((19 if bs == 11 else bs) + 1) // 2}[oname]
  for playout in range(n):
    root, _, _ = oracle.replay(oname, [], seed=5, playout=playout, max_steps=playout % 40, board_size=bs)
    if oracle.state_dump(oname, root, board_size=bs)["terminal"]:
      root = root[:-1]
    packed = bsg.pack_histories(name, [root])
    ret0, moves = ctypes.c_int(), ctypes.c_int()
    C.check(C.lib.bsg_host_fill_playout(gid, packed.ctypes.data, 1, 0, 77, playout,
                                        ctypes.byref(ret0), ctypes.byref(moves)))
    _, o_term, o_ret = oracle.replay(oname, root, seed=77, playout=playout, board_size=bs)
    assert o_term == 1 and ret0.value == o_ret, (name, playout)
    assert moves.value == cells - len(root)
  # a finished root: no move is made, the winner is the root's
  full, _, o_ret = oracle.replay(oname, [], seed=3, playout=1, board_size=bs)
  packed = bsg.pack_histories(name, [full])
  C.check(C.lib.bsg_host_fill_playout(gid, packed.ctypes.data, 1, 0, 1, 1, ctypes.byref(ret0), ctypes.byref(moves)))
  assert ret0.value == o_ret and moves.value == 0
  with pytest.raises(RuntimeError):
    C.check(C.lib.bsg_host_fill_playout(bsg.game_id("connect_four"), packed.ctypes.data, 1, 0, 1, 1, None, None))


def test_plane_table_comes_from_the_abi():
  """bsg_plane_bytes: the packed SoA plane sizes sum to the state size for every game id."""
  from open_spiel_1_b200 import engine as E
  for gid in range(13):
    planes = E.plane_bytes(gid)
    assert planes and sum(planes) == E.state_bytes(gid)
  assert E.plane_bytes(C.CONNECT_FOUR) == [8] and E.plane_bytes(C.BREAKTHROUGH) == [16, 4]


@pytest.mark.parametrize("name", ["connect_four", "breakthrough", "hex", "othello", "y"])
def test_pack_states_interleaves_mirror_states_like_pack_histories(name):
  """evaluator.pack_states: a batch of mirror states travels as their own bitboards (also
  when a state has no history: a connect_four / breakthrough root deserialised from a board
  string), and equals the batch packed from the histories."""
  from open_spiel_1_b200.evaluator import pack_states
  game = bsg.load_game(name)
  rng = np.random.RandomState(3)
  states = []
  for i in range(7):
    s = game.new_initial_state()
    for _ in range(3 * i):
      la = s.legal_actions()
      if not la:
        break
      s.apply_action(la[rng.randint(len(la))])
    states.append(s)
  want = bsg.pack_histories(name, [s.history() for s in states])
  assert np.array_equal(pack_states(game.gid, states), want)
  if name == "connect_four":
    # the same positions through the board-string reader: no history, same bitboards; moves
    # made after the deserialisation are the only history such a state has
    des = [game.deserialize_state(s.serialize()) for s in states if not s.is_terminal()]
    live = [s for s in states if not s.is_terminal()]
    assert all(d.history() == [] for d in des)
    assert np.array_equal(pack_states(game.gid, des), pack_states(game.gid, live))
    for d, s in zip(des, live):
      a = s.legal_actions()[0]
      d.apply_action(a); s.apply_action(a)
    assert np.array_equal(pack_states(game.gid, des), pack_states(game.gid, live))


def test_connect_four_reader_rejects_floating_stones():
  game = bsg.load_game("connect_four")
  ok = ".......\n.......\n.......\n.......\n...o...\n...x...\n"
  assert game.deserialize_state(ok).legal_actions() == list(range(7))
  floating = ".......\n.......\n.......\n...o...\n.......\n...x...\n"
  with pytest.raises(RuntimeError):
    game.deserialize_state(floating)


def test_alpha_zero_cache_survives_eviction_inside_a_batch():
  """AlphaZeroEvaluator.inference_batch with a cache smaller than the batch, and with a hit at
  the LRU end while misses are inserted (evaluator.py:56-75 of the reference is an LRU too)."""
  from open_spiel_1_b200 import evaluator as ev_mod

  class FakeEngine:
    gid = C.TIC_TAC_TOE
    state_bytes = 4

    def to_device(self, buf):
      return np.asarray(buf).copy()

    def observations(self, d_states, n, players, want_mask=False):
      import torch
      w = torch.from_numpy(d_states.view(np.uint32).astype(np.int64))
      return w.reshape(n, 1).double(), torch.ones((n, 9), dtype=torch.uint8)

  class Model:
    calls = 0

    def inference(self, obs, mask):
      import torch
      Model.calls += 1
      return obs / 1e6, torch.full((obs.shape[0], 9), 1.0 / 9, dtype=torch.float64)

  game = bsg.load_game("tic_tac_toe")
  ev = ev_mod.AlphaZeroEvaluator(game, Model(), cache_size=3, engine=FakeEngine())
  states = []
  for a in range(9):
    s = game.new_initial_state()
    s.apply_action(a)
    states.append(s)
  want = np.array([float(s.packed().view(np.uint32)[0]) / 1e6 for s in states])
  v, p = ev.inference_batch(states)                       # 9 misses through a cache of 3
  assert np.allclose(v, want) and p.shape == (9, 9)
  assert ev.cache_info()["size"] == 3
  # states[6] is now the LRU entry: a batch that hits it and brings three misses must not lose it
  v, _ = ev.inference_batch([states[0], states[6], states[1], states[2]])
  assert np.allclose(v, want[[0, 6, 1, 2]])
  assert ev.cache_info()["size"] == 3
  v, _ = ev.inference_batch([states[2], states[2]])       # duplicate keys, one hit each
  assert np.allclose(v, want[[2, 2]])


def test_compact_roots_round_trip_through_the_kernels_decode():
  """connect_four roots as 7 bytes each (bsg_compact_roots) expand to the very packed states
  they came from through ConnectFour::load_compact — the decode the playout kernel runs —
  including finished games (the win flag and the empty-cell count are derived, not stored) and
  a drawn full board; other games' "compact" format is the packed one."""
  game = bsg.load_game("connect_four")
  rng = np.random.RandomState(7)
  hs = []
  for i in range(3000):
    s = game.new_initial_state()
    for _ in range(rng.randint(0, 43)):
      la = s.legal_actions()
      if not la:
        break
      s.apply_action(la[rng.randint(len(la))])
    hs.append(s.history())
  # a drawn game (the reference's DeserializeDraw board, connect_four_test.cc:62-80, as moves)
  draw = "ooxxxoo\nxxoooxx\nooxxxoo\nxxoooxx\nooxxxoo\nxxoooxx\n"
  n = len(hs)
  packed = bsg.pack_histories("connect_four", hs)
  d = game.deserialize_state(draw)
  assert d.is_terminal() and d.returns() == [0.0, 0.0]
  packed = np.concatenate([packed, d.packed()])
  n += 1
  assert C.lib.bsg_compact_state_bits(C.CONNECT_FOUR) == 49
  assert C.lib.bsg_compact_buffer_bytes(C.CONNECT_FOUR, n) == (n * 49 + 7) // 8
  comp = np.full((n * 49 + 7) // 8, 0xA5, np.uint8)        # (the converter clears the buffer itself)
  C.check(C.lib.bsg_compact_roots(C.CONNECT_FOUR, packed.ctypes.data, n, comp.ctypes.data))
  back = np.zeros(n * 8, np.uint8)
  C.check(C.lib.bsg_expand_roots(C.CONNECT_FOUR, comp.ctypes.data, n, back.ctypes.data))
  assert np.array_equal(back, packed)
  # root i is the 49-bit string at bit 49 * i: seven rows of seven columns, bit 7 of every row byte dropped
  bits = np.unpackbits(comp, bitorder="little")
  for i in (0, 1, n // 2, n - 1):
    key = int.from_bytes(packed[8 * i:8 * i + 7].tobytes(), "little")
    want = sum(((key >> (8 * r)) & 0x7F) << (7 * r) for r in range(7))
    got = sum(int(b) << k for k, b in enumerate(bits[49 * i:49 * i + 49]))
    assert got == want
  assert C.lib.bsg_compact_state_bits(C.HEX) == 8 * 64 and C.lib.bsg_compact_state_bits(C.TIC_TAC_TOE) == 8 * 4
