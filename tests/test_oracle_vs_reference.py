"""Cross-checks the C oracle against the UNMODIFIED reference C++ (oracle/_ref/libspiel_ref.so).

Skipped where the reference library was not built (it needs /root/reference at build time).
"""
import os

import numpy as np
import pytest

from oracle_lib import Oracle, Reference, reference_path

pytestmark = pytest.mark.skipif(not os.path.exists(reference_path()),
                                reason="reference library not built")

CASES = [("tic_tac_toe", 11, 400), ("connect_four", 11, 400), ("breakthrough", 11, 150),
         ("hex", 11, 60), ("hex", 5, 200), ("pentago", 11, 300), ("othello", 11, 150),
         ("y", 11, 60), ("y", 9, 200), ("havannah", 11, 150), ("havannah", 4, 400),
         ("go", 9, 200), ("go", 7, 400)]   # y: 11 stands for the default size 19


@pytest.fixture(scope="module")
def libs():
  return Oracle(), Reference()


@pytest.mark.parametrize("game,bs,n", CASES)
def test_replay_and_dumps_match(libs, game, bs, n):
  orc, ref = libs
  rng = np.random.RandomState(123)
  for playout in range(n):
    a_ref, t_ref, r_ref = ref.replay(game, [], seed=99, playout=playout, board_size=bs)
    a_orc, t_orc, r_orc = orc.replay(game, [], seed=99, playout=playout, board_size=bs)
    assert a_ref == a_orc and t_ref == t_orc and r_ref == r_orc
    if playout % 10 == 0:
      for cut in {0, len(a_ref) // 2, max(len(a_ref) - 1, 0), len(a_ref)}:
        for player in (0, 1):
          d_ref = ref.state_dump(game, a_ref[:cut], player, board_size=bs)
          d_orc = orc.state_dump(game, a_ref[:cut], player, board_size=bs)
          assert d_ref["legal"] == d_orc["legal"]
          assert d_ref["current_player"] == d_orc["current_player"]
          assert d_ref["terminal"] == d_orc["terminal"]
          assert d_ref["returns"] == d_orc["returns"]
          assert np.array_equal(d_ref["obs"], d_orc["obs"])
    # explicit r table, from a mid-game root
    root = a_ref[:rng.randint(0, len(a_ref))]
    table = rng.randint(0, 2**32, size=256, dtype=np.uint64).astype(np.uint32)
    assert ref.replay(game, root, r_table=table, board_size=bs) == \
        orc.replay(game, root, r_table=table, board_size=bs)


def test_batch_matches(libs):
  orc, ref = libs
  for game in ("connect_four", "breakthrough"):
    roots = []
    for p in range(50):
      acts, _, _ = orc.replay(game, [], seed=5, playout=p, max_steps=p % 12)
      roots.append(acts)
    o = orc.replay_batch(game, roots, rollouts=3, seed=11, actions_stride=64)
    r = ref.replay_batch(game, roots, rollouts=3, seed=11, actions_stride=64, threads=4)
    for x, y in zip(o, r):
      assert np.array_equal(x, y)


def _log_of(histories, stride):
  log = np.full((len(histories), stride), -1, np.int16)
  lens = np.zeros(len(histories), np.int32)
  for i, h in enumerate(histories):
    log[i, :len(h)] = h
    lens[i] = len(h)
  return log, lens


@pytest.mark.parametrize("game", ["tic_tac_toe", "connect_four", "breakthrough", "hex", "pentago", "othello", "y", "havannah", "go(board_size=9)"])
def test_dense_log_entry_points_equal_the_per_state_ones(libs, game):
  """ref_replay_log / ref_state_dump_log (the BASELINE-size checker calls of the GPU tests) =
  ref_replay_batch / ref_state_dump on the same roots, terminal states included, and the oracle
  taking the same roots through ragged_from_log."""
  from oracle_lib import ragged_from_log
  orc, ref = libs
  roots = []
  for p in range(120):
    acts, _, _ = orc.replay(game, [], seed=6, playout=p, max_steps=(5 * p) % 37)
    roots.append(acts)                       # some of them are terminal (short games)
  full, _, _ = orc.replay(game, [], seed=6, playout=999)
  roots.append(full)                         # one certainly terminal root
  log, lens = _log_of(roots, 256)
  for R in (1, 3):
    a = ref.replay_log(game, log, lens, rollouts=R, seed=17, first_playout=40, threads=3)
    b = ref.replay_batch(game, roots, rollouts=R, seed=17, first_playout=40, threads=2)
    c = orc.replay_batch(game, ragged_from_log(log, lens), rollouts=R, seed=17, first_playout=40)
    for x, y, z in zip(a, b[:3], c[:3]):
      assert np.array_equal(x, y) and np.array_equal(x, z)
  players = (np.arange(len(roots)) % 2).astype(np.uint8)
  d = ref.state_dump_log(game, log, lens, players, threads=3)
  for i, h in enumerate(roots):
    one = ref.state_dump(game, h, int(players[i]))
    assert np.array_equal(d["obs"][i], one["obs"])
    assert np.nonzero(d["mask"][i])[0].tolist() == one["legal"]
    assert d["current_player"][i] == one["current_player"] and bool(d["terminal"][i]) == one["terminal"]
  assert d["terminal"][-1] == 1 and d["mask"][-1].sum() == 0
