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
         ("y", 11, 60), ("y", 9, 200)]   # y: 11 stands for the default size 19


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
