"""CPU test of the GPU-resident search forest's per-tree code (csrc/bsg_forest.cuh).

The functions the forest kernels run per tree are __host__ __device__; bsg_forest_selftest runs
them on host memory next to the host tree of bsg_mcts_* (the tree tests/test_gpu_mcts.py and
tests/test_reference_python.py tie to the reference's mcts.py), both fed the same stand-in leaf
values, and compares every node of every tree — after a first search with uneven budgets and
after a resumed one."""
import numpy as np
import pytest

import open_spiel_1_b200 as bsg
from open_spiel_1_b200 import _cabi as C


@pytest.mark.parametrize("game,history,sims,batch", [
    ("tic_tac_toe", [], 300, 1), ("tic_tac_toe", [4, 0], 400, 8),      # solver proofs reach the root
    ("connect_four", [3, 3, 2], 500, 16), ("connect_four", [], 200, 1),
    ("breakthrough", [], 150, 8), ("hex", [60, 48], 120, 4), ("pentago", [112, 3], 150, 8),
    ("othello", [19, 18], 200, 8), ("y", [40, 41], 120, 4),
    ("havannah", [112, 97], 150, 4), ("go(board_size=9)", [40, 41, 81], 120, 4)])
def test_forest_trees_equal_host_trees(game, history, sims, batch):
  gid = bsg.game_id(game)
  h = np.asarray(history, np.int64)
  for solve in (1, 0):
    rc = C.lib.bsg_forest_selftest(gid, 5, sims, sims // 3, batch, 8, solve, 77 + solve,
                                   h.ctypes.data if len(h) else None, len(h))
    assert rc == 0, "tree %d differs" % (rc - 1) if rc > 0 else "error %d" % rc
