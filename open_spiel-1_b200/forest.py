"""Root-parallel MCTS with the trees themselves on the GPU (csrc/bsg_forest.cuh).

`ForestMCTS` is `parallel.RootParallelMCTS(native=True, trees_per_rank=T)` with the T trees of
a rank resident in device memory: every round one kernel descends all trees, ONE launch of the
playout kernel evaluates all leaves of all trees, one kernel backs the results up; the per-step
merge of the root statistics is two small kernels around one int64 all-reduce.  The host does
no tree work at all, so the number of trees is set by what fills the GPU (thousands), not by
the host's core count.

Tree v of rank r is virtual rank r*T + v, exactly as in RootParallelMCTS: same shuffle seeds,
playout-id ranges and simulation shares, the same per-tree search (the reference's MCTSBot,
python/algorithms/mcts.py:292-443, visit for visit) and the same merge — the merged root equals
RootParallelMCTS's bit for bit (tests/test_gpu_forest.py).
"""
import ctypes

import numpy as np

from . import _cabi as C
from .mcts import SearchNode
from .parallel import _world, shard_range

# upper bounds on the legal actions of one state (children per expansion)
_MAX_CHILDREN = {C.TIC_TAC_TOE: 9, C.CONNECT_FOUR: 7, C.BREAKTHROUGH: 48, C.HEX: 121, C.PENTAGO: 288, C.OTHELLO: 34, C.Y: 190, C.HAVANNAH: 169, C.GO: 82}
_MERGE_FIELDS = 5


class ForestMCTS:
  """Args as parallel.RootParallelMCTS, plus:
    n_trees: trees resident on this rank's GPU.
    node_capacity: nodes per tree (default: enough for every simulation to expand a node).
  `evaluator` supplies the engine, n_rollouts, the Philox seed and the playout-id base."""

  def __init__(self, game, uct_c, max_simulations, evaluator, n_trees, sims_per_step=None,
               solve=True, seed=0, leaf_batch=32, device=None, group=None, node_capacity=None,
               merge="engine"):
    """merge = "engine" (default): the per-step root merge runs inside the library on the
    engine's stream — pack, one ncclAllReduce over the ENGINE'S OWN communicator, install — with
    no host synchronisation (bsg_forest_merge; the communicator is created here from a unique id
    broadcast over torch.distributed, any other transport works).  merge = "torch": the
    collective is torch.distributed.all_reduce around bsg_forest_pack_root / _install_root
    (same result; kept for hosts that want to run the collective themselves)."""
    self.rank, self.world = _world(group)
    self.group = group
    self.merge = merge
    if merge not in ("engine", "torch"):
      raise ValueError("merge must be 'engine' or 'torch'")
    self.device = device
    self.engine = evaluator.engine
    self.trees = int(n_trees)
    self.n_rollouts = int(evaluator.n_rollouts)
    self.total_simulations = int(max_simulations)
    vworld = self.world * self.trees
    shares = [shard_range(self.total_simulations, self.rank * self.trees + v, vworld)
              for v in range(self.trees)]
    self.tree_simulations = np.array([hi - lo for lo, hi in shares], np.int64)
    per_rank_max = -(-self.total_simulations // vworld)
    self.sims_per_step = int(sims_per_step or per_rank_max)
    self.n_steps = max(1, -(-per_rank_max // self.sims_per_step))
    self.num_actions = game.num_distinct_actions()
    self.leaf_batch = max(1, int(leaf_batch))
    per_tree = int(self.tree_simulations.max()) if self.trees else 0
    if node_capacity is None:
      node_capacity = (per_tree + 2) * _MAX_CHILDREN[self.engine.gid] + 8
    h = C.vp()
    vrank0 = self.rank * self.trees
    C.check(C.lib.bsg_forest_create(self.engine._h, self.trees, int(node_capacity),
                                    self.total_simulations + 8, float(uct_c), int(bool(solve)),
                                    self.n_rollouts, self.leaf_batch, int(evaluator.seed),
                                    int(seed) + vrank0, int(evaluator.next_playout) + (vrank0 << 40),
                                    ctypes.byref(h)), self.engine._h)
    self._h = h
    self.all_reduces = 0
    self.last_root = None
    if merge == "engine" and self.world > 1 and C.lib.bsg_comm_ranks(self.engine._h) != self.world:
      self._init_engine_comm()

  def _init_engine_comm(self):
    """ncclGetUniqueId on rank 0, the 128 bytes broadcast over the existing process group,
    ncclCommInitRank on every rank (bsg_comm_*)."""
    import torch
    import torch.distributed as dist
    uid = np.zeros(128, np.uint8)
    if self.rank == 0:
      C.check(C.lib.bsg_comm_unique_id(uid.ctypes.data), self.engine._h)
    backend = dist.get_backend(self.group)
    t = torch.from_numpy(uid)
    if backend == "nccl":
      t = t.to(self.engine.torch_device)
    src = dist.get_global_rank(self.group, 0) if self.group is not None else 0
    dist.broadcast(t, src, group=self.group)
    uid = t.cpu().numpy().copy()
    C.check(C.lib.bsg_comm_init(self.engine._h, self.world, self.rank, uid.ctypes.data), self.engine._h)

  def allreduce_stats(self):
    """{count, mean_us, max_us} of the all-reduces bsg_forest_merge issued since the last call
    (CUDA events on the engine's stream around each ncclAllReduce)."""
    out = np.zeros(3, np.float64)
    self._check(C.lib.bsg_forest_allreduce_stats(self._h, out.ctypes.data))
    return {"count": int(out[0]), "mean_us": float(out[1]), "max_us": float(out[2])}

  def close(self):
    if getattr(self, "_h", None):
      C.lib.bsg_forest_destroy(self._h)
      self._h = None

  def __del__(self):
    try:
      self.close()
    except Exception:
      pass

  def _check(self, rc):
    if rc < 0:
      raise RuntimeError("bsg: " + C.lib.bsg_forest_last_error(self._h).decode())
    return rc

  def info(self):
    """{game_steps, launches, rounds, trees} of this rank's forest so far."""
    out = np.zeros(4, np.float64)
    self._check(C.lib.bsg_forest_info(self._h, out.ctypes.data))
    return {"game_steps": int(out[0]), "launches": int(out[1]), "rounds": int(out[2]), "trees": int(out[3])}

  @property
  def total_steps(self):
    return self.info()["game_steps"]

  def tree_root(self, tree, player):
    """Root of one tree as a SearchNode with its children (statistics as the tree holds them)."""
    n = self.num_actions
    actions, visits = np.zeros(n, np.int64), np.zeros(n, np.int64)
    reward, outcome = np.zeros(n, np.float64), np.zeros(n, np.int32)
    info = np.zeros(4, np.float64)
    k = self._check(C.lib.bsg_forest_tree_root(self._h, int(tree), n, actions.ctypes.data,
                                               visits.ctypes.data, reward.ctypes.data,
                                               outcome.ctypes.data, info.ctypes.data))
    root = SearchNode(None, player, 1)
    root.explore_count = int(info[0])
    if info[1]:
      root.outcome = [float(info[2]), -float(info[2])]
    for i in range(k):
      ch = SearchNode(int(actions[i]), player, 1.0 / k)
      ch.explore_count = int(visits[i])
      ch.total_reward = float(reward[i])
      if outcome[i] != 2:
        v = float(outcome[i])
        ch.outcome = [v, -v] if player == 0 else [-v, v]
      root.children.append(ch)
    return root

  def mcts_search(self, state):
    """Runs the sharded search; returns tree 0's root with the GLOBAL child statistics (equal
    in every tree of every rank)."""
    from .evaluator import _packed_of
    packed = np.ascontiguousarray(_packed_of(self.engine.gid, state))
    left = self.tree_simulations.copy()
    if self.merge == "engine":
      for step in range(self.n_steps):
        ks = np.minimum(left, self.sims_per_step).astype(np.int64)
        left -= ks
        self._check(C.lib.bsg_forest_search(self._h, packed.ctypes.data, 0, ks.ctypes.data,
                                            1 if step else 0))
        self._check(C.lib.bsg_forest_merge(self._h, 0 if step else 1))   # stream-ordered, no sync
        self.all_reduces += 1
    else:
      import torch
      dev = self.engine.torch_device
      base = torch.zeros((self.num_actions, 2), dtype=torch.int64, device=dev)
      for step in range(self.n_steps):
        ks = np.minimum(left, self.sims_per_step).astype(np.int64)
        left -= ks
        self._check(C.lib.bsg_forest_search(self._h, packed.ctypes.data, 0, ks.ctypes.data,
                                            1 if step else 0))
        out = torch.zeros((self.num_actions, _MERGE_FIELDS), dtype=torch.int64, device=dev)
        torch.cuda.synchronize(dev)
        self._check(C.lib.bsg_forest_pack_root(self._h, base.data_ptr(), out.data_ptr()))
        if self.world > 1:
          import torch.distributed as dist
          dist.all_reduce(out, group=self.group)     # int64 sums: order-independent
        self.all_reduces += 1
        torch.cuda.synchronize(dev)
        self._check(C.lib.bsg_forest_install_root(self._h, base.data_ptr(), out.data_ptr()))
        base += out[:, :2]
      torch.cuda.synchronize(dev)
    self.last_root = self.tree_root(0, state.current_player())   # synchronises the engine's stream
    return self.last_root

  def step_with_policy(self, state):
    root = self.mcts_search(state)
    best = max(sorted(root.children, key=lambda c: c.action), key=SearchNode.sort_key)
    action = best.action
    policy = [(a, 1.0 if a == action else 0.0) for a in state.legal_actions(state.current_player())]
    return policy, action

  def step(self, state):
    return self.step_with_policy(state)[1]
