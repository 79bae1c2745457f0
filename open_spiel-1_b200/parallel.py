"""Multi-GPU side of the rollout path: one process per GPU (torch.distributed).

Two things shard:

* Independent playouts (BASELINE configs 2-4).  `shard_range` gives rank r the contiguous
  slice [r*N/W, (r+1)*N/W) of the playout index range.  The Philox counter of a playout is its
  GLOBAL index, so per-playout results are identical for every GPU count; there is no
  data-path collective, only an optional final all-reduce of the four int64 totals
  (`ShardedPlayouts.run`).

* Root-parallel MCTS (BASELINE config 5).  Every rank searches the same root with its own
  tree, random stream and playout-id range; after every search step (a batch of simulations)
  ONE all-reduce(sum) of a fixed int64 buffer merges the root children's statistics
  (`explore_count`, `total_reward` — python/algorithms/mcts.py:116-122) into every rank's root,
  and the final move is `best_child` over the merged statistics (mcts.py:144-162).  Rewards of
  a rollout evaluator are multiples of 1/n_rollouts, so they travel as exact integers
  (reward * n_rollouts): the merge is order-independent and bit-reproducible.  Solver-proven
  outcomes (mcts.py:437) ride in the same buffer as per-value proof counts.

Backend: "nccl" on GPUs (the buffer lives on the rank's device; 40 B per root child, so the
collective is latency-bound and NVLink bandwidth is irrelevant), "gloo" in the CPU tests.
"""
import numpy as np

from .mcts import MCTSBot, NativeMCTSBot, SearchNode


def shard_range(n, rank, world):
  """Contiguous slice [lo, hi) of range(n) owned by `rank` out of `world` (balanced to ±1)."""
  if world < 1 or not 0 <= rank < world:
    raise ValueError("bad rank/world %r/%r" % (rank, world))
  return (n * rank) // world, (n * (rank + 1)) // world


def _dist():
  import torch.distributed as dist
  return dist


def _world(group=None):
  dist = _dist()
  if dist.is_available() and dist.is_initialized():
    return dist.get_rank(group), dist.get_world_size(group)
  return 0, 1


def all_reduce_int64(values, device=None, group=None, op="sum"):
  """One all-reduce of a small int64 vector; returns a numpy array.  World size 1 is a no-op
  (never a CPU detour for the playouts themselves).  `device` = torch device of the buffer
  (the rank's GPU under NCCL, None/cpu under gloo)."""
  values = np.asarray(values, np.int64)
  rank, world = _world(group)
  if world == 1:
    return values.copy()
  import torch
  dist = _dist()
  t = torch.from_numpy(values.copy())
  if device is not None:
    t = t.to(device)
  dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX, group=group)
  return t.cpu().numpy()


class ShardedPlayouts:
  """Runs this rank's shard of `n_roots` roots through an `engine.Engine` and (optionally)
  reduces the totals.  `roots_of(lo, hi)` must return the packed DEVICE states of roots
  lo..hi-1 (capacity hi-lo), so each rank only ever materialises its own shard."""

  def __init__(self, engine, group=None):
    self.engine = engine
    self.group = group
    self.rank, self.world = _world(group)

  def run(self, roots_of, n_roots, rollouts=1, seed=0, reduce=True):
    import torch
    lo, hi = shard_range(n_roots, self.rank, self.world)
    totals = torch.zeros(4, dtype=torch.int64, device=self.engine.torch_device)
    if hi > lo:
      self.engine.playouts(roots_of(lo, hi), hi - lo, rollouts, seed=seed,
                           first_playout=lo * rollouts, totals=totals)
    if reduce and self.world > 1:
      _dist().all_reduce(totals, group=self.group)  # int64 sum: order-independent
    return totals.cpu().numpy()


# Layout of the merge buffer, per root child (index = action id): visit delta, reward delta *
# scale, then how many ranks hold a solver proof "mover loses / draws / wins" for that child.
# Everything is summed, so ONE all-reduce carries statistics and proofs together.
_F_VISITS, _F_REWARD, _F_PROVEN, _F_FIELDS = 0, 1, 2, 5


class RootParallelMCTS:
  """Root-parallel MCTS over `world` ranks with one all-reduce per search step.

  Args:
    game: pyspiel.Game / spiel.Game.
    uct_c, evaluator, solve, child_selection_fn: as `MCTSBot` (mcts.py:199-262).
    max_simulations: total simulations per move, summed over all ranks.
    sims_per_step: simulations each rank runs between two all-reduces (the "search step").
    seed: base seed; rank r shuffles priors with RandomState(seed + r) and — through
      `evaluator.next_playout` when present — draws from its own playout-id range.
    reward_scale: integer that makes every evaluator return an exact integer when multiplied
      (n_rollouts for a rollout evaluator).
    leaf_batch: leaves gathered per kernel launch inside a rank (mcts.MCTSBot.leaf_batch).
    device: torch device of the collective buffer (None = CPU, for gloo).
    native: run each rank's tree in the C-ABI library (mcts.NativeMCTSBot) instead of the
      Python mirror; `evaluator` must then be a GpuRolloutEvaluator.
    trees_per_rank: (native only) K independent trees per rank, each walked by its own host
      thread with its own engine (streams, staging buffers) on the rank's GPU.  One host thread
      walking one tree keeps a B200 ~1 % busy, so this is how a rank
      fills its GPU.  Tree j of rank r is "virtual rank" r*K + j: it gets the seeds, playout-id
      range and simulation share that rank r*K + j would get in a job of world*K ranks with one
      tree each, and the per-step merge sums over all of them — so the merged root statistics
      depend only on world*K, not on how the trees are spread over processes.
    threads: walk the K trees concurrently (default) or one after the other (same result).
  """

  def __init__(self, game, uct_c, max_simulations, evaluator, sims_per_step=None, solve=True,
               seed=0, reward_scale=None, leaf_batch=1, child_selection_fn=SearchNode.uct_value,
               device=None, group=None, native=False, trees_per_rank=1, threads=True):
    self.rank, self.world = _world(group)
    self.group = group
    self.device = device
    self.total_simulations = int(max_simulations)
    self.trees = int(trees_per_rank)
    if self.trees < 1 or (self.trees > 1 and not native):
      raise ValueError("trees_per_rank > 1 needs native=True")
    vworld = self.world * self.trees
    vranks = [self.rank * self.trees + j for j in range(self.trees)]
    shares = [shard_range(self.total_simulations, v, vworld) for v in vranks]
    self.tree_simulations = [hi - lo for lo, hi in shares]
    self.local_simulations = sum(self.tree_simulations)
    # every rank must take the same number of steps (the all-reduce is collective)
    per_rank_max = -(-self.total_simulations // vworld)
    self.sims_per_step = int(sims_per_step or per_rank_max)
    self.n_steps = max(1, -(-per_rank_max // self.sims_per_step))
    if reward_scale is None:
      reward_scale = getattr(evaluator, "n_rollouts", 1)
    self.reward_scale = int(reward_scale)
    self.num_actions = game.num_distinct_actions()
    if native:
      # the tree in the C-ABI library (csrc/bsg_mcts.cuh); `evaluator` supplies engine,
      # n_rollouts and seed, its rollouts run inside the native search
      from .engine import Engine
      self.bots = []
      for j, v in enumerate(vranks):
        eng = evaluator.engine if j == 0 else Engine(evaluator.engine.gid, evaluator.engine.device)
        bot = NativeMCTSBot(game, uct_c, self.tree_simulations[j], eng,
                            n_rollouts=evaluator.n_rollouts, solve=solve, seed=evaluator.seed,
                            shuffle_seed=seed + v, leaf_batch=leaf_batch)
        bot.next_playout = evaluator.next_playout + (v << 40)
        self.bots.append(bot)
    else:
      self.bots = [MCTSBot(game, uct_c, self.local_simulations, evaluator, solve=solve,
                           random_state=np.random.RandomState(seed + self.rank),
                           child_selection_fn=child_selection_fn, leaf_batch=leaf_batch)]
      # disjoint playout-id ranges per rank keep the ranks' Philox streams independent
      if hasattr(evaluator, "next_playout"):
        evaluator.next_playout += self.rank << 40
    self.bot = self.bots[0]
    self._pool = None
    if self.trees > 1 and threads:
      from concurrent.futures import ThreadPoolExecutor
      self._pool = ThreadPoolExecutor(max_workers=self.trees)   # ctypes calls drop the GIL
    self.all_reduces = 0
    self.last_root = None

  @property
  def total_steps(self):
    """Game steps played by this rank's rollouts (native trees)."""
    return sum(b.total_steps for b in self.bots)

  # ---- merge
  def _pack(self, root, base):
    """Deltas of this rank's root-child statistics since `base` + proven outcomes."""
    buf = np.zeros((self.num_actions, _F_FIELDS), np.int64)
    for ch in root.children:
      b = base.get(ch.action, (0, 0))
      scaled = int(round(ch.total_reward * self.reward_scale))
      buf[ch.action, _F_VISITS] = ch.explore_count - b[0]
      buf[ch.action, _F_REWARD] = scaled - b[1]
      if ch.outcome is not None:  # proven value for the mover: -1 / 0 / +1 -> slot 0 / 1 / 2
        buf[ch.action, _F_PROVEN + int(round(ch.outcome[ch.player])) + 1] = 1
    return buf

  def _merge(self, roots, base):
    """One all-reduce: sums the visit/reward deltas and proof counts of all trees of all ranks
    and installs the global statistics into the root children of every local tree."""
    mine = self._pack(roots[0], base)
    for root in roots[1:]:
      mine += self._pack(root, base)
    summed = all_reduce_int64(mine.reshape(-1), self.device, self.group)
    summed = summed.reshape(self.num_actions, _F_FIELDS)
    self.all_reduces += 1
    new_base = {}
    for root in roots:
      visits = 0
      for ch in root.children:
        b = base.get(ch.action, (0, 0))
        n = b[0] + int(summed[ch.action, 0])
        r = b[1] + int(summed[ch.action, 1])
        ch.explore_count = n
        ch.total_reward = r / self.reward_scale
        new_base[ch.action] = (n, r)
        visits += n
        proofs = summed[ch.action, _F_PROVEN:_F_PROVEN + 3]
        if ch.outcome is None and proofs.any():  # exact game values: trees cannot disagree
          v = float(int(np.argmax(proofs)) - 1)
          ch.outcome = [v, -v] if ch.player == 0 else [-v, v]
      if root.children:
        root.explore_count = max(root.explore_count, visits)
    return new_base

  # ---- search
  def _search_tree(self, j, state, root, k):
    bot = self.bots[j]
    root = bot.mcts_search(state, root=root, simulations=k)
    if not root.children and not state.is_terminal():
      bot.expand_root(root, state)  # a tree with 0 simulations still joins the merge
    return root

  def mcts_search(self, state):
    """Runs the sharded search; returns this rank's root with GLOBAL child statistics (equal
    on every rank for visits/rewards)."""
    n_trees = len(self.bots)
    roots = [None] * n_trees
    base = {}
    left = list(self.tree_simulations) if n_trees > 1 else [self.local_simulations]
    for _ in range(self.n_steps):
      ks = [min(self.sims_per_step, l) for l in left]
      left = [l - k for l, k in zip(left, ks)]
      if self._pool is not None:
        futs = [self._pool.submit(self._search_tree, j, state, roots[j], ks[j])
                for j in range(n_trees)]
        roots = [f.result() for f in futs]
      else:
        roots = [self._search_tree(j, state, roots[j], ks[j]) for j in range(n_trees)]
      base = self._merge(roots, base)
    self.last_root = roots[0]
    return roots[0]

  def step_with_policy(self, state):
    root = self.mcts_search(state)
    # identical on every rank: ties in (proven, visits, reward) fall to the lowest action id
    best = max(sorted(root.children, key=lambda c: c.action), key=SearchNode.sort_key)
    action = best.action
    policy = [(a, 1.0 if a == action else 0.0)
              for a in state.legal_actions(state.current_player())]
    return policy, action

  def step(self, state):
    return self.step_with_policy(state)[1]
