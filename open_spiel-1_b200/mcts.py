"""Host-side MCTS driver with the reference's interface, for use where pyspiel is absent.

Same names, argument meaning and results as `open_spiel.python.algorithms.mcts`
(mcts.py:85-443): `SearchNode` (uct_value / puct_value / sort_key / best_child) and
`MCTSBot(game, uct_c, max_simulations, evaluator, solve, random_state, child_selection_fn,
dirichlet_noise, verbose)` with `step`, `step_with_policy`, `mcts_search`.  The tree stays on
the host exactly as in the reference; what changes is the evaluator behind it
(`evaluator.GpuRolloutEvaluator`), plus an optional batched-leaf mode that gathers several
leaves per kernel launch.  The reference's own MCTSBot can be used
instead, unmodified, with the same evaluator object.
"""
import math
import time

import numpy as np

from .spiel import GameType, PlayerId


class SearchNode(object):
  """One edge/state of the search tree (fields as mcts.py:85-122)."""
  __slots__ = ["action", "player", "prior", "explore_count", "total_reward", "outcome",
               "children", "virtual"]

  def __init__(self, action, player, prior):
    self.action = action
    self.player = player
    self.prior = prior
    self.explore_count = 0
    self.total_reward = 0.0
    self.outcome = None
    self.children = []
    self.virtual = 0  # virtual visits of in-flight batched simulations

  def uct_value(self, parent_explore_count, uct_c):
    """Q/N + c * sqrt(ln(parent N) / N); proven nodes return their outcome (mcts.py:124-133)."""
    if self.outcome is not None:
      return self.outcome[self.player]
    if self.explore_count == 0:
      return float("inf")
    exploit = self.total_reward / self.explore_count
    return exploit + uct_c * math.sqrt(math.log(parent_explore_count) / self.explore_count)

  def puct_value(self, parent_explore_count, uct_c):
    """Q/N + c * prior * sqrt(parent N) / (N + 1) (mcts.py:135-142)."""
    if self.outcome is not None:
      return self.outcome[self.player]
    exploit = self.total_reward / self.explore_count if self.explore_count else 0
    return exploit + uct_c * self.prior * math.sqrt(parent_explore_count) / (self.explore_count + 1)

  def sort_key(self):
    """(proven value, visits, reward): the final-move ordering of mcts.py:144-158."""
    proven = 0 if self.outcome is None else self.outcome[self.player]
    return (proven, self.explore_count, self.total_reward)

  def best_child(self):
    return max(self.children, key=SearchNode.sort_key)

  def children_str(self, state=None):
    ordered = sorted(self.children, key=SearchNode.sort_key, reverse=True)
    return "\n".join(c.to_str(state) for c in ordered)

  def to_str(self, state=None):
    if state is not None and self.action is not None:
      name = state.action_to_string(state.current_player(), self.action)
    else:
      name = str(self.action)
    value = self.total_reward / self.explore_count if self.explore_count else 0
    outcome = "{:4.1f}".format(self.outcome[self.player]) if self.outcome else "none"
    return ("{:>6}: player: {}, prior: {:5.3f}, value: {:6.3f}, sims: {:5d}, outcome: {}, "
            "{:3d} children").format(name, self.player, self.prior, value, self.explore_count,
                                     outcome, len(self.children))

  def __str__(self):
    return self.to_str(None)


class PhiloxShuffle(object):
  """`random_state` for MCTSBot whose shuffle() is the Fisher-Yates of the native tree
  (csrc/bsg_mcts.cuh): j = bsg_draw(seed, counter, 0) mod (i + 1) for i = n-1 .. 1.  With it the
  Python mirror and `NativeMCTSBot` build identical trees (the mirror is the specification)."""

  def __init__(self, seed):
    self.seed = int(seed)
    self.counter = 0

  def shuffle(self, items):
    from .engine import draw
    for i in range(len(items) - 1, 0, -1):
      j = draw(self.seed, self.counter, 0) % (i + 1)
      self.counter += 1
      items[i], items[j] = items[j], items[i]

  def dirichlet(self, alpha):
    raise NotImplementedError("PhiloxShuffle has no Dirichlet noise")


class MCTSBot(object):
  """UCT / PUCT search with MCTS-Solver backups (mcts.py:199-443)."""

  def __init__(self, game, uct_c, max_simulations, evaluator, solve=True, random_state=None,
               child_selection_fn=SearchNode.uct_value, dirichlet_noise=None, verbose=False,
               leaf_batch=1):
    game_type = game.get_type()
    if game_type.reward_model.name != GameType.RewardModel.TERMINAL.name:
      raise ValueError("Game must have terminal rewards.")
    if game_type.dynamics.name != GameType.Dynamics.SEQUENTIAL.name:
      raise ValueError("Game must have sequential turns.")
    self._game = game
    self.uct_c = uct_c
    self.max_simulations = max_simulations
    self.evaluator = evaluator
    self.verbose = verbose
    self.solve = solve
    self.max_utility = game.max_utility()
    self._dirichlet_noise = dirichlet_noise
    self._random_state = random_state or np.random.RandomState()
    self._child_selection_fn = child_selection_fn
    self.leaf_batch = max(1, int(leaf_batch))

  def restart_at(self, state):
    pass

  def inform_action(self, state, player_id, action):
    pass  # pyspiel.Bot's default: the search keeps no state between moves

  def step_with_policy(self, state):
    t0 = time.time()
    root = self.mcts_search(state)
    best = root.best_child()
    if self.verbose:
      secs = max(time.time() - t0, 1e-9)
      print("Finished {} sims in {:.3f} secs, {:.1f} sims/s".format(
          root.explore_count, secs, root.explore_count / secs))
      print("Root:\n" + root.to_str(state))
      print("Children:\n" + root.children_str(state))
    action = best.action
    policy = [(a, 1.0 if a == action else 0.0)
              for a in state.legal_actions(state.current_player())]
    return policy, action

  def step(self, state):
    return self.step_with_policy(state)[1]

  # ---- descent
  def _expand(self, node, state, is_root):
    priors = self.evaluator.prior(state)
    if is_root and self._dirichlet_noise:
      eps, alpha = self._dirichlet_noise
      noise = self._random_state.dirichlet([alpha] * len(priors))
      priors = [(a, (1 - eps) * p + eps * x) for (a, p), x in zip(priors, noise)]
    self._random_state.shuffle(priors)  # remove the move-generation order bias (mcts.py:320)
    player = state.current_player()
    node.children = [SearchNode(a, player, p) for a, p in priors]

  def _apply_tree_policy(self, root, state):
    """Descends to a leaf (terminal or never-evaluated node); returns (path, leaf state)."""
    path = [root]
    working = state.clone()
    node = root
    while not working.is_terminal() and (node.explore_count + node.virtual) > 0:
      if not node.children:
        self._expand(node, working, node is root)
      parent_n = node.explore_count + node.virtual
      fn, c = self._child_selection_fn, self.uct_c
      if self.leaf_batch > 1:
        # virtual visits steer concurrent descents apart; values are unchanged when virtual=0
        best, best_v = None, -float("inf")
        for ch in node.children:
          if ch.virtual and ch.outcome is None:
            n = ch.explore_count + ch.virtual
            v = (ch.total_reward - ch.virtual) / n + c * math.sqrt(math.log(max(parent_n, 2)) / n)
          else:
            v = fn(ch, max(parent_n, 1), c)
          if v > best_v:
            best, best_v = ch, v
        chosen = best
      else:
        chosen = max(node.children, key=lambda ch: fn(ch, parent_n, c))
      working.apply_action(chosen.action)
      node = chosen
      path.append(node)
    return path, working

  # ---- backup (MCTS-Solver, max^n for two players; mcts.py:408-439)
  def _backup(self, path, returns, solved, root_player):
    for node in reversed(path):
      who = root_player if node.player == PlayerId.CHANCE else node.player
      node.total_reward += returns[who]
      node.explore_count += 1
      if solved and node.children:
        mover = node.children[0].player
        best = None
        all_solved = True
        for child in node.children:
          if child.outcome is None:
            all_solved = False
          elif best is None or child.outcome[mover] > best.outcome[mover]:
            best = child
        if best is not None and (all_solved or best.outcome[mover] == self.max_utility):
          node.outcome = best.outcome
        else:
          solved = False

  def expand_root(self, root, state):
    """Creates the root's children without running a simulation (root-parallel merge)."""
    if not root.children and not state.is_terminal():
      self._expand(root, state, True)

  def mcts_search(self, state, root=None, simulations=None):
    """Runs max_simulations simulations from `state`; returns the root SearchNode.
    `root` resumes an earlier search of the same state and `simulations` overrides the
    budget of this call (both used by parallel.RootParallelMCTS between merges)."""
    root_player = state.current_player()
    if root is None:
      root = SearchNode(None, root_player, 1)
    budget = self.max_simulations if simulations is None else int(simulations)
    done = 0
    while done < budget:
      if self.leaf_batch == 1:
        path, leaf = self._apply_tree_policy(root, state)
        if leaf.is_terminal():
          returns = leaf.returns()
          path[-1].outcome = returns
          solved = self.solve
        else:
          returns = self.evaluator.evaluate(leaf)
          solved = False
        self._backup(path, returns, solved, root_player)
        done += 1
      else:
        done += self._batched_round(root, state, root_player, budget - done)
      if root.outcome is not None:
        break
    return root

  def _batched_round(self, root, state, root_player, budget):
    """Collects up to leaf_batch leaves (virtual visits keep them apart), evaluates all the
    non-terminal ones in ONE kernel launch, then backs everything up."""
    pending = []
    k = min(self.leaf_batch, budget)
    for _ in range(k):
      path, leaf = self._apply_tree_policy(root, state)
      for node in path:
        node.virtual += 1
      pending.append((path, leaf))
      if leaf.is_terminal() and self.solve:
        break  # a proven leaf may solve the root; resolve it before descending again
    leaves = [leaf for _, leaf in pending if not leaf.is_terminal()]
    values = self.evaluator.evaluate_batch(leaves) if leaves else []
    vi = 0
    for path, leaf in pending:
      for node in path:
        node.virtual -= 1
      if leaf.is_terminal():
        returns = leaf.returns()
        path[-1].outcome = returns
        self._backup(path, returns, self.solve, root_player)
      else:
        self._backup(path, values[vi], False, root_player)
        vi += 1
    return len(pending)


class NativeMCTSBot(object):
  """MCTSBot whose tree lives in the C-ABI library (csrc/bsg_mcts.cuh): same search as
  `MCTSBot(game, uct_c, max_simulations, GpuRolloutEvaluator(engine, n_rollouts, seed),
  solve, random_state=PhiloxShuffle(shuffle_seed), leaf_batch=leaf_batch)` — identical tree,
  visit for visit — without the Python interpreter between two leaves.  The evaluator is the
  engine's rollout evaluator by construction (its rollouts run on the GPU, up to `leaf_batch`
  leaves per kernel launch); child selection is UCT (mcts.py:124-133).

  mcts_search returns a `SearchNode` root whose children carry the tree's statistics; a root
  passed back in (`root=`, as parallel.RootParallelMCTS does after merging) has its children's
  statistics written into the native tree before the search resumes."""

  def __init__(self, game, uct_c, max_simulations, engine, n_rollouts=1, solve=True, seed=0,
               shuffle_seed=0, leaf_batch=1, verbose=False):
    import ctypes
    from . import _cabi as C
    self._C, self._ct = C, ctypes
    self._game = game
    self.engine = engine
    self.uct_c = float(uct_c)
    self.max_simulations = int(max_simulations)
    self.n_rollouts = int(n_rollouts)
    self.solve = bool(solve)
    self.leaf_batch = max(1, int(leaf_batch))
    self.verbose = verbose
    h = ctypes.c_void_p()
    C.check(C.lib.bsg_mcts_create(engine._h, self.uct_c, int(self.solve), self.n_rollouts,
                                  self.leaf_batch, int(seed), int(shuffle_seed), ctypes.byref(h)),
            engine._h)
    self._h = h
    self._cap = game.num_distinct_actions()
    self._last_root = None

  def close(self):
    if getattr(self, "_h", None):
      self._C.lib.bsg_mcts_destroy(self._h)
      self._h = None

  def __del__(self):
    try:
      self.close()
    except Exception:
      pass

  # evaluator-compatible bookkeeping (parallel.RootParallelMCTS offsets the playout ids per rank)
  @property
  def next_playout(self):
    return int(self.info()[6])

  @next_playout.setter
  def next_playout(self, value):
    self._C.check(self._C.lib.bsg_mcts_set_next_playout(self._h, int(value)))

  @property
  def total_steps(self):
    return int(self.info()[4])

  def info(self):
    out = np.zeros(8, np.float64)
    self._C.check(self._C.lib.bsg_mcts_info(self._h, out.ctypes.data))
    return out

  def restart_at(self, state):
    pass

  def inform_action(self, state, player_id, action):
    pass

  def _packed(self, state):
    from .evaluator import _packed_of
    return np.ascontiguousarray(_packed_of(self.engine.gid, state))

  def _read_root(self, state):
    C = self._C
    n = self._cap
    actions, visits = np.zeros(n, np.int64), np.zeros(n, np.int64)
    reward, outcome = np.zeros(n, np.float64), np.zeros(n, np.int32)
    k = C.check(C.lib.bsg_mcts_root_children(self._h, n, actions.ctypes.data, visits.ctypes.data,
                                              reward.ctypes.data, outcome.ctypes.data))
    info = self.info()
    player = int(info[7])
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
    root.total_reward = float(sum(c.total_reward for c in root.children))
    return root

  def _write_root(self, root):
    C = self._C
    for ch in root.children:
      oc = 2 if ch.outcome is None else int(round(ch.outcome[ch.player]))
      C.check(C.lib.bsg_mcts_set_root_child(self._h, int(ch.action), int(ch.explore_count),
                                            float(ch.total_reward), oc))
    C.check(C.lib.bsg_mcts_set_root_explore(self._h, int(root.explore_count)))

  def expand_root(self, root, state):
    if not root.children and not state.is_terminal():
      packed = self._packed(state)
      self._C.check(self._C.lib.bsg_mcts_expand_root(self._h, packed.ctypes.data), self.engine._h)
      fresh = self._read_root(state)
      root.children = fresh.children

  def mcts_search(self, state, root=None, simulations=None):
    C = self._C
    budget = self.max_simulations if simulations is None else int(simulations)
    packed = self._packed(state)
    resume = 0
    if root is not None and root is self._last_root:
      self._write_root(root)
      resume = 1
    rc = C.lib.bsg_mcts_search(self._h, packed.ctypes.data, budget, resume)
    if rc < 0:
      raise RuntimeError("bsg: " + C.lib.bsg_mcts_last_error(self._h).decode())
    new_root = self._read_root(state)
    if root is not None and resume:
      # keep the caller's object alive: copy the fresh statistics into it
      root.explore_count, root.total_reward = new_root.explore_count, new_root.total_reward
      root.outcome, root.children = new_root.outcome, new_root.children
      new_root = root
    self._last_root = new_root
    return new_root

  def step_with_policy(self, state):
    t0 = time.time()
    root = self.mcts_search(state)
    best = root.best_child()
    if self.verbose:
      secs = max(time.time() - t0, 1e-9)
      print("Finished {} sims in {:.3f} secs, {:.1f} sims/s".format(
          root.explore_count, secs, root.explore_count / secs))
    action = best.action
    policy = [(a, 1.0 if a == action else 0.0)
              for a in state.legal_actions(state.current_player())]
    return policy, action

  def step(self, state):
    return self.step_with_policy(state)[1]
