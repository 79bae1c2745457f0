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
