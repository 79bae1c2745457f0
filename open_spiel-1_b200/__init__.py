"""B200-native batched game-simulation engine for OpenSpiel's rollout hot path.

Import as `open_spiel_1_b200` (the repo-root shim of that name loads this directory, whose
on-disk name `open_spiel-1_b200` is not a Python identifier).

  engine.Engine                  device side of the C ABI (include/bsg.h): playouts, observations
  evaluator.GpuRolloutEvaluator  drop-in for mcts.RandomRolloutEvaluator (evaluate / prior)
  spiel.load_game / Game / State host mirror of the pyspiel surface the path touches
  mcts.MCTSBot / SearchNode      host tree with the reference's interface (+ batched leaves)
  parallel.RootParallelMCTS      root-parallel search, one NCCL all-reduce per search step
  evaluate_bots.evaluate_bots    the reference's bot-vs-bot loop (+ play_match over many roots)
  replay.DeviceReplayBuffer      the AlphaZero learner's replay buffer as device tensors
  forest.ForestMCTS              root-parallel MCTS with the trees resident on the GPU
  actor.ForestActor              AlphaZero-style self-play, many games in flight, records on the device

The CUDA extension (libbsg.so) is mandatory: importing this package without it raises.
"""
from . import _cabi  # noqa: F401  (raises ImportError when libbsg.so is missing)
from .engine import Engine, game_id, pack_histories, draw
from .evaluator import Evaluator, GpuRolloutEvaluator, AlphaZeroEvaluator
from .spiel import load_game, Game, State, GameType, PlayerId
from .mcts import MCTSBot, SearchNode, NativeMCTSBot, PhiloxShuffle
from .parallel import RootParallelMCTS, ShardedPlayouts, shard_range
from .evaluate_bots import evaluate_bots, play_match, UniformRandomBot
from .replay import DeviceReplayBuffer, TrainInput
from .actor import ForestActor

__all__ = ["Engine", "game_id", "pack_histories", "draw", "Evaluator", "GpuRolloutEvaluator", "AlphaZeroEvaluator",
           "load_game", "Game", "State", "GameType", "PlayerId", "MCTSBot", "SearchNode", "NativeMCTSBot", "PhiloxShuffle",
           "RootParallelMCTS", "ShardedPlayouts", "shard_range", "evaluate_bots", "play_match", "UniformRandomBot", "DeviceReplayBuffer", "TrainInput", "ForestActor"]
