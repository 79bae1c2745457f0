#!/usr/bin/env python
"""bench_mcts.py — secondary benchmark: the rollout evaluator behind MCTS (BASELINE configs 1, 5).

Not the driver's headline (that is bench.py); it measures how the drop-in evaluator behaves
behind the search loop the reference calls it from (e):

  cfg1  tic_tac_toe, 1000 simulations/move, n_rollouts = 1, uct_c = 2 (the defaults of
        python/examples/mcts.py:64-66): self-play games, simulations/s
          * reference C++ MCTSBot + RandomRolloutEvaluator (algorithms/mcts.cc), one bot per
            host thread (oracle/_ref/libspiel_ref.so);
          * this repo: MCTSBot mirror + GpuRolloutEvaluator, one leaf per kernel launch
            (leaf_batch = 1, exactly the reference's call pattern) and with batched leaves;
          * this repo: the native host tree (NativeMCTSBot, csrc/bsg_mcts.cuh), same search.
  cfg5  connect_four, 256 rollouts/leaf, root-parallel over the ranks of the job with ONE
        int64 all-reduce per search step (parallel.RootParallelMCTS): simulations/s and
        playouts/s of one move search from the empty board, whole job.

  python bench_mcts.py [--cfg 1|5|all] [--sims N] [--leaf-batch B] [--games G]
  torchrun --nproc-per-node N bench_mcts.py --cfg 5      (one rank per GPU, NCCL)

Rank 0 prints one JSON line per configuration.  Search results are checked inside the run:
cfg5 asserts that all ranks hold identical merged root statistics.
"""
import argparse
import json
import os
import sys
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

import numpy as np  # noqa: E402


def dist_env():
  return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
          int(os.environ.get("LOCAL_RANK", "0")))


def selfplay(bsg, game_name, bot_of, n_games):
  """Plays n_games of self-play with one bot for both sides; returns (moves, sims, secs)."""
  game = bsg.load_game(game_name)
  moves = sims = 0
  t0 = time.perf_counter()
  for _ in range(n_games):
    bot = bot_of(game)
    state = game.new_initial_state()
    while not state.is_terminal():
      root = bot.mcts_search(state)
      sims += root.explore_count
      state.apply_action(root.best_child().action)
      moves += 1
  return moves, sims, time.perf_counter() - t0


def run_cfg1(args, bsg, local):
  from oracle_lib import Reference, reference_path
  out = {"config": "cfg1: tic_tac_toe MCTSBot + RandomRolloutEvaluator, %d sims/move, n_rollouts=1, "
                   "uct_c=2, self-play from the initial state" % args.sims,
         "unit": "simulations/s"}
  if os.path.exists(reference_path()):
    ref = Reference()
    threads = os.cpu_count() or 1
    secs, mv, sims, _ = ref.mcts_selfplay("tic_tac_toe", 20, args.sims, 1, threads=1)
    out["reference_cpp_1_thread"] = sims / secs
    secs, mv, sims, _ = ref.mcts_selfplay("tic_tac_toe", 20, args.sims, 1, threads=threads)
    out["reference_cpp_all_threads"] = sims / secs
    out["host_threads"] = threads
  eng = bsg.Engine("tic_tac_toe", local)
  from oracle_lib import reference_python
  mods = reference_python()
  if mods is not None:
    # the reference's own python/algorithms/mcts.py (unmodified bytecode, oracle/_ref) over its
    # own compiled game: (a) exactly as python/examples/mcts.py runs it, with its CPU
    # RandomRolloutEvaluator; (b) with this repo's GPU evaluator dropped into `evaluator`
    pyspiel, ref_mcts = mods
    rgame = pyspiel.load_game("tic_tac_toe")
    for key, ev in (("reference_python_mcts_py_cpu_evaluator",
                     ref_mcts.RandomRolloutEvaluator(1, np.random.RandomState(7))),
                    ("reference_python_mcts_py_gpu_evaluator",
                     bsg.GpuRolloutEvaluator(eng, n_rollouts=1, seed=7))):
      sims = 0
      t0 = time.perf_counter()
      for g in range(args.games):
        bot = ref_mcts.MCTSBot(rgame, 2.0, args.sims, ev, solve=True,
                               random_state=np.random.RandomState(3 + g))
        state = rgame.new_initial_state()
        while not state.is_terminal():
          root = bot.mcts_search(state)
          sims += root.explore_count
          state.apply_action(root.best_child().action)
      out[key] = sims / (time.perf_counter() - t0)
  for lb in (1, args.leaf_batch):
    ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=1, seed=7)
    bot_of = lambda game: bsg.MCTSBot(game, 2.0, args.sims, ev, solve=True,
                                      random_state=np.random.RandomState(3), leaf_batch=lb)
    selfplay(bsg, "tic_tac_toe", bot_of, 1)  # warm-up
    mv, sims, secs = selfplay(bsg, "tic_tac_toe", bot_of, args.games)
    out["ours_python_tree_leaf_batch_%d" % lb] = sims / secs
  for lb in (1, args.leaf_batch):
    bot_of = lambda game: bsg.NativeMCTSBot(game, 2.0, args.sims, eng, n_rollouts=1, solve=True,
                                            seed=7, shuffle_seed=3, leaf_batch=lb)
    selfplay(bsg, "tic_tac_toe", bot_of, 1)
    mv, sims, secs = selfplay(bsg, "tic_tac_toe", bot_of, max(args.games, 20))
    out["ours_native_tree_leaf_batch_%d" % lb] = sims / secs
  eng.close()
  return out


def run_cfg5(args, bsg, local, rank, world):
  import torch
  dist = None
  if world > 1:
    import torch.distributed as dist
    if not dist.is_initialized():
      dist.init_process_group("nccl", device_id=torch.device("cuda", local))
  dev = torch.device("cuda", local)
  game = bsg.load_game("connect_four")
  eng = bsg.Engine("connect_four", local)
  n_roll = 256
  total_sims = args.sims5 * world * args.trees         # weak scaling: fixed sims per tree
  ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=n_roll, seed=11)
  if args.forest:
    from open_spiel_1_b200.forest import ForestMCTS
    rp = ForestMCTS(game, 2.0, total_sims, ev, args.trees, sims_per_step=args.sims_per_step,
                    solve=True, seed=5, leaf_batch=args.leaf_batch, device=dev)
  else:
    rp = bsg.RootParallelMCTS(game, 2.0, total_sims, ev, sims_per_step=args.sims_per_step,
                              solve=True, seed=5, leaf_batch=args.leaf_batch, device=dev,
                              native=args.native, trees_per_rank=args.trees)
  state = game.new_initial_state()
  rp.mcts_search(state)                                # warm-up search
  torch.cuda.synchronize(dev)
  if dist is not None:
    dist.barrier()
  steps0, ar0 = (rp.total_steps if (args.native or args.forest) else ev.total_steps), rp.all_reduces
  t0 = time.perf_counter()
  root = rp.mcts_search(state)
  torch.cuda.synchronize(dev)
  secs = time.perf_counter() - t0
  stats = np.array([[c.action, c.explore_count, int(round(c.total_reward * n_roll))]
                    for c in sorted(root.children, key=lambda c: c.action)], np.int64)
  steps_now = rp.total_steps if (args.native or args.forest) else ev.total_steps
  t = torch.tensor([secs, float(steps_now - steps0)], dtype=torch.float64, device=dev)
  same = True
  if dist is not None:
    mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    secs, game_steps = float(mx[0]), float(sm[1])
    ref = torch.from_numpy(stats).to(dev); mine = ref.clone()
    dist.broadcast(ref, 0)
    flag = torch.tensor([int(torch.equal(ref, mine))], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    same = bool(flag.item())
  else:
    game_steps = float(t[1])
  assert same, "ranks disagree on the merged root statistics"
  visits = int(stats[:, 1].sum())
  out = {"config": "cfg5: connect_four root-parallel MCTS from the empty board, %d rollouts/leaf, "
                   "%d simulations/move/tree x %d tree(s)/rank x %d rank(s), %d sims per tree per "
                   "all-reduce step, leaf_batch %d"
                   % (n_roll, args.sims5, args.trees, world, args.sims_per_step, args.leaf_batch),
         "tree": ("GPU-resident forest (csrc/bsg_forest.cuh)" if args.forest else
                  ("native (csrc/bsg_mcts.cuh)" if args.native else "python mirror")),
         "n_gpus": world, "trees_per_rank": args.trees, "simulations": visits, "seconds": secs,
         "simulations_per_s": visits / secs, "playouts_per_s": visits * n_roll / secs,
         "game_steps_per_s": game_steps / secs, "all_reduces": rp.all_reduces - ar0,
         "merged_root_identical_on_all_ranks": same,
         "best_action": int(max(sorted(root.children, key=lambda c: c.action),
                                key=bsg.SearchNode.sort_key).action)}
  if rank == 0 and world == 1:
    from oracle_lib import Reference, reference_path
    if os.path.exists(reference_path()):
      threads = os.cpu_count() or 1
      r = Reference()
      # one search of the same shape per thread (the reference has no root-parallel mode)
      # bounded sample: one self-play game of 256 simulations/move per thread (a few seconds)
      secs_r, mv, sims, _ = r.mcts_selfplay("connect_four", 1, 256, n_roll, threads=threads)
      out["reference_cpp_all_threads_simulations_per_s"] = sims / secs_r
      out["host_threads"] = threads
  eng.close()
  return out


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--cfg", default="all")
  ap.add_argument("--sims", type=int, default=1000, help="cfg1 simulations per move")
  ap.add_argument("--games", type=int, default=3, help="cfg1 self-play games (ours)")
  ap.add_argument("--sims5", type=int, default=16384, help="cfg5 simulations per move per tree")
  ap.add_argument("--sims-per-step", type=int, default=4096)
  ap.add_argument("--leaf-batch", type=int, default=256)
  ap.add_argument("--trees", type=int, default=1,
                  help="cfg5: independent trees (host threads + engines) per rank, native only")
  ap.add_argument("--forest", type=int, default=0,
                  help="cfg5: 1 = the --trees trees of a rank live on the GPU (forest.ForestMCTS)")
  ap.add_argument("--native", type=int, default=1, help="cfg5: 1 = native host tree, 0 = Python mirror")
  args = ap.parse_args()
  rank, world, local = dist_env()
  import torch
  torch.cuda.set_device(local)
  import open_spiel_1_b200 as bsg
  lines = []
  if args.cfg in ("1", "all") and rank == 0 and world == 1:
    lines.append(run_cfg1(args, bsg, local))
  if args.cfg in ("5", "all"):
    res = run_cfg5(args, bsg, local, rank, world)
    if rank == 0:
      lines.append(res)
  if world > 1:
    import torch.distributed as dist
    if dist.is_initialized():
      dist.barrier()
      dist.destroy_process_group()
  for line in lines:
    print(json.dumps(line))
  return 0


if __name__ == "__main__":
  sys.exit(main())
