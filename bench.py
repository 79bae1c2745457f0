#!/usr/bin/env python
"""bench.py — headline benchmark of the rollout hot path (BASELINE.json configs[1]).

Workload at every N: per GPU, 16 Mi independent uniform-random connect_four playouts, one per
synthetic mid-game root (d ~ U{8..20} random moves from the empty board), i.e. weak scaling.
A "step" is one pass of the playout kernel over that batch.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--game G] [--playouts P]

For N > 1 launch under torchrun (one rank per GPU); rank 0 prints ONE JSON line.
  value      game-steps/s, whole job, roots already resident in HBM (CUDA events, max over ranks)
  e2e        same metric through the C-ABI host-buffer call bsg_playouts_host: pinned host
             roots -> H2D -> kernel -> D2H totals, all inside the timed region
  roofline   integer-issue roofline of the playout kernel (SURVEY.md §8d budget of 96
             thread-instructions per connect_four step over the integer issue rate measured
             live by bsg_int_peak), plus the HBM roofline of the observation writer
  cpu_baseline  the unmodified reference C++ (oracle/_ref/libspiel_ref.so) on the box's cores,
             on a bounded sample of the same kind of roots
`--impl reference` times only that CPU reference arm (rank 0) and prints its own line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

import numpy as np  # noqa: E402

OPS_PER_STEP = {"tic_tac_toe": 48, "connect_four": 96, "breakthrough": 140, "hex": 200,
                # pentago (not in SURVEY 8d; same counting rules): legal 4, rng 25, mod 4, 64-bit
                # rank-select 24, place 2, two quadrant turns 24, two 4-direction run-of-five
                # tests on 64-bit boards 96, outcome/loop 6
                "pentago": 185,
                # othello: parallel-prefix move generation 4 axes x 22 64-bit ops = 176, popc 2,
                # rng 25, mod 4, 64-bit rank-select 24, flips 8 directions x 19 64-bit ops = 304,
                # update 8, terminal test 6, loop 6
                "othello": 555,
                # y 19 (11-word sets over the 19 x 19 grid): legal 11, popc 11, rng 25, mod 4,
                # 11-word rank-select 50, place (one-hot 22, neighbour labels 44, set updates 55)
                # 121, amortised flood ~1.0 ring x 198 (11 words x (label 3 + within 5 + ring 10)),
                # loop 4
                "y": 424,
                # havannah 8 (8-word sets over the 15 x 15 grid, union-find groups): legal 8, popc 8, rng 25,
                # mod 4, 8-word rank-select 40, place 8, six neighbour tests 36, ~1.3 joins x 30 (two finds
                # + link), win test 10, amortised ring search 20, loop 4
                "havannah": 202,
                # go 9 x 9 (3-word sets): legal points 3 dilations 60 + ~1 chain flood 90, popc 3, rng 25, mod 4,
                # rank-select 25, place 6, four neighbour chains (flood + liberties) 4 x 60, hash 30 + probe of
                # the position set 10, loop 4
                "go(board_size=9)": 497}
ROOT_SEED = 20240611
METRIC = "connect_four game-steps/s (uniform-random playouts from mid-game roots)"


def parse_args():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=10)
  ap.add_argument("--warmup", type=int, default=3)
  ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
  ap.add_argument("--game", default="connect_four")
  ap.add_argument("--playouts", type=int, default=1 << 24, help="playouts per GPU per step")
  ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU baseline sample budget")
  ap.add_argument("--no-cpu-baseline", action="store_true")
  ap.add_argument("--no-extras", action="store_true", help="skip obs / other-game side numbers")
  ap.add_argument("--no-quality", action="store_true",
                  help="skip the one-tree quality yardstick of cfg5 and the cfg1 block (tens of thousands of "
                       "tiny launches: keeps an ncu launch list readable)")
  return ap.parse_args()


class NvmlClockSampler:
  """SM clock and throttle reasons of one GPU, read through NVML from a thread every millisecond
  of the timed region (the region is ~12 ms: nvidia-smi -lms 100 often delivers its first sample
  after it, always so on an eight-GPU box).  start() returns False when NVML is unavailable."""

  def __init__(self, torch_device):
    self.dev = torch_device
    self.samples = []
    self.stop_flag = False
    self.thread = None

  def start(self):
    try:
      import pynvml
      import torch
      pynvml.nvmlInit()
      p = torch.cuda.get_device_properties(self.dev)
      bus = "%08X:%02X:%02X.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
      h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
      self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
      pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
    except Exception:
      return False
    names = [("hw_slowdown", pynvml.nvmlClocksThrottleReasonHwSlowdown),
             ("hw_thermal_slowdown", pynvml.nvmlClocksThrottleReasonHwThermalSlowdown),
             ("sw_thermal_slowdown", pynvml.nvmlClocksThrottleReasonSwThermalSlowdown),
             ("sw_power_cap", pynvml.nvmlClocksThrottleReasonSwPowerCap)]

    def pump():
      while not self.stop_flag:
        try:
          mhz = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
          bits = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
          self.samples.append((float(mhz), [n for n, b in names if bits & b]))
        except Exception:
          pass
        time.sleep(0.001)

    self.thread = threading.Thread(target=pump, daemon=True)
    self.thread.start()
    return True

  def stop(self):
    self.stop_flag = True
    if self.thread:
      self.thread.join(timeout=1)
    if not self.samples:
      return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
    reasons = sorted({r for _, rs in self.samples for r in rs})
    return {"sm_mhz": float(np.median([m for m, _ in self.samples])), "sm_max_mhz": self.max_mhz,
            "reasons": reasons, "samples": len(self.samples), "source": "nvml, 1 ms period"}


class ClockSampler:
  """Samples nvidia-smi clocks and throttle reasons during the timed region (the fallback when
  NVML cannot be loaded into this process)."""
  QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
           "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
           "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

  def __init__(self, gpu_index):
    self.gpu = gpu_index
    self.proc = None
    self.lines = []

  def start(self):
    try:
      self.proc = subprocess.Popen(
          ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.QUERY,
           "--format=csv,noheader,nounits", "-lms", "100"],
          stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
      self.thread = threading.Thread(target=self._pump, daemon=True)
      self.thread.start()
    except Exception:
      self.proc = None

  def _pump(self):
    for line in self.proc.stdout:
      self.lines.append(line.strip())

  def stop(self):
    if not self.proc:
      return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    time.sleep(0.15)
    self.proc.terminate()
    try:
      self.proc.wait(timeout=2)
    except Exception:
      self.proc.kill()
    sm, mx, reasons = [], [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for line in self.lines:
      f = [x.strip() for x in line.split(",")]
      if len(f) < 9:
        continue
      try:
        sm.append(float(f[1])); mx.append(float(f[2]))
      except ValueError:
        continue
      for name, val in zip(names, f[5:9]):
        if val.lower().startswith("active"):
          reasons.add(name)
    if not sm:
      return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
    return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)),
            "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
  rank = int(os.environ.get("RANK", "0"))
  world = int(os.environ.get("WORLD_SIZE", "1"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  return rank, world, local


def load_reference():
  from oracle_lib import Oracle, Reference, reference_path
  if os.path.exists(reference_path()):
    return Reference(), "reference"
  return Oracle(), "port"


def sample_histories(game, n, seed):
  """Bounded sample of synthetic roots WITH histories for the CPU arm, built by the oracle
  under the same recipe as the GPU roots (d ~ U{dmin..dmax}; stop before a terminal move)."""
  from oracle_lib import Oracle
  from importlib import import_module
  depths = {"tic_tac_toe": (0, 4), "connect_four": (8, 20), "breakthrough": (0, 40), "hex": (0, 60),
            "pentago": (0, 20), "othello": (0, 40), "y": (0, 80), "havannah": (0, 60), "go(board_size=9)": (0, 60)}
  dmin, dmax = depths[game]
  orc = Oracle()
  rng = np.random.RandomState(seed & 0x7FFFFFFF)
  depth = rng.randint(dmin, dmax + 1, size=n)
  hist = []
  for i in range(n):
    acts, term, _ = orc.replay(game, [], seed=seed, playout=i, max_steps=int(depth[i]))
    hist.append(acts[:-1] if term else acts)
  return hist


def cpu_reference_run(game, hist, seconds, threads):
  """Times the reference's own C++ playout loop (benchmark_game.cc shape) on `threads` cores
  for about `seconds`; returns (steps/s, playouts/s, playouts run)."""
  ref, kind = load_reference()
  # calibrate on a small batch, then size the sample to the time budget
  n0 = 2000 * max(threads, 1)
  if kind == "reference":
    secs, steps, _ = ref.random_playouts(game, hist, n0, seed=1, threads=threads)
  else:
    t0 = time.time(); steps, _ = ref.random_playouts(game, hist, n0, seed=1); secs = time.time() - t0
    threads = 1
  rate = n0 / max(secs, 1e-6)
  n = int(max(n0, min(rate * seconds, 5e8)))
  if kind == "reference":
    secs, steps, _ = ref.random_playouts(game, hist, n, seed=2, threads=threads)
  else:
    t0 = time.time(); steps, _ = ref.random_playouts(game, hist, n, seed=2); secs = time.time() - t0
  return steps / secs, n / secs, n, kind, threads, secs


def run_reference_arm(args):
  """--impl reference: the reference's own CPU implementation of the path, all host threads."""
  rank, world, _ = dist_env()
  if rank != 0:
    return 0
  threads = os.cpu_count() or 1
  hist = sample_histories(args.game, 4096, ROOT_SEED + 2)
  per_step_seconds = max(1.0, min(10.0, 60.0 / max(args.steps + args.warmup, 1)))
  vals, last = [], None
  for i in range(args.warmup + args.steps):
    last = cpu_reference_run(args.game, hist, per_step_seconds, threads)
    if i >= args.warmup:
      vals.append(last)
  steps_s = float(np.mean([v[0] for v in vals]))
  playouts_s = float(np.mean([v[1] for v in vals]))
  kind, used = last[3], last[4]
  sample = "%d playouts/step from 4096 mid-game %s roots (d~U{8..20}), ~%.1f s/step" % (
      last[2], args.game, per_step_seconds)
  line = {
      "impl": "reference", "metric": METRIC, "value": steps_s, "unit": "game-steps/s",
      "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
      "ms_per_step": 1e3 * float(np.mean([v[5] for v in vals])), "higher_is_better": True,
      "scaling": "weak", "vs_baseline": None, "dtype": "int64", "dtype_note": "int64 actions over enum cell arrays (the reference C++ State classes)",
      "data": "synthetic", "playouts_per_s": playouts_s,
      "config": {"workload": "connect_four: 16Mi playouts from synthetic mid-game roots "
                             "(bounded CPU sample)", "game": args.game,
                 "root_depth": "U{8..20}", "rollouts_per_root": 1},
      "cpu_baseline": {"value": steps_s, "unit": "game-steps/s", "cores": used, "kind": kind,
                       "sample": sample},
      "e2e": {"value": steps_s, "unit": "game-steps/s", "h2d_bytes_per_step": 0,
              "d2h_bytes_per_step": 0},
      "gpu_launches": 0,
  }
  print(json.dumps(line))
  return 0


CFG5_TOTAL_SIMS = 1 << 20      # BASELINE cfg5: 1 Mi simulations per move, summed over all GPUs
CFG5_ROLLOUTS = 256            # rollouts per leaf
CFG5_SIMS_PER_TREE = 256       # => 4096 trees in the job, 4096 / N per GPU
CFG5_SIMS_PER_STEP = 64        # simulations per tree between two all-reduces => 4 per move
CFG5_LEAF_BATCH = 8


def cfg5_roots(bsg):
  """The empty board and 16 fixed mid-game connect_four roots (d ~ U{8..20} random legal moves,
  host PRNG seeded with ROOT_SEED + 5, terminal states rejected) as mirror states; the same on
  every rank."""
  game = bsg.load_game("connect_four")
  rng = np.random.RandomState(ROOT_SEED + 5)
  roots = [game.new_initial_state()]
  while len(roots) < 17:
    s = game.new_initial_state()
    for _ in range(int(rng.randint(8, 21))):
      la = s.legal_actions()
      if not la:
        break
      s.apply_action(la[int(rng.randint(len(la)))])
    if not s.is_terminal():
      roots.append(s)
  return game, roots


def best_of(bsg, root):
  """(best action, its Q for the mover, visits) by the reference's final-move rule
  (mcts.py:144-162 sort_key; ties to the lowest action id)."""
  best = max(sorted(root.children, key=lambda c: c.action), key=bsg.SearchNode.sort_key)
  q = best.total_reward / best.explore_count if best.explore_count else 0.0
  return int(best.action), float(q), int(best.explore_count)


def cfg5_single_tree_reference(bsg, local):
  """Quality yardstick of cfg5, rank 0 only: ONE native tree (csrc/bsg_mcts.cuh = the
  reference's MCTSBot search) with the whole 1 Mi-simulation budget on each of the 16 mid-game
  roots, 256 rollouts per leaf."""
  game, roots = cfg5_roots(bsg)
  eng = bsg.Engine("connect_four", local)
  out = []
  t0 = time.perf_counter()
  for i, st in enumerate(roots[1:]):
    bot = bsg.NativeMCTSBot(game, 2.0, CFG5_TOTAL_SIMS, eng, n_rollouts=CFG5_ROLLOUTS, solve=True,
                            seed=ROOT_SEED + 6, shuffle_seed=100 + i, leaf_batch=256)
    out.append(best_of(bsg, bot.mcts_search(st)))
    bot.close()
  secs = time.perf_counter() - t0
  eng.close()
  return {"best": out, "seconds": secs,
          "simulations_per_s": CFG5_TOTAL_SIMS * len(out) / secs}


def run_cfg5(bsg, torch, dev, local, rank, world, dist, single_tree):
  """BASELINE cfg5: connect_four root-parallel MCTS, 1 Mi simulations per move in the whole job,
  256 rollouts per leaf, ONE int64 ncclAllReduce per search step on the engine's own stream
  (bsg_forest_merge).  The job always holds 4096 trees of 256 simulations (4096 / N per GPU):
  tree v is "virtual rank" v whatever the GPU count, so the merged root must not depend on N."""
  from open_spiel_1_b200.forest import ForestMCTS
  game, roots = cfg5_roots(bsg)
  trees = CFG5_TOTAL_SIMS // CFG5_SIMS_PER_TREE // world
  eng = bsg.Engine("connect_four", local)
  ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=CFG5_ROLLOUTS, seed=ROOT_SEED + 7)
  forest = ForestMCTS(game, 2.0, CFG5_TOTAL_SIMS, ev, trees, sims_per_step=CFG5_SIMS_PER_STEP,
                      solve=True, seed=5, leaf_batch=CFG5_LEAF_BATCH, device=dev)

  def sync():
    torch.cuda.synchronize(dev)
    if dist is not None:
      dist.barrier()
      torch.cuda.synchronize(dev)

  forest.mcts_search(roots[0])          # warm-up move
  forest.allreduce_stats()
  steps0, launches0 = forest.total_steps, forest.info()["launches"]
  secs, results, sims = [], [], 0
  for st in roots:
    sync()
    t0 = time.perf_counter()
    root = forest.mcts_search(st)       # returns after the engine's stream has drained
    secs.append(time.perf_counter() - t0)
    results.append(root)
    sims += sum(c.explore_count for c in root.children) + trees * world   # + each tree's first visit (the root)
  ar = forest.allreduce_stats()
  game_steps = forest.total_steps - steps0
  launches = forest.info()["launches"] - launches0 + 4 * forest.n_steps * len(roots)
  # identical merged root on every rank
  tab = np.array([[c.action, c.explore_count, int(round(c.total_reward * CFG5_ROLLOUTS))]
                  for r in results for c in sorted(r.children, key=lambda c: c.action)], np.int64)
  tot_secs = float(sum(secs))
  same = True
  if dist is not None:
    t = torch.tensor([tot_secs, float(game_steps)], dtype=torch.float64, device=dev)
    mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    tot_secs, game_steps = float(mx[0]), float(sm[1])
    ref = torch.from_numpy(tab).to(dev); mine = ref.clone()
    dist.broadcast(ref, 0)
    flag = torch.tensor([int(torch.equal(ref, mine))], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    same = bool(flag.item())
  best = [best_of(bsg, r) for r in results]
  out = {
      "workload": "connect_four root-parallel MCTS: %d simulations per move in the whole job, %d "
                  "rollouts/leaf, from the empty board and 16 fixed mid-game roots (d~U{8..20}); "
                  "%d trees x %d simulations per GPU, leaf_batch %d, %d simulations per tree per "
                  "search step" % (CFG5_TOTAL_SIMS, CFG5_ROLLOUTS, trees, CFG5_SIMS_PER_TREE,
                                   CFG5_LEAF_BATCH, CFG5_SIMS_PER_STEP),
      "n_gpus": world, "trees_per_gpu": trees, "moves": len(roots),
      "simulations": sims, "seconds": tot_secs,
      "ms_per_move": 1e3 * tot_secs / len(roots),
      "simulations_per_s": sims / tot_secs,
      "playouts_per_s": game_steps and sims * CFG5_ROLLOUTS / tot_secs,
      "game_steps_per_s": game_steps / tot_secs,
      "all_reduces": ar["count"] if world > 1 else 0, "search_steps": forest.n_steps * len(roots),
      "allreduce_mean_us": ar["mean_us"], "allreduce_max_us": ar["max_us"],
      "collective": ("ncclAllReduce(int64 sum, %d B) on the engine's own communicator and stream "
                     "(bsg_forest_merge), CUDA events around it" % (7 * 5 * 8)) if world > 1 else
                    "world size 1: no collective",
      "merged_root_identical_on_all_ranks": same,
      "best_actions": [b[0] for b in best],
      "root_table_checksum": int(np.bitwise_xor.reduce((tab * np.array([1, 1000003, 7919], np.int64)).sum(1))),
      "trade_off": "1 Mi simulations per move = trees x simulations per tree: 4096 x 256 keeps a "
                   "real budget per tree (the alternative, 16384 x 64, fills the GPU better and searches "
                   "worse); one tree with the whole budget is the quality yardstick below",
      "gpu_launches": int(launches),
  }
  if single_tree is not None:
    agree = sum(1 for a, b in zip(best[1:], single_tree["best"]) if a[0] == b[0])
    dq = [abs(a[1] - b[1]) for a, b in zip(best[1:], single_tree["best"])]
    out["quality_vs_one_1Mi_simulation_tree"] = {
        "roots": len(single_tree["best"]), "best_action_agreement": agree,
        "mean_abs_diff_of_best_child_value": float(np.mean(dq)), "max_abs_diff": float(np.max(dq)),
        "single_tree_best_actions": [b[0] for b in single_tree["best"]],
        "single_tree_simulations_per_s": single_tree["simulations_per_s"],
        "note": "one native host tree (the reference's MCTSBot search, csrc/bsg_mcts.cuh) with all "
                "1 Mi simulations per root, 256 rollouts/leaf, leaf_batch 256, on one GPU"}
  forest.close()
  # weak-scaling companion: the SAME per-GPU forest at every N (4096 trees x 256 simulations per
  # GPU, so N Mi simulations per move in the job), one all-reduce per search step as above
  if world > 1:
    trees_w = CFG5_TOTAL_SIMS // CFG5_SIMS_PER_TREE
    ev_w = bsg.GpuRolloutEvaluator(eng, n_rollouts=CFG5_ROLLOUTS, seed=ROOT_SEED + 7)
    fw = ForestMCTS(game, 2.0, CFG5_TOTAL_SIMS * world, ev_w, trees_w, sims_per_step=CFG5_SIMS_PER_STEP,
                    solve=True, seed=5, leaf_batch=CFG5_LEAF_BATCH, device=dev)
    fw.mcts_search(roots[0])
    fw.allreduce_stats()
    secs_w, sims_w = 0.0, 0
    for st in roots[:5]:
      sync()
      t0 = time.perf_counter()
      root = fw.mcts_search(st)
      secs_w += time.perf_counter() - t0
      sims_w += sum(c.explore_count for c in root.children) + trees_w * world
    arw = fw.allreduce_stats()
    if dist is not None:
      t = torch.tensor([secs_w], dtype=torch.float64, device=dev)
      dist.all_reduce(t, op=dist.ReduceOp.MAX)
      secs_w = float(t[0])
    out["weak_scaling"] = {
        "workload": "the 1-GPU forest on every GPU: %d trees x %d simulations per GPU = %d simulations "
                    "per move in the job, 5 roots" % (trees_w, CFG5_SIMS_PER_TREE, CFG5_TOTAL_SIMS * world),
        "simulations": sims_w, "seconds": secs_w, "simulations_per_s": sims_w / secs_w,
        "ms_per_move": 1e3 * secs_w / 5, "all_reduces": arw["count"], "allreduce_mean_us": arw["mean_us"]}
    fw.close()
  eng.close()
  return out


def cpu_mcts_baselines():
  """the reference's own MCTS on the box's host cores.  C++
  MCTSBot + RandomRolloutEvaluator (algorithms/mcts.cc) one bot per thread for the cfg1 and cfg5
  shapes (ref_mcts_selfplay), and the unmodified python/algorithms/mcts.py for cfg1 exactly as
  python/examples/mcts.py runs it.  Bounded samples; rates in simulations/s."""
  from oracle_lib import Reference, reference_path, reference_python
  out = {"kind": "reference", "unit": "simulations/s"}
  if not os.path.exists(reference_path()):
    return {"kind": "port", "unavailable": "oracle/_ref/libspiel_ref.so not built"}
  ref = Reference()
  threads = os.cpu_count() or 1
  out["cores"] = threads
  secs, mv, sims, _ = ref.mcts_selfplay("tic_tac_toe", 20, 1000, 1, threads=1)
  out["cfg1_cpp_1_thread"] = sims / secs
  secs, mv, sims, _ = ref.mcts_selfplay("tic_tac_toe", 20, 1000, 1, threads=threads)
  out["cfg1_cpp_all_threads"] = sims / secs
  secs, mv, sims, _ = ref.mcts_selfplay("connect_four", 1, 128, CFG5_ROLLOUTS, threads=threads)
  out["cfg5_cpp_all_threads"] = sims / secs
  out["cfg5_cpp_playouts_per_s"] = sims * CFG5_ROLLOUTS / secs
  out["sample"] = ("cfg1: 20 self-play games of tic_tac_toe per thread, 1000 sims/move, n_rollouts=1, "
                   "uct_c=2; cfg5 shape: one self-play game of connect_four per thread at 128 "
                   "sims/move x 256 rollouts/leaf (the reference has no root-parallel mode: one "
                   "independent bot per thread, rate summed)")
  mods = reference_python()
  if mods is not None:
    pyspiel, ref_mcts = mods
    rgame = pyspiel.load_game("tic_tac_toe")
    ev = ref_mcts.RandomRolloutEvaluator(1, np.random.RandomState(7))
    sims = 0
    t0 = time.perf_counter()
    for g in range(2):
      bot = ref_mcts.MCTSBot(rgame, 2.0, 1000, ev, solve=True, random_state=np.random.RandomState(3 + g))
      state = rgame.new_initial_state()
      while not state.is_terminal():
        root = bot.mcts_search(state)
        sims += root.explore_count
        state.apply_action(root.best_child().action)
    out["cfg1_python_mcts_py_1_process"] = sims / (time.perf_counter() - t0)
  return out


def run_ours(args):
  import torch
  import open_spiel_1_b200 as bsg
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS

  rank, world, local = dist_env()
  torch.cuda.set_device(local)
  dev = torch.device("cuda", local)
  game = args.game
  # ---- rank 0 only, BEFORE the process group exists (the other ranks wait in the rendezvous,
  # not in an NCCL barrier that would keep their GPUs spinning): the CPU arms and the one-tree
  # quality yardstick of cfg5
  pre = {}
  if rank == 0:
    if not args.no_cpu_baseline:
      pre["cpu_baseline"] = cpu_baseline_block(args)
      try:
        pre["cpu_baseline"]["mcts"] = cpu_mcts_baselines()
      except Exception as ex:
        pre["cpu_baseline"]["mcts"] = {"error": repr(ex)}
    if not args.no_extras and not args.no_quality and game == "connect_four":
      try:
        pre["single_tree"] = cfg5_single_tree_reference(bsg, local)
        if world == 1:
          import bench_mcts
          pre["mcts_cfg1"] = bench_mcts.run_cfg1(argparse.Namespace(sims=1000, games=3, leaf_batch=256), bsg, local)
      except Exception as ex:
        pre["pre_error"] = repr(ex)
  if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
  else:
    dist = None
  eng = bsg.Engine(game, local)
  n = int(args.playouts)
  dmin, dmax = ROOT_DEPTHS[game]
  # distinct roots per rank (weak scaling): root seed offset by rank
  states, _ = make_roots(eng, n, dmin, dmax, seed=ROOT_SEED + 2 + 1000 * rank)
  ret0 = torch.zeros(n, dtype=torch.int8, device=dev)
  length = torch.zeros(n, dtype=torch.uint8, device=dev)
  totals = torch.zeros(4, dtype=torch.int64, device=dev)
  first = rank * (1 << 40)  # disjoint Philox playout ids per rank

  def barrier():
    torch.cuda.synchronize(dev)
    if dist is not None:
      dist.barrier()
    torch.cuda.synchronize(dev)

  def reduce_max_sum(vals):
    t = torch.tensor(vals, dtype=torch.float64, device=dev)
    if dist is None:
      return t.tolist(), t.tolist()
    mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    return mx.tolist(), sm.tolist()

  def one_step(i):
    eng.playouts(states, n, 1, seed=ROOT_SEED, first_playout=first + i * n, ret0=ret0,
                 length=length, totals=totals)

  for i in range(args.warmup):
    one_step(i)
  barrier()
  totals.zero_()
  launches0 = eng.launches
  sampler = NvmlClockSampler(dev)
  if rank == 0 and not sampler.start():
    sampler = ClockSampler(local)
    sampler.start()
  barrier()
  ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  ev0.record()
  for i in range(args.steps):
    one_step(args.warmup + i)
  ev1.record()
  barrier()
  ms = ev0.elapsed_time(ev1)
  clocks = sampler.stop() if rank == 0 else None
  gpu_launches = eng.launches - launches0
  tot = totals.cpu().numpy().astype(np.int64)
  mx, sm = reduce_max_sum([ms, float(tot[0]), float(tot[1])])
  ms_max, playouts, steps = mx[0], sm[1], sm[2]
  value = steps / (ms_max * 1e-3)

  e2e = e2e_block(args, eng, torch, dev, states, n, first, barrier, reduce_max_sum, world)

  extras = {}
  if not args.no_extras:
    extras.update(side_numbers(args, eng, torch, dev, rank, world, reduce_max_sum))
    if game == "connect_four":
      try:
        extras["mcts_cfg5"] = run_cfg5(bsg, torch, dev, local, rank, world, dist, pre.get("single_tree"))
      except Exception as ex:   # a side block must never kill the headline line
        extras["mcts_cfg5"] = {"error": repr(ex)}
  if rank == 0:
    extras.update(roofline_block(args, eng, value / world, ms_max / args.steps, n))
  if dist is not None:
    dist.barrier()
    dist.destroy_process_group()
  if rank != 0:
    return 0
  line = {
      "metric": METRIC, "value": value, "unit": "game-steps/s", "n_gpus": world,
      "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
      "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
      "dtype": "u64", "dtype_note": "u64 bitboards handled as u32 word pairs; integer only", "data": "synthetic",
      "playouts_per_s": playouts / (ms_max * 1e-3),
      "mean_steps_per_playout": steps / max(playouts, 1.0),
      "config": {"workload": "connect_four: %d independent random playouts per GPU per step from "
                             "synthetic mid-game roots, 1 playout/root" % n,
                 "game": game, "playouts_per_gpu": n, "rollouts_per_root": 1,
                 "root_depth": "U{%d..%d} random moves from the initial state" % (dmin, dmax),
                 "root_seed": ROOT_SEED + 2, "playout_seed": ROOT_SEED,
                 "rng": "philox4x32-10, counter=(playout id, step>>2), action=legal[r mod n]",
                 "l2_policy": "inputs larger than L2 (%d MB of roots per step)" % (states.numel() >> 20),
                 "parallelism": "playouts sharded by root across %d GPU(s), no data-path collective" % world},
      "clocks": clocks,
      "e2e": e2e,
      "gpu_launches": gpu_launches,
  }
  line.update(extras)
  line.pop("_peak_int", None)
  if "cpu_baseline" in pre:
    line["cpu_baseline"] = pre["cpu_baseline"]
  if "mcts_cfg1" in pre:
    line["mcts_cfg1"] = pre["mcts_cfg1"]
  if "pre_error" in pre:
    line["pre_error"] = pre["pre_error"]
  print(json.dumps(line))
  return 0


def time_kernel(torch, dev, fn, reps=5, warm=2):
  for _ in range(warm):
    fn()
  torch.cuda.synchronize(dev)
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record()
  for _ in range(reps):
    fn()
  e1.record()
  torch.cuda.synchronize(dev)
  return e0.elapsed_time(e1) / reps


def e2e_block(args, eng, torch, dev, states, n, first, barrier, reduce_max_sum, world):
  """The same metric through the host-buffer C-ABI call (pinned host roots, copies inside the
  timed region), max over ranks; plus what bounds it: the H2D bandwidth of the very same pinned
  buffer measured in the same job at this N, and a variant that returns per-root results (what
  evaluate_batch consumes) instead of the 32-byte totals."""
  host_roots = torch.empty(states.numel(), dtype=torch.uint8, pin_memory=True)
  host_roots.copy_(states)
  torch.cuda.synchronize(dev)
  host_np = host_roots.numpy()
  reps = max(2, min(args.steps, 5))
  for i in range(2):
    eng.playouts_host(host_np, n, 1, seed=ROOT_SEED, first_playout=first + i * n)
  barrier()
  t0 = time.perf_counter()
  game_steps = 0
  for i in range(reps):
    t, _, _ = eng.playouts_host(host_np, n, 1, seed=ROOT_SEED, first_playout=first + (args.warmup + i) * n)
    game_steps += int(t[1])
  torch.cuda.synchronize(dev)
  secs = time.perf_counter() - t0
  mx, sm = reduce_max_sum([secs, float(game_steps)])
  out = {"value": sm[1] / mx[0], "unit": "game-steps/s",
         "h2d_bytes_per_step": int(states.numel()), "d2h_bytes_per_step": 32,
         "api": "bsg_playouts_host (C ABI, pinned host roots in, totals out)"}
  # the same through the compact root format (connect_four: 49 bits per root instead of 64): the
  # H2D copy is what bounds the call, so this is the API path a throughput caller takes
  try:
    comp_np = eng.compact_roots(host_np, n)
    if comp_np.nbytes < host_np.nbytes:
      comp = torch.empty(comp_np.nbytes, dtype=torch.uint8, pin_memory=True)
      comp.numpy()[:] = comp_np
      c_np = comp.numpy()
      for i in range(2):
        eng.playouts_host_compact(c_np, n, 1, seed=ROOT_SEED, first_playout=first + i * n)
      barrier()
      t0 = time.perf_counter()
      game_steps_c = 0
      for i in range(reps):
        t, _, _ = eng.playouts_host_compact(c_np, n, 1, seed=ROOT_SEED, first_playout=first + (args.warmup + i) * n)
        game_steps_c += int(t[1])
      torch.cuda.synchronize(dev)
      secs_c = time.perf_counter() - t0
      mxc, smc = reduce_max_sum([secs_c, float(game_steps_c)])
      out["packed8"] = {"value": out["value"], "h2d_bytes_per_step": out["h2d_bytes_per_step"], "api": out["api"]}
      out["value"] = smc[1] / mxc[0]
      out["h2d_bytes_per_step"] = int(comp_np.nbytes)
      out["api"] = ("bsg_playouts_host_compact (C ABI, pinned host roots in the 49-bit compact format in, totals "
                    "out); same playouts and totals as the packed call (checked: %s)" % (game_steps_c == game_steps))
      del comp
  except Exception as ex:
    out["compact_error"] = repr(ex)
  # H2D alone: the same pinned buffer, all ranks at once (what caps e2e when N GPUs share one
  # host's memory system)
  scratch = torch.empty_like(states)
  scratch.copy_(host_roots, non_blocking=True)
  barrier()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record()
  for _ in range(reps):
    scratch.copy_(host_roots, non_blocking=True)
  e1.record()
  torch.cuda.synchronize(dev)
  h2d_ms = e0.elapsed_time(e1) / reps
  mx, _ = reduce_max_sum([h2d_ms])
  gbs = states.numel() / (mx[0] * 1e-3) / 1e9
  out["h2d_only_gbs"] = gbs * world
  out["h2d_only_gbs_per_gpu"] = gbs
  out["h2d_only_ceiling_game_steps_per_s"] = sm[1] / reps / (mx[0] * 1e-3)
  out["numa_nodes"] = numa_summary(eng.device)
  del scratch
  # the same call with the roots in WRITE-COMBINED page-locked memory (bsg_host_alloc)
  try:
    from open_spiel_1_b200.engine import HostBuffer
    wc = HostBuffer(states.numel(), write_combined=True)
    wc.array[:] = host_np
    for i in range(2):
      eng.playouts_host(wc.array, n, 1, seed=ROOT_SEED, first_playout=first + i * n)
    barrier()
    t0 = time.perf_counter()
    game_steps = 0
    for i in range(reps):
      t, _, _ = eng.playouts_host(wc.array, n, 1, seed=ROOT_SEED, first_playout=first + (args.warmup + i) * n)
      game_steps += int(t[1])
    torch.cuda.synchronize(dev)
    secs = time.perf_counter() - t0
    mxw, smw = reduce_max_sum([secs, float(game_steps)])
    out["write_combined_roots"] = {"value": smw[1] / mxw[0], "unit": "game-steps/s",
                                   "api": "bsg_playouts_host, roots in cudaHostAllocWriteCombined memory (bsg_host_alloc)"}
    wc.close()
  except Exception as ex:
    out["write_combined_roots"] = {"error": repr(ex)}
  # per-playout results: (int)Returns()[0] and the move count of every playout copied back into
  # page-locked host arrays (2 B per playout) — what evaluate_batch consumes
  try:
    h_ret = torch.zeros(n, dtype=torch.int8, pin_memory=True)
    h_len = torch.zeros(n, dtype=torch.uint8, pin_memory=True)
    r_np, l_np = h_ret.numpy(), h_len.numpy()
    comp_np = eng.compact_roots(host_np, n)
    comp = torch.empty(comp_np.nbytes, dtype=torch.uint8, pin_memory=True)
    comp.numpy()[:] = comp_np
    c_np = comp.numpy()
    for i in range(2):
      eng.playouts_host_compact(c_np, n, 1, seed=ROOT_SEED, first_playout=first + i * n, ret0=r_np, length=l_np)
    barrier()
    t0 = time.perf_counter()
    game_steps = 0
    for i in range(reps):
      t, _, _ = eng.playouts_host_compact(c_np, n, 1, seed=ROOT_SEED, first_playout=first + (args.warmup + i) * n,
                                          ret0=r_np, length=l_np)
      game_steps += int(t[1])
    torch.cuda.synchronize(dev)
    secs = time.perf_counter() - t0
    mx, sm = reduce_max_sum([secs, float(game_steps)])
    out["per_root_results"] = {"value": sm[1] / mx[0], "unit": "game-steps/s",
                               "h2d_bytes_per_step": int(comp_np.nbytes), "d2h_bytes_per_step": 2 * n + 32,
                               "checksum_ok": bool(int(l_np.astype(np.int64).sum()) == int(t[1])),
                               "api": "bsg_playouts_host_compact (C ABI): pinned compact host roots in, int8 return + "
                                      "uint8 length per playout out, copied back behind each chunk's kernel"}
  except Exception as ex:
    out["per_root_results"] = {"error": repr(ex)}
  del host_roots
  return out


def numa_summary(device):
  """NUMA node of this GPU and the node count of the host, from sysfs (best effort)."""
  try:
    import glob
    nodes = len(glob.glob("/sys/devices/system/node/node[0-9]*"))
    bus = subprocess.check_output(["nvidia-smi", "-i", str(device), "--query-gpu=pci.bus_id",
                                   "--format=csv,noheader"], text=True).strip().lower()
    path = "/sys/bus/pci/devices/%s/numa_node" % bus[-12:]
    node = open(path).read().strip() if os.path.exists(path) else "?"
    return {"host_nodes": nodes, "gpu_node": node}
  except Exception:
    return None


def cpu_baseline_block(args):
  threads = os.cpu_count() or 1
  hist = sample_histories(args.game, 4096, ROOT_SEED + 2)
  v_all = cpu_reference_run(args.game, hist, args.cpu_seconds, threads)
  v_one = cpu_reference_run(args.game, hist, min(args.cpu_seconds, 4.0), 1)
  return {
      "value": v_all[0], "unit": "game-steps/s", "cores": v_all[4], "kind": v_all[3],
      "sample": "%d playouts from 4096 synthetic mid-game roots (d~U{8..20}), %.1f s, reference "
                "State::Clone/LegalActions/ApplyAction loop, one std::thread per core" % (v_all[2], v_all[5]),
      "single_core_value": v_one[0], "playouts_per_s": v_all[1],
      "cpu_model": cpu_model(),
  }


def roofline_block(args, eng, steps_per_s_per_gpu, kernel_ms, n):
  """Integer-issue roofline of the playout kernel; rank 0."""
  # integer-issue peak measured live on this GPU: best of ALU-only, IMAD-only and mixed chains
  kinds = {0: "LOP3+IADD3 (ALU pipe)", 1: "IMAD (FMA pipe)", 2: "SHF+LOP3+IMAD 2:1",
           3: "LOP3+IMAD 1:1"}
  peaks = {k: eng.int_peak(k, 4000) for k in kinds}
  peak_int = max(peaks.values())
  ops = OPS_PER_STEP[args.game]
  achieved = steps_per_s_per_gpu * ops
  return {"roofline": {
      "bound": "int_issue", "kernel": "playout_kernel<ConnectFour,false>",
      "achieved": achieved / 1e12, "peak": peak_int / 1e12, "unit": "T thread-instr/s",
      "frac": achieved / peak_int, "traffic": ncu_traffic_bytes("playout_kernel<ConnectFour, 0"),
      "traffic_source": "constant from the newest committed ncu --set full capture under profiles/ "
                        "(dram__bytes_read.sum + dram__bytes_write.sum of one launch), not measured by this run",
      "algorithmic_ops_per_step": ops,
      "peak_source": "bsg_int_peak microbenchmark, live: " + ", ".join(
          "%s=%.2fT" % (kinds[k], peaks[k] / 1e12) for k in kinds),
      "kernel_ms": kernel_ms,
      "note": "the path is integer/branchy by design (north_star): the bound is the integer "
              "issue rate, not HBM or tensor; algorithmic HBM bytes are 8 B root + 2 B results per "
              "playout (~1 B/step)",
      "hbm_achieved_gbs": n * (eng.state_bytes + 2) / (kernel_ms * 1e-3) / 1e9,
  }, "_peak_int": peak_int}


def side_numbers(args, eng, torch, dev, rank, world, reduce_max_sum):
  """Side numbers on EVERY rank (weak scaling, reduced: max time, summed work): observation
  writer roofline, cfg3 trajectory write-out, cfg3 / cfg4 and the other games' playout rates."""
  import open_spiel_1_b200 as bsg
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  out, side = {}, {}
  peaks_path = os.path.join(REPO, "MEASURED_PEAKS.json")
  hbm_peak, hbm_src = 6650.0, "fallback (B200_PROFILING.md)"
  if os.path.exists(peaks_path):
    with open(peaks_path) as f:
      hbm_peak, hbm_src = float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
  peak_int = max(eng.int_peak(k, 2000) for k in (2, 3))
  try:
    # observation writer: HBM-write bound (breakthrough, 784 algorithmic B/state)
    e_bt = bsg.Engine("breakthrough", eng.device)
    n_obs = 1 << 22
    st, _ = make_roots(e_bt, n_obs, 0, 40, seed=ROOT_SEED + 3 + 1000 * rank)
    obs = torch.empty((n_obs, 3, 8, 8), dtype=torch.float32, device=dev)
    ms = time_kernel(torch, dev, lambda: e_bt.observations(st, n_obs, None, out=obs))
    mx, _ = reduce_max_sum([ms])
    gbs = n_obs * 784 / (mx[0] * 1e-3) / 1e9
    out["obs_roofline"] = {"bound": "hbm", "kernel": "expand_kernel<Breakthrough,false>",
                           "achieved": gbs, "peak": hbm_peak, "unit": "GB/s",
                           "frac": gbs / hbm_peak, "traffic": ncu_traffic_bytes("expand_kernel<Breakthrough, 0"),
                           "traffic_source": "constant from the newest committed ncu capture under profiles/",
                           "peak_source": hbm_src, "per_gpu": True, "n_gpus": world,
                           "states": n_obs, "obs_per_s": n_obs * world / (mx[0] * 1e-3),
                           "algorithmic_bytes_per_state": 784}
    del obs, st
    # BASELINE cfg3: breakthrough playouts + ObservationTensor / legal-mask write-out of every
    # visited state (the AlphaZero actor's records), all on the device
    n_tr = 1 << 18
    st, _ = make_roots(e_bt, n_tr, 0, 40, seed=ROOT_SEED + 5 + 1000 * rank)
    rec = {}

    def run_traj():
      rec["t"] = e_bt.trajectories(st, n_tr, 1, seed=9, want_actions=True)
    ms = time_kernel(torch, dev, run_traj, reps=3, warm=2)
    total = rec["t"]["total"]
    mx, sm = reduce_max_sum([ms, float(total)])
    out["trajectory_cfg3"] = {
        "workload": "breakthrough: %d playouts per GPU, observation [3,8,8] f32 + legal mask [768] u8 + "
                    "action of every decision point" % n_tr,
        "n_gpus": world, "records": sm[1], "ms": mx[0], "records_per_s": sm[1] / (mx[0] * 1e-3),
        "bytes_written_per_record": 768 + 768 + 20,
        "output_gb_per_s": sm[1] * (768 + 768 + 20) / (mx[0] * 1e-3) / 1e9}
    rec.clear()
    del st
    e_bt.close()
    torch.cuda.empty_cache()
    # the AlphaZero actor proper (alpha_zero.py:196-228): MCTS self-play, many games in flight,
    # policy = root visit counts, records written into a device replay buffer
    try:
      from open_spiel_1_b200.actor import ForestActor
      e_c4 = bsg.Engine("connect_four", eng.device)
      g_c4 = bsg.load_game("connect_four")
      n_games, sims, roll = 4096, 64, 16
      ev = bsg.GpuRolloutEvaluator(e_c4, n_rollouts=roll, seed=ROOT_SEED + 8)
      actor = ForestActor(g_c4, ev, n_games, uct_c=2.0, max_simulations=sims, temperature=1.0,
                          temperature_drop=10, solve=False, seed=1000 * rank, sample_seed=ROOT_SEED + 9 + rank,
                          leaf_batch=8)
      rbuf = bsg.DeviceReplayBuffer(1 << 18, e_c4.obs_size, e_c4.num_actions, dev)
      actor.play_batch(buffer=rbuf)
      torch.cuda.synchronize(dev)
      t0 = time.perf_counter()
      res = actor.play_batch(buffer=rbuf)
      torch.cuda.synchronize(dev)
      secs = time.perf_counter() - t0
      mx, sm = reduce_max_sum([secs, float(res["total"])])
      out["trajectory_actor"] = {
          "workload": "connect_four MCTS self-play: %d games in flight per GPU, %d simulations/move, "
                      "%d rollouts/leaf, temperature 1, temperature_drop 10; records (observation, "
                      "legal mask, visit-count policy, final return) into a device replay buffer"
                      % (n_games, sims, roll),
          "n_gpus": world, "records": sm[1], "seconds": mx[0], "records_per_s": sm[1] / mx[0],
          "games_per_s": n_games * world / mx[0], "moves": int(res["moves"]),
          "simulations_per_s": sm[1] * sims / mx[0]}
      actor.close()
      e_c4.close()
    except Exception as ex:
      out["trajectory_actor"] = {"error": repr(ex)}
    for g, npl in (("tic_tac_toe", 1 << 24), ("breakthrough", 1 << 22), ("hex", 1 << 20),
                   ("pentago", 1 << 22), ("othello", 1 << 21), ("y", 1 << 21), ("havannah", 1 << 20), ("go(board_size=9)", 1 << 20)):
      e_g = bsg.Engine(g, eng.device)
      dmin, dmax = ROOT_DEPTHS[g]
      st, _ = make_roots(e_g, npl, dmin, dmax, seed=ROOT_SEED + 4 + 1000 * rank)
      tot = torch.zeros(4, dtype=torch.int64, device=dev)
      ms = time_kernel(torch, dev, lambda: e_g.playouts(st, npl, 1, seed=7, totals=tot), reps=3, warm=1)
      t = tot.cpu().numpy()
      mx, sm = reduce_max_sum([ms, float(t[1]) / 4.0])
      rate = sm[1] / (mx[0] * 1e-3)
      side[g] = {"playouts_per_gpu": npl, "n_gpus": world, "ms": mx[0], "game_steps_per_s": rate,
                 "playouts_per_s": npl * world / (mx[0] * 1e-3),
                 "int_roofline_frac": rate / world * OPS_PER_STEP[g] / peak_int}
      if g in ("hex", "y"):
        # value-only launch (bsg_playout_io.value_only): same returns, plain stones to a full
        # board + one reading of the winner; what an Evaluator needs
        tot.zero_()
        ms_v = time_kernel(torch, dev, lambda: e_g.playouts(st, npl, 1, seed=7, totals=tot, value_only=True),
                           reps=3, warm=1)
        mxv, _ = reduce_max_sum([ms_v])
        side[g]["value_only"] = {"ms": mxv[0], "playouts_per_s": npl * world / (mxv[0] * 1e-3),
                                 "speedup_in_playouts": mx[0] / mxv[0],
                                 "sum_ret0_equal": bool(int(tot.cpu().numpy()[2]) == int(t[2]))}
      e_g.close()
  except Exception as ex:  # side numbers must never kill the headline line
    side["error"] = repr(ex)
  out["other_games"] = side
  return out


def ncu_traffic_bytes(kernel_substr):
  """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the kernel, from the newest
  committed ncu summary (profiles/r*_kernels.csv, written by profiles/summarize_ncu.py)."""
  import csv
  import glob
  for path in sorted(glob.glob(os.path.join(REPO, "profiles", "r*_kernels.csv")), reverse=True):
    try:
      rows = list(csv.reader(open(path)))
      hdr = rows[0]
      rd = [i for i, h in enumerate(hdr) if h.startswith("dram__bytes_read.sum")][0]
      wr = [i for i, h in enumerate(hdr) if h.startswith("dram__bytes_write.sum")][0]
      scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
      for r in rows[1:]:
        if kernel_substr in r[0]:
          return (float(r[rd]) * scale[hdr[rd].split("[")[1].rstrip("]")] +
                  float(r[wr]) * scale[hdr[wr].split("[")[1].rstrip("]")])
    except Exception:
      continue
  return None


def cpu_model():
  try:
    with open("/proc/cpuinfo") as f:
      for line in f:
        if line.startswith("model name"):
          return line.split(":", 1)[1].strip()
  except Exception:
    pass
  return "unknown"


def main():
  args = parse_args()
  if args.impl == "reference":
    return run_reference_arm(args)
  return run_ours(args)


if __name__ == "__main__":
  sys.exit(main())
