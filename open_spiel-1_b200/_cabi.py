"""ctypes binding of include/bsg.h (the C ABI of the engine).

The library is built in-tree by `__graft_entry__.build()` (or `make -C csrc`) as
`open_spiel-1_b200/libbsg.so`.  There is no Python/CPU stand-in: if the library is missing,
importing this module raises, and every device entry point raises when no GPU is present.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BSG_LIB") or os.path.join(_HERE, "libbsg.so")  # BSG_LIB: developer override (A/B builds)

vp = ctypes.c_void_p
i32, i64, u32, u64 = ctypes.c_int32, ctypes.c_int64, ctypes.c_uint32, ctypes.c_uint64

OK = 0
ERR_BAD_GAME, ERR_BAD_ARG, ERR_ILLEGAL_ACTION, ERR_CUDA, ERR_NO_DEVICE, ERR_STUCK_LANE, ERR_NCCL = (
    -1, -2, -3, -4, -5, -6, -7)
TERMINAL_PLAYER = -4

TIC_TAC_TOE, CONNECT_FOUR, BREAKTHROUGH, HEX, HEX5_HOST_ONLY, PENTAGO, OTHELLO, Y, Y9_HOST_ONLY, HAVANNAH, HAVANNAH4_HOST_ONLY, GO, GO7_HOST_ONLY = 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12
DEVICE_GAMES = (TIC_TAC_TOE, CONNECT_FOUR, BREAKTHROUGH, HEX, PENTAGO, OTHELLO, Y, HAVANNAH, GO)   # games with CUDA kernels


class PlayoutIO(ctypes.Structure):
  """struct bsg_playout_io (include/bsg.h)."""
  _fields_ = [
      ("max_steps", i32),
      ("step_limit", vp),
      ("r_table", vp),
      ("r_stride", i32),
      ("ret0", vp),
      ("length", vp),
      ("terminal", vp),
      ("actions", vp),
      ("actions_stride", i32),
      ("root_sum_ret0", vp),
      ("root_sum_steps", vp),
      ("final_states", vp),
      ("final_before_terminal", i32),
      ("totals", vp),
      ("state_log", vp),
      ("state_log_cap", i64),
      ("state_log_off", vp),
      ("root_playout_base", vp),
      ("value_only", i32),
      ("state_log_stride", i64),
  ]


class ActorIO(ctypes.Structure):
  """struct bsg_actor_io (include/bsg.h)."""
  _fields_ = [("states", vp), ("alive", vp), ("length", vp), ("outcome", vp), ("n_alive_dev", vp),
              ("state_log", vp), ("policy_log", vp), ("est_log", vp), ("action_log", vp), ("weights", vp),
              ("n_games", i32), ("max_moves", i32), ("move", i32), ("temperature_drop", i32),
              ("temperature", ctypes.c_double), ("sample_seed", u64), ("game_id_base", u64)]


class ActorEmitIO(ctypes.Structure):
  """struct bsg_actor_emit_io (include/bsg.h)."""
  _fields_ = [("state_log", vp), ("policy_log", vp), ("est_log", vp), ("action_log", vp),
              ("length", vp), ("outcome", vp), ("offsets", vp), ("out_states", vp), ("out_policy", vp),
              ("out_value", vp), ("out_est", vp), ("out_action", vp), ("out_player", vp),
              ("n_games", i32), ("max_moves", i32), ("total", i64)]


def _load():
  if not os.path.exists(LIB_PATH):
    raise ImportError(
        "%s not found: build the CUDA extension first (python -c 'import __graft_entry__ as g; "
        "g.build()' or make -C open_spiel-1_b200/csrc). There is no CPU fallback." % LIB_PATH)
  lib = ctypes.CDLL(LIB_PATH)
  sig = {
      "bsg_version": (ctypes.c_int, []),
      "bsg_error_string": (ctypes.c_char_p, [ctypes.c_int]),
      "bsg_game_id": (ctypes.c_int, [ctypes.c_char_p]),
      "bsg_num_distinct_actions": (ctypes.c_int, [ctypes.c_int]),
      "bsg_max_game_length": (ctypes.c_int, [ctypes.c_int]),
      "bsg_observation_size": (ctypes.c_int, [ctypes.c_int]),
      "bsg_observation_shape": (ctypes.c_int, [ctypes.c_int, vp]),
      "bsg_state_bytes": (ctypes.c_int, [ctypes.c_int]),
      "bsg_plane_bytes": (ctypes.c_int, [ctypes.c_int, vp, ctypes.c_int]),
      "bsg_pack_histories": (ctypes.c_int, [ctypes.c_int, vp, vp, i64, vp, i64]),
      "bsg_host_state_info": (ctypes.c_int, [ctypes.c_int, vp, i64, i64, vp, vp, vp]),
      "bsg_host_legal_actions": (ctypes.c_int, [ctypes.c_int, vp, i64, i64, vp, ctypes.c_int]),
      "bsg_host_apply_action": (ctypes.c_int, [ctypes.c_int, vp, i64, i64, i64]),
      "bsg_host_observation": (ctypes.c_int, [ctypes.c_int, vp, i64, i64, ctypes.c_int, vp]),
      "bsg_host_legal_mask": (ctypes.c_int, [ctypes.c_int, vp, i64, i64, vp]),
      "bsg_host_replay_step": (ctypes.c_int, [ctypes.c_int, vp, i64, i64, u32, vp]),
      "bsg_host_fill_playout": (ctypes.c_int, [ctypes.c_int, vp, i64, i64, ctypes.c_uint64, ctypes.c_uint64, vp, vp]),
      "bsg_host_to_string": (ctypes.c_int, [ctypes.c_int, vp, i64, i64, vp, ctypes.c_int]),
      "bsg_host_serialize": (ctypes.c_int, [ctypes.c_int, vp, i64, i64, vp, ctypes.c_int]),
      "bsg_host_deserialize": (ctypes.c_int, [ctypes.c_int, ctypes.c_char_p, vp, i64, i64]),
      "bsg_draw": (u32, [u64, u64, u32]),
      "bsg_create": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.POINTER(vp)]),
      "bsg_destroy": (None, [vp]),
      "bsg_last_error": (ctypes.c_char_p, [vp]),
      "bsg_engine_game": (ctypes.c_int, [vp]),
      "bsg_engine_device": (ctypes.c_int, [vp]),
      "bsg_engine_sm_count": (ctypes.c_int, [vp]),
      "bsg_engine_set_value_only": (ctypes.c_int, [vp, ctypes.c_int]),
      "bsg_playouts": (ctypes.c_int, [vp, vp, i64, i32, u64, u64, ctypes.POINTER(PlayoutIO), vp]),
      "bsg_observations": (ctypes.c_int, [vp, vp, i64, vp, vp, vp, vp]),
      "bsg_playouts_host": (ctypes.c_int, [vp, vp, i64, i32, u64, u64, vp, vp, vp]),
      "bsg_observations_host": (ctypes.c_int, [vp, vp, i64, vp, vp, vp]),
      "bsg_playouts_host_results": (ctypes.c_int, [vp, vp, i64, i32, u64, u64, vp, vp, vp]),
      "bsg_compact_state_bits": (ctypes.c_int, [ctypes.c_int]),
      "bsg_compact_buffer_bytes": (ctypes.c_int64, [ctypes.c_int, ctypes.c_int64]),
      "bsg_compact_roots": (ctypes.c_int, [ctypes.c_int, vp, i64, vp]),
      "bsg_expand_roots": (ctypes.c_int, [ctypes.c_int, vp, i64, vp]),
      "bsg_playouts_host_compact": (ctypes.c_int, [vp, vp, i64, i32, u64, u64, vp, vp, vp]),
      "bsg_int_peak": (ctypes.c_int, [vp, ctypes.c_int, ctypes.c_int, vp]),
      "bsg_noalloc_max_playouts": (ctypes.c_int, [ctypes.c_int]),
      "bsg_playouts_value_only_noalloc": (ctypes.c_int, [vp, vp, i64, i32, u64, u64, vp, vp, vp, vp, vp, vp]),
      "bsg_mcts_create": (ctypes.c_int, [vp, ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                          u64, u64, ctypes.POINTER(vp)]),
      "bsg_mcts_destroy": (None, [vp]),
      "bsg_mcts_last_error": (ctypes.c_char_p, [vp]),
      "bsg_mcts_search": (ctypes.c_int, [vp, vp, i64, ctypes.c_int]),
      "bsg_mcts_expand_root": (ctypes.c_int, [vp, vp]),
      "bsg_mcts_root_children": (ctypes.c_int, [vp, ctypes.c_int, vp, vp, vp, vp]),
      "bsg_mcts_info": (ctypes.c_int, [vp, vp]),
      "bsg_mcts_set_root_child": (ctypes.c_int, [vp, i64, i64, ctypes.c_double, i32]),
      "bsg_mcts_set_root_explore": (ctypes.c_int, [vp, i64]),
      "bsg_mcts_set_next_playout": (ctypes.c_int, [vp, u64]),
      "bsg_forest_create": (ctypes.c_int, [vp, i32, i32, i64, ctypes.c_double, ctypes.c_int, ctypes.c_int,
                                           ctypes.c_int, u64, u64, u64, ctypes.POINTER(vp)]),
      "bsg_forest_destroy": (None, [vp]),
      "bsg_forest_last_error": (ctypes.c_char_p, [vp]),
      "bsg_forest_search": (ctypes.c_int, [vp, vp, i64, vp, ctypes.c_int]),
      "bsg_forest_pack_root": (ctypes.c_int, [vp, vp, vp]),
      "bsg_forest_install_root": (ctypes.c_int, [vp, vp, vp]),
      "bsg_forest_tree_root": (ctypes.c_int, [vp, i32, ctypes.c_int, vp, vp, vp, vp, vp]),
      "bsg_forest_info": (ctypes.c_int, [vp, vp]),
      "bsg_trajectory_scan": (ctypes.c_int, [vp, vp, i64, vp, vp, vp]),
      "bsg_trajectory_compact": (ctypes.c_int, [vp, vp, i64, i64, vp, i64, vp, i32, vp, vp, vp, vp, vp]),
      "bsg_forest_search_roots": (ctypes.c_int, [vp, vp, vp, i64]),
      "bsg_forest_actor_step": (ctypes.c_int, [vp, ctypes.POINTER(ActorIO), vp]),
      "bsg_actor_emit": (ctypes.c_int, [vp, ctypes.POINTER(ActorEmitIO)]),
      "bsg_engine_stream": (vp, [vp]),
      "bsg_host_alloc": (vp, [ctypes.c_size_t, ctypes.c_int]),
      "bsg_host_free": (None, [vp]),
      "bsg_comm_unique_id": (ctypes.c_int, [vp]),
      "bsg_comm_init": (ctypes.c_int, [vp, ctypes.c_int, ctypes.c_int, vp]),
      "bsg_comm_destroy": (ctypes.c_int, [vp]),
      "bsg_comm_ranks": (ctypes.c_int, [vp]),
      "bsg_forest_merge": (ctypes.c_int, [vp, ctypes.c_int]),
      "bsg_forest_allreduce_stats": (ctypes.c_int, [vp, vp]),
      "bsg_forest_selftest": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_int, ctypes.c_int, ctypes.c_int, u64, vp, ctypes.c_int]),
  }
  for name, (res, args) in sig.items():
    fn = getattr(lib, name)
    fn.restype = res
    fn.argtypes = args
  return lib, sorted(sig)


lib, EXPORTS = _load()


def check(rc, engine=None):
  """Maps an error code to RuntimeError, the reference's error convention from Python
  (SpielFatalError -> SpielException -> RuntimeError, pyspiel.cc:71-78,839-840)."""
  if rc >= 0:
    return rc
  msg = lib.bsg_error_string(rc).decode()
  if engine:
    detail = lib.bsg_last_error(engine).decode()
    if detail:
      msg = "%s (%s)" % (msg, detail)
  raise RuntimeError("bsg: " + msg)
