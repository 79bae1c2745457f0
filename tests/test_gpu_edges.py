"""Edge cases of the playout path through the C ABI (SURVEY.md §8b "Error convention", §4):
terminal roots, ragged batches, single-root / many-rollout calls, empty input, step limits,
a root without legal actions, bad arguments.  Checked against the oracle where there is
something to compute and against the documented return codes where there is not."""
import ctypes

import numpy as np
import pytest

from oracle_lib import Oracle

pytestmark = pytest.mark.gpu
GAMES = ["tic_tac_toe", "connect_four", "breakthrough", "hex", "pentago", "othello", "y", "havannah", "go(board_size=9)"]


@pytest.fixture(scope="module")
def torch():
  import torch as t
  return t


@pytest.fixture(scope="module")
def bsg():
  import open_spiel_1_b200 as m
  return m


def random_history(oracle, game, rng, moves):
  """A history of up to `moves` random legal moves (stops at a terminal state)."""
  acts = []
  for _ in range(moves):
    dump = oracle.state_dump(game, acts)
    if dump["terminal"]:
      break
    la = dump["legal"]
    acts.append(int(la[rng.randint(len(la))]))
  return acts


def outputs(torch, n, dev):
  return (torch.full((n,), 99, dtype=torch.int8, device=dev),
          torch.full((n,), 99, dtype=torch.uint8, device=dev),
          torch.full((n,), 99, dtype=torch.uint8, device=dev),
          torch.zeros(4, dtype=torch.int64, device=dev))


@pytest.mark.parametrize("game", GAMES)
def test_terminal_and_ragged_roots(game, bsg, torch):
  """A batch mixing finished games, one-move-from-the-end positions and fresh boards: a
  terminal root is a playout of length 0 that returns the state's own result;
  every other playout equals the oracle's replay."""
  oracle = Oracle()
  rng = np.random.RandomState(8)
  hists = []
  for i in range(96):
    h = random_history(oracle, game, rng, 400)            # runs to the end: terminal
    if i % 3 == 1:
      h = h[:-1]                                          # one move before the end
    elif i % 3 == 2:
      h = h[:rng.randint(0, 4)]                           # (almost) fresh board
    hists.append(h)
  n, R = len(hists), 3
  eng = bsg.Engine(game, 0)
  roots = eng.to_device(bsg.pack_histories(game, hists))
  ret0, length, term, totals = outputs(torch, n * R, eng.torch_device)
  sums = torch.zeros(n, dtype=torch.int32, device=eng.torch_device)
  eng.playouts(roots, n, R, seed=31, first_playout=1000, ret0=ret0, length=length, terminal=term,
               totals=totals, root_sum_ret0=sums)
  torch.cuda.synchronize()
  o_len, o_ret, o_term, _ = oracle.replay_batch(game, hists, rollouts=R, seed=31, first_playout=1000)
  assert np.array_equal(ret0.cpu().numpy(), o_ret)
  assert np.array_equal(length.cpu().numpy(), o_len)
  assert np.array_equal(term.cpu().numpy(), o_term)
  for i in range(0, n, 3):                                # the finished games
    own = oracle.state_dump(game, hists[i])["returns"][0]
    assert (length.cpu().numpy()[i * R:(i + 1) * R] == 0).all()
    assert (ret0.cpu().numpy()[i * R:(i + 1) * R] == own).all()
  t = totals.cpu().numpy()
  assert t[0] == n * R and t[1] == o_len.sum() and t[2] == o_ret.sum() and t[3] == 0
  assert np.array_equal(sums.cpu().numpy(), o_ret.reshape(n, R).sum(1))
  eng.close()


@pytest.mark.parametrize("game", GAMES)
def test_one_root_many_rollouts_and_one_playout(game, bsg, torch):
  oracle = Oracle()
  h = random_history(oracle, game, np.random.RandomState(2), 5)
  eng = bsg.Engine(game, 0)
  root = eng.to_device(bsg.pack_histories(game, [h]))
  for R in (1, 1000):
    ret0, length, term, totals = outputs(torch, R, eng.torch_device)
    eng.playouts(root, 1, R, seed=4, first_playout=2**40 + 3, ret0=ret0, length=length,
                 terminal=term, totals=totals)
    torch.cuda.synchronize()
    o_len, o_ret, _, _ = oracle.replay_batch(game, [h], rollouts=R, seed=4, first_playout=2**40 + 3)
    assert np.array_equal(ret0.cpu().numpy(), o_ret) and np.array_equal(length.cpu().numpy(), o_len)
    assert term.cpu().numpy().all()
    # the host-buffer entry point gives the same sums
    tot, rr, rs = eng.playouts_host(bsg.pack_histories(game, [h]), 1, R, seed=4,
                                    first_playout=2**40 + 3, want_root_sums=True)
    assert tot[0] == R and tot[1] == o_len.sum() == rs[0] and tot[2] == o_ret.sum() == rr[0]
  eng.close()


def test_empty_input_and_bad_arguments(bsg, torch):
  """n_roots = 0 is a no-op on the device entry point; malformed calls come back as error
  codes with a message (never a crash, never an exception across the ABI)."""
  from open_spiel_1_b200 import _cabi as C
  eng = bsg.Engine("connect_four", 0)
  root = eng.to_device(bsg.pack_histories("connect_four", [[3]]))
  totals = torch.zeros(4, dtype=torch.int64, device=eng.torch_device)
  eng.playouts(root, 0, 1, totals=totals)                 # nothing to do
  torch.cuda.synchronize()
  assert totals.cpu().numpy().tolist() == [0, 0, 0, 0]
  with pytest.raises(RuntimeError):
    eng.playouts(root, 1, 0, totals=totals)               # rollouts < 1
  with pytest.raises(RuntimeError):
    eng.playouts(None, 1, 1, totals=totals)               # no roots
  with pytest.raises(RuntimeError):
    eng.playouts(root, 1 << 20, 1 << 13, totals=totals)   # >= 2^32 - 2^24 playouts in one call
  with pytest.raises(RuntimeError):
    eng.playouts_host(np.zeros(8, np.uint8), 0, 1)        # the host entry point needs >= 1 root
  msg = C.lib.bsg_last_error(eng._h).decode()
  assert msg                                              # the message of the last failure
  assert C.lib.bsg_playouts(None, None, 1, 1, 0, 0, None, None) == C.ERR_BAD_ARG
  with pytest.raises(RuntimeError):
    bsg.Engine("chess", 0)                                # unsupported game string
  with pytest.raises(RuntimeError):
    bsg.pack_histories("connect_four", [[3, 3, 3, 3, 3, 3, 3]])   # seventh stone in a column
  # the engine is still usable after the failures
  eng.playouts(root, 1, 4, totals=totals)
  torch.cuda.synchronize()
  assert totals.cpu().numpy()[0] == 4
  eng.close()


@pytest.mark.parametrize("game", ["connect_four", "breakthrough"])
def test_step_limits_stop_playouts_early(game, bsg, torch):
  """max_steps / per-playout step_limit (the synthetic-root generator of bench.py): a stopped
  playout reports terminal = 0, length = its limit and the state it stopped in."""
  oracle = Oracle()
  n = 512
  eng = bsg.Engine(game, 0)
  roots = eng.to_device(bsg.pack_histories(game, [[]] * n))
  limit = torch.from_numpy(np.random.RandomState(1).randint(0, 12, n).astype(np.uint8)).to(eng.torch_device)
  ret0, length, term, totals = outputs(torch, n, eng.torch_device)
  actions = torch.full((n, 16), -1, dtype=torch.int16, device=eng.torch_device)
  final = torch.zeros(n * eng.state_bytes, dtype=torch.uint8, device=eng.torch_device)
  eng.playouts(roots, n, 1, seed=12, ret0=ret0, length=length, terminal=term, totals=totals,
               step_limit=limit, actions=actions, final_states=final)
  torch.cuda.synchronize()
  lim, ln, acts = limit.cpu().numpy(), length.cpu().numpy(), actions.cpu().numpy()
  tm, rt = term.cpu().numpy(), ret0.cpu().numpy()
  hists = [[int(a) for a in acts[i, :ln[i]]] for i in range(n)]
  stopped = 0
  for i in range(n):                                       # the oracle under the same limit
    o_acts, o_term, o_ret = oracle.replay(game, [], seed=12, playout=i, max_steps=int(lim[i]))
    assert hists[i] == list(o_acts) and tm[i] == o_term and rt[i] == o_ret
    if not o_term:
      assert ln[i] == lim[i] and rt[i] == 0
      stopped += 1
  assert stopped > n // 2
  # the states they stopped in: legal histories, and exactly those positions
  assert np.array_equal(final.cpu().numpy(), bsg.pack_histories(game, hists))
  eng.close()


def test_root_without_legal_action_is_reported(bsg, torch):
  """A breakthrough root that is not terminal and has no legal action (only a malformed root
  can be: see the note in csrc/bsg_games.cuh) is counted in totals[3], plays no move, and the
  host entry point returns BSG_ERR_STUCK_LANE."""
  from open_spiel_1_b200 import _cabi as C
  eng = bsg.Engine("breakthrough", 0)
  good = bsg.pack_histories("breakthrough", [[]])
  packed = np.zeros(2 * 20, np.uint8)                     # capacity 2: planes of 16 B and 4 B
  boards = packed[:32].view(np.uint64)
  boards[0:2] = good[:16].view(np.uint64)                 # root 0: the initial position
  boards[2], boards[3] = 0, 0xFFFF000000000000            # root 1: black to move, no black piece
  meta = packed[32:].view(np.uint32)
  meta[0], meta[1] = good[16:20].view(np.uint32)[0], 0    # mover 0, not terminal
  roots = eng.to_device(packed)
  ret0, length, term, totals = outputs(torch, 2, eng.torch_device)
  eng.playouts(roots, 2, 1, seed=3, ret0=ret0, length=length, terminal=term, totals=totals)
  torch.cuda.synchronize()
  assert totals.cpu().numpy()[3] == 1 and totals.cpu().numpy()[0] == 2
  assert length.cpu().numpy()[1] == 0 and term.cpu().numpy()[1] == 0 and ret0.cpu().numpy()[1] == 0
  assert term.cpu().numpy()[0] == 1 and length.cpu().numpy()[0] > 0
  out = np.zeros(4, np.int64)
  rc = C.lib.bsg_playouts_host(eng._h, packed.ctypes.data, 2, 1, 3, 0, out.ctypes.data, None, None)
  assert rc == C.ERR_STUCK_LANE and out[3] == 1
  eng.close()
