"""Synthetic root states: states reached by d uniform-random legal moves from
the initial position, d ~ U{dmin..dmax}, never terminal.

Generated on the device by the playout kernel itself in its limited-steps mode: playout i
plays step_limit[i] = d_i moves under the Philox contract (seed, playout id i) and writes the
state it stops in; a playout that would end the game writes its last non-terminal state
instead.  With `want_histories` the action log comes back too, so the oracle can rebuild the
very same roots from histories (parity tests).
"""
import numpy as np

ROOT_DEPTHS = {  # cfg2 U{8..20}, cfg3 U{0..40}, cfg4 U{0..60}
    "tic_tac_toe": (0, 4),
    "connect_four": (8, 20),
    "breakthrough": (0, 40),
    "hex": (0, 60),
    "pentago": (0, 20),
    "othello": (0, 40),
    "y": (0, 80),
    "havannah": (0, 60),
    "go(board_size=9)": (0, 60),
}


def make_roots(engine, n, dmin, dmax, seed, want_histories=False, chunk=1 << 22, want_log=False):
  """Returns (states, histories): `states` is a device uint8 tensor holding n packed states
  (SoA, capacity n); `histories` is a list of action lists or None.  With `want_log` the
  histories come back as a dense action log instead of Python lists — (actions int16
  [n, max(dmax, 1)] numpy, root_len int32 [n] numpy): row i holds the root_len[i] moves that
  lead to root i — which is how BASELINE-size batches (16 Mi roots) reach a CPU checker."""
  import torch
  dev = engine.torch_device
  sb = engine.state_bytes
  init = engine.to_device(_initial(engine))
  rng = np.random.RandomState(seed & 0x7FFFFFFF)
  depth = rng.randint(dmin, dmax + 1, size=n).astype(np.uint8)
  d_depth = torch.from_numpy(depth).to(dev)
  states = torch.empty(n * sb, dtype=torch.uint8, device=dev)
  length = torch.empty(n, dtype=torch.uint8, device=dev)
  term = torch.empty(n, dtype=torch.uint8, device=dev)
  acts = None
  if want_histories or want_log:
    acts = torch.full((n, max(int(dmax), 1)), -1, dtype=torch.int16, device=dev)
  # one root (the initial position), n rollouts: playout i is root i of the result
  engine.playouts(init, 1, n, seed=seed, first_playout=0, step_limit=d_depth, length=length,
                  terminal=term, actions=acts, final_states=states, final_before_terminal=True)
  torch.cuda.synchronize(dev)
  histories = None
  if want_log:
    ln = length.cpu().numpy().astype(np.int32)
    histories = (acts.cpu().numpy(), ln - (term.cpu().numpy() != 0).astype(np.int32))
  elif want_histories:
    a = acts.cpu().numpy()
    ln = length.cpu().numpy().astype(np.int64)
    tm = term.cpu().numpy()
    histories = []
    for i in range(n):
      k = int(ln[i]) - (1 if tm[i] else 0)  # the terminal move was not kept
      histories.append(a[i, :k].astype(np.int64).tolist())
  return states, histories


def _initial(engine):
  from .engine import pack_histories
  return pack_histories(engine.gid, [[]])
