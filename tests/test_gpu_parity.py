"""GPU parity tests (run with -m gpu on the B200 box).  Everything goes through the C ABI.

Checker: the unmodified reference (oracle/_ref/libspiel_ref.so) when it travelled with the
repo, otherwise the C oracle.  Bar: bit-exact — action sequences, lengths, terminal flags,
integer returns, observation planes, legal masks.
"""
import os

import numpy as np
import pytest

from oracle_lib import Oracle, Reference, reference_path

pytestmark = pytest.mark.gpu

GAMES = ["tic_tac_toe", "connect_four", "breakthrough", "hex", "pentago", "othello", "y", "havannah", "go(board_size=9)"]
# replay-parity sizes: >= 1e5 ROOTS per game, two rollouts per root where
# the checker is cheap so that the rollouts > 1 indexing is in the same comparison
N_ROOTS = {g: 100000 for g in GAMES}
ROLLOUTS = {"tic_tac_toe": 2, "connect_four": 2, "breakthrough": 1, "hex": 1, "pentago": 2, "othello": 1, "y": 1, "havannah": 1, "go(board_size=9)": 1}


@pytest.fixture(scope="module")
def checker():
  if os.path.exists(reference_path()):
    return Reference()
  return Oracle()


@pytest.fixture(scope="module")
def torch():
  import torch as t
  assert t.cuda.is_available()
  return t


def _engine(game):
  import open_spiel_1_b200 as bsg
  return bsg.Engine(game, 0)


def _roots(eng, game, n, seed):
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  dmin, dmax = ROOT_DEPTHS[game]
  return make_roots(eng, n, dmin, dmax, seed, want_histories=True)


@pytest.mark.parametrize("game", GAMES)
def test_replay_bit_exact(game, checker, torch):
  """>= 1e5 playouts per game from synthetic mid-game roots: every action, length, terminal
  flag and integer return equals the checker's under the same Philox draws."""
  import open_spiel_1_b200 as bsg
  eng = _engine(game)
  n, R = N_ROOTS[game], ROLLOUTS[game]
  states, hist = _roots(eng, game, n, seed=20240611)
  # the device-made roots equal the host-packed roots of the same histories, byte for byte
  assert np.array_equal(states.cpu().numpy(), bsg.pack_histories(game, hist))
  dev = eng.torch_device
  stride = eng.max_len if game != "breakthrough" else 209
  ret0 = torch.zeros(n * R, dtype=torch.int8, device=dev)
  length = torch.zeros(n * R, dtype=torch.uint8, device=dev)
  term = torch.zeros(n * R, dtype=torch.uint8, device=dev)
  acts = torch.full((n * R, stride), -1, dtype=torch.int16, device=dev)
  totals = torch.zeros(4, dtype=torch.int64, device=dev)
  rs = torch.zeros(n, dtype=torch.int32, device=dev)
  rl = torch.zeros(n, dtype=torch.int32, device=dev)
  eng.playouts(states, n, R, seed=77, first_playout=5, ret0=ret0, length=length, terminal=term,
               actions=acts, totals=totals, root_sum_ret0=rs, root_sum_steps=rl)
  torch.cuda.synchronize()
  o_len, o_ret, o_term, o_acts = checker.replay_batch(game, hist, rollouts=R, seed=77,
                                                      first_playout=5, actions_stride=stride)
  assert np.array_equal(length.cpu().numpy().astype(np.int32), o_len)
  assert np.array_equal(ret0.cpu().numpy(), o_ret)
  assert np.array_equal(term.cpu().numpy(), o_term)
  assert np.array_equal(acts.cpu().numpy(), o_acts)
  t = totals.cpu().numpy()
  assert t.tolist() == [n * R, int(o_len.sum()), int(o_ret.astype(np.int64).sum()), 0]
  assert np.array_equal(rs.cpu().numpy(), o_ret.reshape(n, R).astype(np.int32).sum(1))
  assert np.array_equal(rl.cpu().numpy(), o_len.reshape(n, R).sum(1))
  # the fast kernel variant (no action log) gives the same per-playout results
  ret0b = torch.zeros_like(ret0)
  lengthb = torch.zeros_like(length)
  eng.playouts(states, n, R, seed=77, first_playout=5, ret0=ret0b, length=lengthb)
  torch.cuda.synchronize()
  assert torch.equal(ret0, ret0b) and torch.equal(length, lengthb)
  eng.close()


@pytest.mark.parametrize("game", GAMES)
def test_replay_with_explicit_r_table(game, checker, torch):
  """Replay mode driven by pre-generated per-step integers (action = legal[r mod n])."""
  eng = _engine(game)
  n = 2000
  states, hist = _roots(eng, game, n, seed=5)
  dev = eng.torch_device
  stride = 209 if game == "breakthrough" else eng.max_len
  rng = np.random.RandomState(99)
  table = rng.randint(0, 2**32, size=(n, stride), dtype=np.uint64).astype(np.uint32)
  d_table = torch.from_numpy(table.view(np.int32)).to(dev)
  ret0 = torch.zeros(n, dtype=torch.int8, device=dev)
  length = torch.zeros(n, dtype=torch.uint8, device=dev)
  acts = torch.full((n, stride), -1, dtype=torch.int16, device=dev)
  eng.playouts(states, n, 1, seed=1, ret0=ret0, length=length, actions=acts, r_table=d_table)
  torch.cuda.synchronize()
  g_acts, g_len, g_ret = acts.cpu().numpy(), length.cpu().numpy(), ret0.cpu().numpy()
  for i in range(0, n, 7):
    a, term, r = checker.replay(game, hist[i], r_table=table[i])
    assert g_len[i] == len(a) and g_ret[i] == r and term == 1
    assert g_acts[i, :len(a)].tolist() == a
  eng.close()


@pytest.mark.parametrize("game", GAMES)
def test_observations_and_masks_bit_exact(game, checker, torch):
  """ObservationTensor planes and LegalActionsMask of 1e5 states per game, every one of them
  against the COMPILED REFERENCE's State::ObservationTensor / LegalActions in the same process
  (ref_state_dump_log; hex edge-label planes hex.cc:55-121,149-184,266-277), then 1e4 of them
  also against the C restatement, then terminal states (the reference keeps the *Win marks)."""
  import open_spiel_1_b200 as bsg
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  oracle = Oracle()
  eng = _engine(game)
  n = 100000
  dmin, dmax = ROOT_DEPTHS[game]
  states, (log, lens) = make_roots(eng, n, dmin, dmax, seed=42, want_log=True)
  dev = eng.torch_device
  pl = (np.arange(n) % 2).astype(np.uint8)
  players = torch.from_numpy(pl).to(dev)
  obs, mask = eng.observations(states, n, players, want_mask=True)
  torch.cuda.synchronize()
  obs = obs.cpu().numpy().reshape(n, -1)
  mask = mask.cpu().numpy()
  assert set(np.unique(obs).tolist()) <= {0.0, 1.0}
  cells = {"y": 190, "havannah": 169}.get(game, eng.obs_size // eng.obs_shape[0])   # y, havannah: cut-off corners
  if game.startswith("go"):   # black / white / empty planes, plus a plane of ones when white is to move
    assert np.all(obs[:, :3 * 81].sum(1) == 81) and set(np.unique(obs[:, 3 * 81:].sum(1)).tolist()) <= {0.0, 81.0}
  else:
    assert np.all(obs.sum(1) == cells)  # exactly one plane per board cell
  obs8 = obs.astype(np.uint8)
  if isinstance(checker, Reference):
    d = checker.state_dump_log(game, log, lens, pl, threads=os.cpu_count() or 1)
    assert np.array_equal(obs8, d["obs"])
    assert np.array_equal(mask, d["mask"])
    assert not d["terminal"].any()
  hist = [log[i, :lens[i]].astype(np.int64).tolist() for i in range(n)]
  step = 1 if not isinstance(checker, Reference) else 10
  for i in range(0, n, step):             # the C restatement
    d = oracle.state_dump(game, hist[i], i % 2)
    assert np.array_equal(obs8[i], d["obs"]), i
    assert np.nonzero(mask[i])[0].tolist() == d["legal"], i
  # host-buffer entry point gives the same bytes
  k = 1000
  packed = bsg.pack_histories(game, hist[:k])
  obs_h, mask_h = eng.observations_host(packed, k, pl[:k], True)
  assert np.array_equal(obs_h.reshape(k, -1), obs[:k]) and np.array_equal(mask_h, mask[:k])
  # terminal states: 3000 finished games (hex / y: the winning cell carries both edge labels)
  term_hist = []
  for p in range(300):
    a, _, _ = oracle.replay(game, [], seed=9, playout=p)
    term_hist.append(a)
  _, _, _, t_acts = oracle.replay_batch(game, [[]], rollouts=3000, seed=10, actions_stride=256)
  t_lens = (t_acts >= 0).sum(1).astype(np.int32)
  term_hist += [t_acts[i, :t_lens[i]].astype(np.int64).tolist() for i in range(3000)]
  nt = len(term_hist)
  packed = bsg.pack_histories(game, term_hist)
  tp = (np.arange(nt) % 2).astype(np.uint8)
  obs_t, mask_t = eng.observations_host(packed, nt, tp, True)
  obs_t = obs_t.reshape(nt, -1).astype(np.uint8)
  assert mask_t.sum() == 0
  if isinstance(checker, Reference):
    t_log = np.full((nt, 256), -1, np.int16)
    for i, h in enumerate(term_hist):
      t_log[i, :len(h)] = h
    d = checker.state_dump_log(game, t_log, np.array([len(h) for h in term_hist], np.int32), tp,
                               threads=os.cpu_count() or 1)
    assert d["terminal"].all() and np.array_equal(obs_t, d["obs"]) and d["mask"].sum() == 0
  for i in range(300):
    d = oracle.state_dump(game, term_hist[i], int(tp[i]))
    assert np.array_equal(obs_t[i], d["obs"])
  eng.close()


@pytest.mark.parametrize("game", GAMES)
def test_native_rng_win_rates_within_binomial_ci(game, checker, torch):
  """>= 1e6 playouts from the initial state: player-0 win / draw / loss rates within a
  99.9 % binomial interval (z = 3.29) of the checker's rates under ITS OWN RNG stream
  (std::mt19937 for the reference; distinct Philox seed for the oracle)."""
  import open_spiel_1_b200 as bsg
  eng = _engine(game)
  dev = eng.torch_device
  n_gpu = 1 << 20
  root = eng.to_device(bsg.pack_histories(game, [[]]))
  ret0 = torch.zeros(n_gpu, dtype=torch.int8, device=dev)
  length = torch.zeros(n_gpu, dtype=torch.uint8, device=dev)
  eng.playouts(root, 1, n_gpu, seed=20240611, ret0=ret0, length=length)
  torch.cuda.synchronize()
  r = ret0.cpu().numpy()
  # >= 1e6 reference playouts per game where the reference plays them in
  # seconds (all host cores); breakthrough / hex / y at what ~5 s of the reference buys
  n_ref = {"tic_tac_toe": 1 << 20, "connect_four": 1 << 20, "breakthrough": 200000, "hex": 100000,
           "pentago": 400000, "othello": 40000, "y": 60000, "havannah": 60000, "go(board_size=9)": 60000}[game]
  if isinstance(checker, Reference):
    # reference's own mt19937 stream; per-outcome counts via replay batch are Philox-tied, so
    # use random_playouts for the mean and replay_batch (different seed) for the outcome split
    _, _, sum_ret = checker.random_playouts(game, [[]], n_ref, seed=4242, threads=os.cpu_count() or 1)
    mean_ref = sum_ret / n_ref
  else:
    _, sum_ret = checker.random_playouts(game, [[]], n_ref, seed=4242)
    mean_ref = sum_ret / n_ref
  mean_gpu = float(r.astype(np.float64).mean())
  var = float(r.astype(np.float64).var()) + 1e-9
  half = 3.29 * np.sqrt(var / n_gpu + var / n_ref)
  assert abs(mean_gpu - mean_ref) <= half, (mean_gpu, mean_ref, half)
  # outcome split against an independently seeded checker replay
  if isinstance(checker, Reference):
    o_len, o_ret, _, _ = checker.replay_batch(game, [[]], rollouts=n_ref, seed=987654321,
                                              threads=os.cpu_count() or 1)
  else:
    o_len, o_ret, _, _ = checker.replay_batch(game, [[]], rollouts=n_ref, seed=987654321)
  for outcome in (1, 0, -1):
    p_gpu = float((r == outcome).mean())
    p_ref = float((o_ret == outcome).mean())
    p = (p_gpu * n_gpu + p_ref * n_ref) / (n_gpu + n_ref)
    half = 3.29 * np.sqrt(p * (1 - p) * (1.0 / n_gpu + 1.0 / n_ref)) + 1e-12
    assert abs(p_gpu - p_ref) <= half, (outcome, p_gpu, p_ref, half)
  mean_len_gpu = float(length.cpu().numpy().astype(np.float64).mean())
  assert abs(mean_len_gpu - float(o_len.mean())) < 0.05 * float(o_len.mean())
  eng.close()


@pytest.mark.parametrize("game,n,max_len,min_len,n_check", [
    ("connect_four", 1 << 24, 42, 1, 1 << 24),     # BASELINE cfg2: 16 Mi playouts, ALL replayed
    ("breakthrough", 1 << 22, 209, 1, 1 << 21),    # cfg3: 4 Mi playouts per GPU, 2 Mi replayed
    ("hex", 1 << 20, 121, 1, 1 << 20),             # cfg4: 1 Mi playouts per GPU, all replayed
    ("pentago", 1 << 22, 36, 1, 1 << 20),
    ("othello", 1 << 22, 120, 1, 1 << 16),
    ("y", 1 << 20, 190, 1, 1 << 18),
    ("havannah", 1 << 20, 169, 1, 1 << 17),
    ("go(board_size=9)", 1 << 18, 162, 1, 1 << 17),
])
def test_full_size_properties(game, n, max_len, min_len, n_check, checker, torch):
  """BASELINE configuration sizes.  (1) The lean kernel's per-playout results (length, return)
  against the COMPILED REFERENCE replaying the same playouts from the same roots — all 16 Mi of
  cfg2, 2 Mi of cfg3, all 1 Mi of cfg4 (ref_replay_log, all host cores; the roots reach it as
  the dense action log of the root generator).  (2) Size-independent properties: checksum of
  checksums (totals vs per-playout outputs), every playout terminal, length bounds,
  idempotence, invariance under splitting the launch."""
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  from oracle_lib import ragged_from_log
  eng = _engine(game)
  dev = eng.torch_device
  dmin, dmax = ROOT_DEPTHS[game]
  states, (log, root_len) = make_roots(eng, n, dmin, dmax, seed=20240611 + 2, want_log=True)
  ret0 = torch.zeros(n, dtype=torch.int8, device=dev)
  length = torch.zeros(n, dtype=torch.uint8, device=dev)
  term = torch.zeros(n, dtype=torch.uint8, device=dev)
  totals = torch.zeros(4, dtype=torch.int64, device=dev)
  eng.playouts(states, n, 1, seed=1, ret0=ret0, length=length, terminal=term, totals=totals)
  torch.cuda.synchronize()
  t = totals.cpu().numpy()
  assert t[0] == n and t[3] == 0
  assert int(length.to(torch.int64).sum()) == t[1]          # checksum of per-playout lengths
  assert int(ret0.to(torch.int64).sum()) == t[2]
  assert bool(term.all())                                    # every playout reached a terminal
  assert int(length.max()) <= max_len and int(length.min()) >= min_len
  if game in ("breakthrough", "hex", "y"):
    assert bool((ret0 != 0).all())                           # no draws in breakthrough / hex
  # idempotence / schedule independence: a split launch with first_playout offsets reproduces
  # the same per-playout results (every SoA plane must be sliced for the second half)
  ret0b = torch.zeros_like(ret0)
  lengthb = torch.zeros_like(length)
  half = n // 2
  import open_spiel_1_b200 as bsg
  from open_spiel_1_b200 import _cabi as C
  host = states.cpu().numpy()
  sb = eng.state_bytes
  planes = bsg.engine.plane_bytes(eng.gid)
  assert sum(planes) == sb

  def take(lo, hi):
    parts, off = [], 0
    for pb in planes:
      parts.append(host[off + lo * pb: off + hi * pb])
      off += n * pb
    return np.concatenate(parts)

  lo_half = eng.to_device(take(0, half))
  hi_half = eng.to_device(take(half, n))
  eng.playouts(lo_half, half, 1, seed=1, first_playout=0, ret0=ret0b[:half], length=lengthb[:half])
  eng.playouts(hi_half, n - half, 1, seed=1, first_playout=half, ret0=ret0b[half:], length=lengthb[half:])
  torch.cuda.synchronize()
  assert torch.equal(ret0, ret0b) and torch.equal(length, lengthb)
  # (1) the compiled reference replays the playouts themselves
  g_len, g_ret = length.cpu().numpy(), ret0.cpu().numpy()
  if isinstance(checker, Reference):
    o_len, o_ret, o_term = checker.replay_log(game, log[:n_check], root_len[:n_check], seed=1,
                                              first_playout=0, threads=os.cpu_count() or 1)
  else:   # no reference library on this box: the C restatement on a smaller prefix
    n_check = min(n_check, 1 << 17)
    o_len, o_ret, o_term, _ = checker.replay_batch(game, ragged_from_log(log[:n_check], root_len[:n_check]),
                                                   seed=1, first_playout=0)
  assert np.array_equal(g_len[:n_check].astype(np.int32), o_len)
  assert np.array_equal(g_ret[:n_check], o_ret)
  assert o_term.all()
  eng.close()


@pytest.mark.parametrize("game,n", [("connect_four", 800000), ("breakthrough", 330000),
                                    ("hex", 110000), ("tic_tac_toe", 1100000),
                                    ("pentago", 300000), ("othello", 330000), ("y", 60000), ("havannah", 70000), ("go(board_size=9)", 25000)])
def test_host_buffer_chunked_path_equals_device_path(game, n, torch):
  """bsg_playouts_host cuts large inputs into chunks on two streams (H2D overlapping the
  kernel); totals and per-root sums must equal one device-buffer launch over the same roots,
  for pinned and for pageable host memory, and for every SoA plane layout."""
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  eng = _engine(game)
  dev = eng.torch_device
  dmin, dmax = ROOT_DEPTHS[game]
  states, _ = make_roots(eng, n, dmin, dmax, seed=77)
  assert states.numel() >= 4 << 20                       # several chunks
  R = 2
  totals = torch.zeros(4, dtype=torch.int64, device=dev)
  rr = torch.zeros(n, dtype=torch.int32, device=dev)
  rs = torch.zeros(n, dtype=torch.int32, device=dev)
  eng.playouts(states, n, R, seed=5, first_playout=123, root_sum_ret0=rr, root_sum_steps=rs,
               totals=totals)
  torch.cuda.synchronize()
  want_t, want_r, want_s = totals.cpu().numpy(), rr.cpu().numpy(), rs.cpu().numpy()
  assert want_t[0] == n * R and want_t[3] == 0
  pageable = states.cpu().numpy()
  pinned = torch.empty(states.numel(), dtype=torch.uint8, pin_memory=True)
  pinned.copy_(states)
  torch.cuda.synchronize()
  for host in (pageable, pinned.numpy()):
    t, r, s = eng.playouts_host(host, n, R, seed=5, first_playout=123, want_root_sums=True)
    assert np.array_equal(t, want_t)
    assert np.array_equal(r, want_r) and np.array_equal(s, want_s)
  eng.close()


@pytest.mark.parametrize("game,n", [("connect_four", 700000), ("breakthrough", 250000), ("hex", 70000),
                                    ("tic_tac_toe", 1100000), ("connect_four", 37), ("hex", 5)])
def test_host_results_entry_point_equals_device_path(game, n, torch):
  """bsg_playouts_host_results: per-playout return and move count in host memory (several
  chunks on two streams with the copy back behind each kernel; one chunk for the small cases)
  equal the device-buffer launch, for pageable and page-locked caller buffers."""
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  eng = _engine(game)
  dev = eng.torch_device
  dmin, dmax = ROOT_DEPTHS[game]
  states, _ = make_roots(eng, n, dmin, dmax, seed=78)
  R = 2
  ret0 = torch.zeros(n * R, dtype=torch.int8, device=dev)
  length = torch.zeros(n * R, dtype=torch.uint8, device=dev)
  totals = torch.zeros(4, dtype=torch.int64, device=dev)
  eng.playouts(states, n, R, seed=6, first_playout=321, ret0=ret0, length=length, totals=totals)
  torch.cuda.synchronize()
  want_t, want_r, want_l = totals.cpu().numpy(), ret0.cpu().numpy(), length.cpu().numpy()
  host = states.cpu().numpy()
  t, r, l = eng.playouts_host_results(host, n, R, seed=6, first_playout=321)       # pageable in and out
  assert np.array_equal(t, want_t) and np.array_equal(r, want_r) and np.array_equal(l, want_l)
  pin_in = torch.empty(states.numel(), dtype=torch.uint8, pin_memory=True)
  pin_in.copy_(states)
  pr = torch.zeros(n * R, dtype=torch.int8, pin_memory=True)
  pl = torch.zeros(n * R, dtype=torch.uint8, pin_memory=True)
  torch.cuda.synchronize()
  t, r, l = eng.playouts_host_results(pin_in.numpy(), n, R, seed=6, first_playout=321,
                                      ret0=pr.numpy(), length=pl.numpy())          # page-locked, in place
  assert np.array_equal(t, want_t)
  assert np.array_equal(pr.numpy(), want_r) and np.array_equal(pl.numpy(), want_l)
  eng.close()


@pytest.mark.parametrize("n,R", [(900001, 1), (5, 3), (70000, 2)])
def test_compact_root_entry_point_equals_device_path(n, R, torch):
  """bsg_playouts_host_compact: connect_four roots as 49 bits each — totals and per-playout
  results equal the device launch over the packed roots, terminal roots included, for pageable
  and page-locked buffers; a game without a compact format takes its packed roots."""
  import open_spiel_1_b200 as bsg
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  eng = _engine("connect_four")
  dev = eng.torch_device
  states, _ = make_roots(eng, n, 0, 42, seed=81)            # depth 0..42: long roots end by themselves
  oracle = Oracle()
  host = states.cpu().numpy().copy()
  for i in range(min(n, 4)):                                 # and a few finished games for certain
    a, _, _ = oracle.replay("connect_four", [], seed=3, playout=i)
    host[8 * i: 8 * i + 8] = bsg.pack_histories("connect_four", [a])
  states = eng.to_device(host)
  ret0 = torch.zeros(n * R, dtype=torch.int8, device=dev)
  length = torch.zeros(n * R, dtype=torch.uint8, device=dev)
  totals = torch.zeros(4, dtype=torch.int64, device=dev)
  eng.playouts(states, n, R, seed=6, first_playout=55, ret0=ret0, length=length, totals=totals)
  torch.cuda.synchronize()
  comp = eng.compact_roots(host, n)
  assert comp.nbytes == (49 * n + 7) // 8 and np.array_equal(eng.expand_roots(comp, n), host)
  t, r, l = eng.playouts_host_compact(comp, n, R, seed=6, first_playout=55, want_results=True)
  assert np.array_equal(t, totals.cpu().numpy())
  assert np.array_equal(r, ret0.cpu().numpy()) and np.array_equal(l, length.cpu().numpy())
  t2, _, _ = eng.playouts_host_compact(comp, n, R, seed=6, first_playout=55)       # totals only
  assert np.array_equal(t2, t)
  pin = torch.empty(comp.nbytes, dtype=torch.uint8, pin_memory=True)
  pin.numpy()[:] = comp
  pr = torch.zeros(n * R, dtype=torch.int8, pin_memory=True)
  pl = torch.zeros(n * R, dtype=torch.uint8, pin_memory=True)
  t3, _, _ = eng.playouts_host_compact(pin.numpy(), n, R, seed=6, first_playout=55, ret0=pr.numpy(), length=pl.numpy())
  assert np.array_equal(t3, t) and np.array_equal(pr.numpy(), r) and np.array_equal(pl.numpy(), l)
  eng.close()
  e2 = _engine("hex")
  st, _ = make_roots(e2, 1000, 0, 60, seed=2)
  hh = st.cpu().numpy()
  want, _, _ = e2.playouts_host(hh, 1000, 2, seed=4)
  got, _, _ = e2.playouts_host_compact(e2.compact_roots(hh, 1000), 1000, 2, seed=4)
  assert np.array_equal(want, got)
  e2.close()


@pytest.mark.parametrize("game", GAMES)
def test_small_host_calls_equal_device_path(game, torch):
  """bsg_playouts_host below 64 KB of roots takes the lean path (one stream, one memset for
  totals + work counter + root sums, one copy back): same totals and per-root sums as the
  device launch, call after call on one engine, with and without root sums, and next to a
  large chunked call."""
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  eng = _engine(game)
  dev = eng.torch_device
  dmin, dmax = ROOT_DEPTHS[game]
  for n, R in ((1, 1), (1, 64), (33, 5), (64 * 1024 // eng.state_bytes, 2), (64 * 1024 // eng.state_bytes + 1, 2)):
    states, _ = make_roots(eng, n, dmin, dmax, seed=100 + n)
    totals = torch.zeros(4, dtype=torch.int64, device=dev)
    rr = torch.zeros(n, dtype=torch.int32, device=dev)
    rs = torch.zeros(n, dtype=torch.int32, device=dev)
    eng.playouts(states, n, R, seed=8, first_playout=77, root_sum_ret0=rr, root_sum_steps=rs, totals=totals)
    torch.cuda.synchronize()
    host = states.cpu().numpy()
    for _ in range(2):
      t, r, s_ = eng.playouts_host(host, n, R, seed=8, first_playout=77, want_root_sums=True)
      assert np.array_equal(t, totals.cpu().numpy())
      assert np.array_equal(r, rr.cpu().numpy()) and np.array_equal(s_, rs.cpu().numpy())
      t, r, s_ = eng.playouts_host(host, n, R, seed=8, first_playout=77)
      assert np.array_equal(t, totals.cpu().numpy()) and r is None
  eng.close()


@pytest.mark.parametrize("game", GAMES)
def test_trajectory_records_match_oracle(game, torch):
  """Engine.trajectories: every decision point of every playout — observation of the player to
  move, legal mask, action taken — equals the oracle's replay of the same playout ids
  (the AlphaZero actor's record, alpha_zero.py:196-228)."""
  import open_spiel_1_b200 as bsg
  oracle = Oracle()
  eng = _engine(game)
  roots = []
  for p in range(12):
    acts, _, _ = oracle.replay(game, [], seed=5, playout=p, max_steps=(3 * p) % 11)
    roots.append(acts if not oracle.state_dump(game, acts)["terminal"] else acts[:-1])
  R = 3
  d_roots = eng.to_device(bsg.pack_histories(game, roots))
  tr = eng.trajectories(d_roots, len(roots), R, seed=21, first_playout=500)
  torch.cuda.synchronize()
  o_len, o_ret, o_term, o_acts = oracle.replay_batch(game, roots, rollouts=R, seed=21,
                                                     first_playout=500, actions_stride=256)
  assert np.array_equal(tr["length"].cpu().numpy(), o_len)
  assert np.array_equal(tr["ret0"].cpu().numpy(), o_ret)
  assert tr["total"] == int(o_len.sum())
  obs = tr["obs"].cpu().numpy().reshape(tr["total"], -1).astype(np.uint8)
  mask = tr["mask"].cpu().numpy()
  actions = tr["actions"].cpu().numpy()
  owner, step = tr["playout"].cpu().numpy(), tr["step"].cpu().numpy()
  rng = np.random.RandomState(1)
  picks = rng.choice(tr["total"], size=min(tr["total"], 150), replace=False)
  for i in picks:
    p, t = int(owner[i]), int(step[i])
    hist = list(roots[p // R]) + [int(a) for a in o_acts[p, :t]]
    assert actions[i] == o_acts[p, t]
    d = oracle.state_dump(game, hist)
    dd = oracle.state_dump(game, hist, d["current_player"])
    assert np.array_equal(obs[i], dd["obs"]), (game, p, t)
    assert np.nonzero(mask[i])[0].tolist() == dd["legal"], (game, p, t)
  eng.close()


def test_device_replay_buffer_takes_trajectories(torch):
  """trajectory records go straight into a device replay buffer with the
  semantics of the reference's Buffer (alpha_zero.py:83-104: keeps the newest, counts what
  passed through, samples without replacement) and TrainInput fields (model.py:80-91)."""
  import open_spiel_1_b200 as bsg
  oracle = Oracle()
  game = "connect_four"
  eng = _engine(game)
  roots = eng.to_device(bsg.pack_histories(game, [[3, 3], [0, 1, 2]]))
  rec = eng.trajectories(roots, 2, rollouts=64, seed=5)
  total = rec["total"]
  buf = bsg.DeviceReplayBuffer(total + 100, eng.obs_size, eng.num_actions, eng.torch_device)
  buf.extend_from_trajectories(rec)
  assert len(buf) == total == buf.total_seen
  got = buf.newest(total)
  assert got.observation.is_cuda and got.observation.shape == (total, eng.obs_size)
  assert torch.equal(got.observation, rec["obs"].reshape(total, -1))
  assert torch.equal(got.legals_mask, rec["mask"].to(torch.bool))
  # policy = uniform over the legal actions; value = the playout's return for player 0
  n_legal = got.legals_mask.sum(dim=1, keepdim=True).to(torch.float32)
  assert torch.equal(got.policy, got.legals_mask.to(torch.float32) / n_legal)
  o_len, o_ret, _, _ = oracle.replay_batch(game, [[3, 3], [0, 1, 2]], rollouts=64, seed=5)
  want_value = np.repeat(o_ret.astype(np.float32), o_len)
  assert np.array_equal(got.value.cpu().numpy().ravel(), want_value)
  batch = buf.sample(256, generator=torch.Generator(device=eng.torch_device).manual_seed(1))
  assert batch.value.shape == (256, 1) and batch.policy.shape == (256, eng.num_actions)
  assert torch.allclose(batch.policy.sum(dim=1), torch.ones(256, device=eng.torch_device))
  # the oldest fall out, exactly like data[:-max_size] = [] in the reference
  small = bsg.DeviceReplayBuffer(100, eng.obs_size, eng.num_actions, eng.torch_device)
  small.extend_from_trajectories(rec)
  small.extend(got.observation[:30], got.legals_mask[:30], got.policy[:30], got.value[:30])
  assert len(small) == 100 and small.total_seen == total + 30
  tail = small.newest(100)
  want_obs = torch.cat([got.observation[total - 70:], got.observation[:30]])
  assert torch.equal(tail.observation, want_obs)
  with pytest.raises(ValueError):
    small.sample(101)
  eng.close()


@pytest.mark.parametrize("game,n,R", [("hex", 12500, 8), ("y", 6250, 16)])
def test_value_only_playouts_return_the_exact_results(game, n, R, torch):
  """bsg_playout_io.value_only (hex, y): plain stones to a full board and one reading of the
  winner give, playout for playout, the ret0 / per-root sums / totals[2] of the exact launch
  (which the tests above tie to the oracle); the moves are the empty cells of the root.
  Terminal roots are in the batch; other games refuse the flag."""
  import open_spiel_1_b200 as bsg
  eng = _engine(game)
  dev = eng.torch_device
  states, hist = _roots(eng, game, n, seed=99)
  # make the last roots terminal ones
  oracle = Oracle()
  k = 64
  for i in range(k):
    a, _, _ = oracle.replay(game, [], seed=8, playout=i)
    hist[n - 1 - i] = a
  states = eng.to_device(bsg.pack_histories(game, hist))
  outs = []
  for vo in (False, True):
    ret0 = torch.zeros(n * R, dtype=torch.int8, device=dev)
    length = torch.zeros(n * R, dtype=torch.uint8, device=dev)
    term = torch.zeros(n * R, dtype=torch.uint8, device=dev)
    totals = torch.zeros(4, dtype=torch.int64, device=dev)
    rs = torch.zeros(n, dtype=torch.int32, device=dev)
    rl = torch.zeros(n, dtype=torch.int32, device=dev)
    eng.playouts(states, n, R, seed=31, first_playout=1 << 33, ret0=ret0, length=length, terminal=term,
                 totals=totals, root_sum_ret0=rs, root_sum_steps=rl, value_only=vo)
    torch.cuda.synchronize()
    outs.append((ret0, length, term, totals.cpu().numpy(), rs, rl))
  (r_a, l_a, t_a, tot_a, rs_a, rl_a), (r_b, l_b, t_b, tot_b, rs_b, rl_b) = outs
  assert torch.equal(r_a, r_b) and torch.equal(rs_a, rs_b)
  assert bool((r_b != 0).all()) and bool(t_b.all())
  assert tot_a[0] == tot_b[0] == n * R and tot_a[2] == tot_b[2] and tot_b[3] == 0
  cells = {"hex": 121, "y": 190}[game]
  empties = np.array([0 if i >= n - k else cells - len(h) for i, h in enumerate(hist)], np.int64)
  assert np.array_equal(l_b.cpu().numpy().reshape(n, R).astype(np.int64), np.repeat(empties[:, None], R, 1))
  assert np.array_equal(rl_b.cpu().numpy().astype(np.int64), empties * R) and tot_b[1] == int(empties.sum()) * R
  assert bool((l_a.to(torch.int64) <= l_b.to(torch.int64)).all())
  # totals-only launch (what the evaluator asks for)
  totals = torch.zeros(4, dtype=torch.int64, device=dev)
  eng.playouts(states, n, R, seed=31, first_playout=1 << 33, totals=totals, value_only=True)
  torch.cuda.synchronize()
  assert np.array_equal(totals.cpu().numpy(), tot_b)
  eng.close()
  e2 = _engine("connect_four")
  root = e2.to_device(bsg.pack_histories("connect_four", [[]]))
  with pytest.raises(RuntimeError):
    e2.playouts(root, 1, 4, totals=torch.zeros(4, dtype=torch.int64, device=dev), value_only=True)
  with pytest.raises(RuntimeError):
    eng2 = _engine(game)
    eng2.playouts(states, n, 1, totals=torch.zeros(4, dtype=torch.int64, device=dev), max_steps=3, value_only=True)
  e2.close()


@pytest.mark.parametrize("game,rollouts", [("connect_four", 96), ("connect_four", 128), ("connect_four", 160),
                                           ("connect_four", 256), ("go(board_size=9)", 130)])
def test_per_root_playout_bases_give_the_plain_launchs_playouts(game, rollouts, torch):
  """The launch with a playout-id base per root (forest leaves, evaluator batches) finds a
  playout's root from the chunk's first root and a comparison (rollouts >= 128) or a division
  (below); with base[i] = F + i * rollouts it must play exactly the playouts of the plain launch
  — the other code path — result for result, chunks that straddle two roots included."""
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  eng = _engine(game)
  dev = eng.torch_device
  n_roots, R, F = 777, rollouts, (1 << 35) + 12345
  dmin, dmax = ROOT_DEPTHS[game]
  states, _ = make_roots(eng, n_roots, dmin, dmax, seed=77)
  n = n_roots * R
  outs = []
  for with_base in (False, True):
    ret0 = torch.zeros(n, dtype=torch.int8, device=dev)
    length = torch.zeros(n, dtype=torch.uint8, device=dev)
    sums = torch.zeros(n_roots, dtype=torch.int32, device=dev)
    steps = torch.zeros(n_roots, dtype=torch.int32, device=dev)
    totals = torch.zeros(4, dtype=torch.int64, device=dev)
    base = None
    if with_base:
      base = (torch.arange(n_roots, dtype=torch.int64, device=dev) * R + F).view(torch.int64)
    eng.playouts(states, n_roots, R, seed=5, first_playout=0 if with_base else F, ret0=ret0, length=length,
                 root_sum_ret0=sums, root_sum_steps=steps, totals=totals, root_playout_base=base)
    torch.cuda.synchronize()
    outs.append((ret0, length, sums, steps, totals))
  for a, b in zip(*outs):
    assert torch.equal(a, b)
  assert int(outs[0][4][0]) == n
  eng.close()
