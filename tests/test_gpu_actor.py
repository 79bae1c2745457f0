"""Self-play actor on the forest (actor.ForestActor, csrc/bsg_actor.cuh) against the
reference's actor loop (python/algorithms/alpha_zero/alpha_zero.py:196-228) played the
reference's way: one game at a time, one search per move by the host MCTSBot mirror (which
tests/test_reference_python.py pins to the reference's own mcts.py), records filed as the
learner files them (alpha_zero.py:366-369)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
UCT_C = 2.0


@pytest.fixture(scope="module")
def bsg():
  import open_spiel_1_b200 as m
  return m


@pytest.mark.parametrize("game,n_games,sims,rollouts,batch,drop,solve", [
    ("connect_four", 5, 48, 4, 4, 6, False),
    ("tic_tac_toe", 7, 60, 2, 1, 2, True),        # solver proofs decide best_child
    ("breakthrough", 3, 24, 2, 4, 1000, False),   # every move sampled
    ("hex", 2, 20, 2, 4, 0, False),               # every move best_child; 9 planes per record
    ("othello", 2, 16, 2, 4, 8, False),           # passes
])
def test_actor_records_equal_the_reference_loop(bsg, game, n_games, sims, rollouts, batch, drop, solve):
  import torch
  from open_spiel_1_b200.actor import ForestActor, reference_style_game
  g = bsg.load_game(game)
  eng = bsg.Engine(game, 0)
  ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=rollouts, seed=31)
  actor = ForestActor(g, ev, n_games, uct_c=UCT_C, max_simulations=sims, temperature=1.0,
                      temperature_drop=drop, solve=solve, seed=17, sample_seed=99, leaf_batch=batch,
                      max_moves=g.max_game_length() if game != "othello" else 128)
  buf = bsg.DeviceReplayBuffer(100000, eng.obs_size, eng.num_actions, eng.torch_device)
  out = actor.play_batch(buffer=buf, want_extras=True)
  total = out["total"]
  assert len(buf) == total == buf.total_seen == int(out["lengths"].sum())
  obs = out["observation"].cpu().numpy()
  mask = out["legals_mask"].cpu().numpy()
  policy = out["policy"].cpu().numpy()
  value = out["value"].cpu().numpy().ravel()
  action = out["action"].cpu().numpy()
  player = out["player"].cpu().numpy()
  est = out["estimate"].cpu().numpy()
  # the buffer holds the very same rows
  got = buf.newest(total)
  assert torch.equal(got.observation, out["observation"]) and torch.equal(got.policy, out["policy"])
  assert torch.equal(got.legals_mask, out["legals_mask"]) and torch.equal(got.value, out["value"])
  row = 0
  for v in range(n_games):
    ev_v = bsg.GpuRolloutEvaluator(eng, n_rollouts=rollouts, seed=31)
    ev_v.next_playout = v << 40
    bot = bsg.MCTSBot(g, UCT_C, sims, ev_v, solve=solve, random_state=bsg.PhiloxShuffle(17 + v),
                      leaf_batch=batch)
    records, returns = reference_style_game(g, bot, 1.0, drop, 99, v)
    assert len(records) == out["lengths"][v] and int(returns[0]) == out["returns"][v]
    assert out["offsets"][v] == row
    for (r_obs, r_player, r_mask, r_action, r_policy, r_value) in records:
      assert action[row] == r_action and player[row] == r_player, (game, v, row)
      assert np.array_equal(obs[row], np.asarray(r_obs, np.float32))
      assert np.array_equal(mask[row], np.asarray(r_mask, bool))
      assert np.array_equal(policy[row], r_policy.astype(np.float32))
      assert est[row] == np.float32(r_value)
      assert value[row] == returns[0]
      row += 1
  assert row == total
  # a second batch continues the game ids and the trees' streams: different games, valid records
  out2 = actor.play_batch()
  assert out2["total"] > 0 and actor.games_played == 2 * n_games
  assert torch.all(out2["policy"].sum(1).sub(1).abs() < 1e-5)
  actor.close()
  eng.close()


def test_actor_fills_a_wrapping_buffer(bsg):
  """Records beyond the ring's end take the extend() path: the newest survive, as
  data[:-max_size] = [] in the reference's Buffer (alpha_zero.py:83-104)."""
  import torch
  from open_spiel_1_b200.actor import ForestActor
  g = bsg.load_game("tic_tac_toe")
  eng = bsg.Engine("tic_tac_toe", 0)
  ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=2, seed=3)
  actor = ForestActor(g, ev, 64, max_simulations=16, temperature_drop=3, leaf_batch=4)
  buf = bsg.DeviceReplayBuffer(300, eng.obs_size, eng.num_actions, eng.torch_device)
  a = actor.play_batch(buffer=buf)
  b = actor.play_batch(buffer=buf)
  assert buf.total_seen == a["total"] + b["total"] and len(buf) == 300
  tail = buf.newest(min(300, b["total"]))
  assert torch.equal(tail.observation, b["observation"][-tail.observation.shape[0]:])
  assert torch.equal(tail.value, b["value"][-tail.value.shape[0]:])
  actor.close()
  eng.close()
