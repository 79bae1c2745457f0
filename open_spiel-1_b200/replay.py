"""Device-resident replay buffer for the AlphaZero data path.

The reference's learner keeps `TrainInput(observation, legals_mask, policy, value)` tuples in a
Python list (`alpha_zero.py:83-104` Buffer: fixed size, keeps the newest, `sample(count)`
without replacement; filled at `alpha_zero.py:366-369` with value = the trajectory's
returns[0]) and stacks a sample into numpy arrays for every training step
(`model.py:80-91`).  Here the same records never leave the GPU: `Engine.trajectories` writes
observation planes and legal masks of every decision point with the CUDA expand kernel, and
this ring of preallocated device tensors takes them as they are.  torch is used for memory and
indexing only.
"""
import collections


class TrainInput(collections.namedtuple("TrainInput", "observation legals_mask policy value")):
  """One batch for the learner, as model.py:80-91 but device tensors: observation [n, C*H*W]
  float32, legals_mask [n, A] bool, policy [n, A] float32, value [n, 1] float32."""


class DeviceReplayBuffer(object):
  """A fixed size buffer that keeps the newest records (alpha_zero.py:83-104), on the GPU."""

  def __init__(self, max_size, obs_size, num_actions, device):
    import torch
    self.max_size = int(max_size)
    self.device = device
    self.observation = torch.zeros((self.max_size, int(obs_size)), dtype=torch.float32, device=device)
    self.legals_mask = torch.zeros((self.max_size, int(num_actions)), dtype=torch.bool, device=device)
    self.policy = torch.zeros((self.max_size, int(num_actions)), dtype=torch.float32, device=device)
    self.value = torch.zeros((self.max_size, 1), dtype=torch.float32, device=device)
    self.total_seen = 0   # records that have passed through
    self._size = 0
    self._head = 0        # next slot to write

  def __len__(self):
    return self._size

  def extend(self, observation, legals_mask, policy, value):
    """Appends n records (device tensors, first dimension n); the oldest fall out."""
    import torch
    n = int(observation.shape[0])
    self.total_seen += n
    if n == 0:
      return
    observation = observation.reshape(n, -1)
    value = value.reshape(n, 1).to(torch.float32)
    if n >= self.max_size:      # only the newest max_size survive
      observation, legals_mask = observation[-self.max_size:], legals_mask[-self.max_size:]
      policy, value = policy[-self.max_size:], value[-self.max_size:]
      n = self.max_size
    idx = (torch.arange(n, device=self.device) + self._head) % self.max_size
    self.observation[idx] = observation.to(torch.float32)
    self.legals_mask[idx] = legals_mask.to(torch.bool)
    self.policy[idx] = policy.to(torch.float32)
    self.value[idx] = value
    self._head = (self._head + n) % self.max_size
    self._size = min(self.max_size, self._size + n)

  def extend_from_trajectories(self, rec):
    """Takes the dict `Engine.trajectories` returns: one record per decision point, policy =
    the behaviour policy of the playouts (uniform over the legal actions), value = the
    playout's final return for player 0 (alpha_zero.py:366-369)."""
    import torch
    if not rec["total"]:
      return
    mask = rec["mask"].to(torch.bool)
    policy = mask.to(torch.float32)
    policy = policy / policy.sum(dim=1, keepdim=True)
    value = rec["ret0"].to(torch.float32)[rec["playout"]]
    self.extend(rec["obs"], mask, policy, value)

  def newest(self, count):
    """The `count` newest records, oldest first (what `data[-count:]` is in the reference)."""
    import torch
    count = min(int(count), self._size)
    idx = (torch.arange(count, device=self.device) + (self._head - count)) % self.max_size
    return TrainInput(self.observation[idx], self.legals_mask[idx], self.policy[idx], self.value[idx])

  def sample(self, count, generator=None):
    """`count` distinct records, uniformly at random (random.sample in the reference)."""
    import torch
    if count > self._size:
      raise ValueError("Sample larger than population")
    pick = torch.randperm(self._size, device=self.device, generator=generator)[:int(count)]
    idx = (pick + (self._head - self._size)) % self.max_size
    return TrainInput(self.observation[idx], self.legals_mask[idx], self.policy[idx], self.value[idx])
