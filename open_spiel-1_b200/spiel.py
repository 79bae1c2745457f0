"""Host-side mirror of the slice of the pyspiel surface that the rollout path touches.

`mcts.py` in the reference needs only: game.get_type().{reward_model,dynamics},
game.max_utility(), state.{clone,is_terminal,is_chance_node,legal_actions,apply_action,
returns,current_player,history,observation_tensor,legal_actions_mask,action_to_string}
(pyspiel.cc:292-396).  The classes here give those names over the engine's
packed bitboard states through the C ABI host helpers, so the drop-in evaluator and the
MCTS mirror can be exercised where pyspiel itself is not installed (e.g. the GPU box).
A real `pyspiel.State` works with the evaluator just as well (it is read through
`state.history()`), these classes are not required.
"""
import ctypes
import enum

import numpy as np

from . import _cabi as C
from . import engine as _engine


class PlayerId(enum.IntEnum):
  """pyspiel.PlayerId (spiel.h:40-50)."""
  INVALID = -3
  TERMINAL = -4
  CHANCE = -1
  SIMULTANEOUS = -2


class GameType:
  """The fields of pyspiel.GameType that MCTSBot checks (mcts.py:240-245)."""

  class RewardModel(enum.Enum):
    REWARDS = 0
    TERMINAL = 1

  class Dynamics(enum.Enum):
    SIMULTANEOUS = 0
    SEQUENTIAL = 1

  def __init__(self, short_name):
    self.short_name = short_name
    self.reward_model = GameType.RewardModel.TERMINAL
    self.dynamics = GameType.Dynamics.SEQUENTIAL
    self.max_num_players = 2
    self.min_num_players = 2


_CANONICAL = {C.TIC_TAC_TOE: "tic_tac_toe()", C.CONNECT_FOUR: "connect_four()",
              C.BREAKTHROUGH: "breakthrough(columns=8,rows=8)", C.HEX: "hex(board_size=11)",
              C.HEX5_HOST_ONLY: "hex(board_size=5)", C.PENTAGO: "pentago(ansi_color_output=False)", C.OTHELLO: "othello()",
              C.Y: "y(ansi_color_output=False,board_size=19)", C.Y9_HOST_ONLY: "y(ansi_color_output=False,board_size=9)",
              C.HAVANNAH: "havannah(ansi_color_output=False,board_size=8)",
              C.HAVANNAH4_HOST_ONLY: "havannah(ansi_color_output=False,board_size=4)",
              C.GO: "go(board_size=9,handicap=0,komi=7.5)", C.GO7_HOST_ONLY: "go(board_size=7,handicap=0,komi=4.5)"}
_SHORT = {C.TIC_TAC_TOE: "tic_tac_toe", C.CONNECT_FOUR: "connect_four",
          C.BREAKTHROUGH: "breakthrough", C.HEX: "hex", C.HEX5_HOST_ONLY: "hex",
          C.PENTAGO: "pentago", C.OTHELLO: "othello", C.Y: "y", C.Y9_HOST_ONLY: "y",
          C.HAVANNAH: "havannah", C.HAVANNAH4_HOST_ONLY: "havannah", C.GO: "go", C.GO7_HOST_ONLY: "go"}


class Game:
  """Mirror of pyspiel.Game for the four supported games (load with `load_game`)."""

  def __init__(self, game_string):
    game_string = str(game_string).strip()
    self.gid = _engine.game_id(game_string)
    self._type = GameType(_SHORT[self.gid])
    # Game::ToString (spiel.cc:565-569) echoes the parameters the game was loaded with,
    # sorted by key
    name, _, rest = game_string.partition("(")
    params = sorted(p.strip() for p in rest.rstrip(")").split(",") if p.strip())
    self._string = "%s(%s)" % (name, ",".join(params))

  def __str__(self):
    return self._string

  def __eq__(self, other):
    # pyspiel.Game.__eq__ compares the game strings (pyspiel.cc:383-386)
    return isinstance(other, Game) and str(self) == str(other)

  def __hash__(self):
    return hash(self._string)

  def __reduce__(self):
    # pickle support as pyspiel.cc:387-396: the game string, reloaded with load_game
    return (load_game, (self._string,))

  def canonical_string(self):
    """The game string with every parameter spelled out."""
    return _CANONICAL[self.gid]

  def get_type(self):
    return self._type

  def new_initial_state(self):
    return State(self)

  def num_distinct_actions(self):
    return _engine.num_distinct_actions(self.gid)

  def max_game_length(self):
    return _engine.max_game_length(self.gid)

  def observation_tensor_shape(self):
    return _engine.observation_shape(self.gid)

  def observation_tensor_size(self):
    return _engine.observation_size(self.gid)

  def num_players(self):
    return 2

  def min_utility(self):
    return -1.0

  def max_utility(self):
    return 1.0

  def utility_sum(self):
    return 0.0

  def deserialize_state(self, text):
    """Game::DeserializeState: board strings for connect_four (connect_four.cc:229-279) and
    breakthrough (breakthrough.cc:399-435, side to move stays black as in the reference), one
    action per line for tic_tac_toe and hex (spiel.cc:396-426)."""
    if self.gid in (C.CONNECT_FOUR, C.BREAKTHROUGH):
      packed = np.zeros(_engine.state_bytes(self.gid), np.uint8)
      C.check(C.lib.bsg_host_deserialize(self.gid, text.encode(), packed.ctypes.data, 1, 0))
      return State(self, packed, _history=None, _history_known=False)
    state = self.new_initial_state()
    for line in text.split("\n"):
      if line.strip():
        state.apply_action(int(line))
    return state


def load_game(game_string):
  """pyspiel.load_game for the supported game strings; RuntimeError otherwise
  (the reference's SpielFatalError surfaces as RuntimeError, pyspiel.cc:839-840)."""
  return Game(game_string)


class State:
  """Mirror of pyspiel.State over one packed bitboard state (host memory)."""

  def __init__(self, game, _packed=None, _history=None, _history_known=True):
    self._game = game
    self._gid = game.gid
    if _packed is None:
      _packed = _engine.pack_histories(self._gid, [[]])
    self._packed = _packed
    self._history = list(_history) if _history is not None else []
    # a state deserialised from a board string has no history, exactly as in the reference
    self._history_known = _history_known

  # -- plumbing
  def get_game(self):
    return self._game

  def packed(self):
    """The packed state (numpy uint8, capacity 1) the device kernels consume."""
    return self._packed

  def clone(self):
    return State(self._game, self._packed.copy(), self._history, self._history_known)

  def __reduce__(self):
    # pickle support as pyspiel.cc:352-360: the state travels as SerializeGameAndState text
    return (_state_from_text, (serialize_game_and_state(self._game, self),))

  def history(self):
    return list(self._history)

  def _info(self):
    cur, term, ret0 = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    C.check(C.lib.bsg_host_state_info(self._gid, self._packed.ctypes.data, 1, 0,
                                      ctypes.byref(cur), ctypes.byref(term), ctypes.byref(ret0)))
    return cur.value, term.value, ret0.value

  # -- the pyspiel.State surface used by mcts.py
  def current_player(self):
    return self._info()[0]

  def is_terminal(self):
    return bool(self._info()[1])

  def is_chance_node(self):
    return False

  def returns(self):
    r = float(self._info()[2])
    return [r, -r]

  def legal_actions(self, player=None):
    cur, term, _ = self._info()
    if term or (player is not None and player != cur):
      return []  # empty for the non-mover, as the reference (basic_tests.cc:84-106)
    out = np.zeros(320, np.int64)
    n = C.check(C.lib.bsg_host_legal_actions(self._gid, self._packed.ctypes.data, 1, 0,
                                             out.ctypes.data, 320))
    return out[:n].tolist()

  def legal_actions_mask(self, player=None):
    cur, term, _ = self._info()
    n = self._game.num_distinct_actions()
    if term or (player is not None and player != cur):
      return [0] * n
    out = np.zeros(n, np.uint8)
    C.check(C.lib.bsg_host_legal_mask(self._gid, self._packed.ctypes.data, 1, 0, out.ctypes.data))
    return out.tolist()

  def apply_action(self, action):
    C.check(C.lib.bsg_host_apply_action(self._gid, self._packed.ctypes.data, 1, 0, int(action)))
    self._history.append(int(action))

  def child(self, action):
    c = self.clone()
    c.apply_action(action)
    return c

  def observation_tensor(self, player=None):
    if player is None:
      player = self.current_player()
    if player < 0 or player > 1:
      raise RuntimeError("bsg: observation_tensor needs player 0 or 1, got %d" % player)
    out = np.zeros(self._game.observation_tensor_size(), np.float32)
    C.check(C.lib.bsg_host_observation(self._gid, self._packed.ctypes.data, 1, 0, int(player),
                                       out.ctypes.data))
    return out.tolist()

  def action_to_string(self, player, action):
    return str(action)

  def _text(self, fn):
    buf = ctypes.create_string_buffer(4096)
    n = C.check(fn(self._gid, self._packed.ctypes.data, 1, 0, buf, 4096))
    return buf.raw[:n].decode()

  def to_string(self):
    """State::ToString in the reference's format (e.g. connect_four.cc:166-175)."""
    return self._text(C.lib.bsg_host_to_string)

  def __str__(self):
    return self.to_string()

  def serialize(self):
    """State::Serialize: the board for connect_four / breakthrough, the action history (one
    action per line, spiel.cc:375-384) for tic_tac_toe / hex."""
    if self._gid in (C.CONNECT_FOUR, C.BREAKTHROUGH):
      return self._text(C.lib.bsg_host_serialize)
    return "".join("%d\n" % a for a in self._history) if self._history else "\n"


def serialize_game_and_state(game, state):
  """SerializeGameAndState (spiel.cc:428-447)."""
  return ("# Automatically generated by OpenSpiel SerializeGameAndState\n[Meta]\nVersion: 1\n\n"
          "[Game]\n%s\n[State]\n%s\n" % (game, state.serialize()))


def _state_from_text(text):
  return deserialize_game_and_state(text)[1]


def deserialize_game_and_state(text):
  """DeserializeGameAndState (spiel.cc:449-495): (game, state) from the three sections."""
  sections = {"[Meta]": [], "[Game]": [], "[State]": []}
  cur = None
  for line in text.split("\n"):
    if not line or line.startswith("#"):
      continue
    if line in sections:
      cur = line
    elif cur is None:
      raise RuntimeError("bsg: text before the first section")
    else:
      sections[cur].append(line)
  game = load_game("\n".join(sections["[Game]"]))
  return game, game.deserialize_state("\n".join(sections["[State]"]))
