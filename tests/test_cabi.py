"""The C-ABI library loads and exports every symbol include/bsg.h declares (no GPU calls)."""
import ctypes
import os
import re

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
  with open(os.path.join(REPO, "include", "bsg.h")) as f:
    text = f.read()
  text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
  return sorted(set(re.findall(r"\b(bsg_[a-z0-9_]+)\s*\(", text)))


def test_exports_every_declared_symbol():
  from open_spiel_1_b200 import _cabi
  lib = ctypes.CDLL(_cabi.LIB_PATH)
  names = declared_symbols()
  assert len(names) >= 26
  for name in names:
    assert hasattr(lib, name), name
  assert set(_cabi.EXPORTS) == set(names)
  assert lib.bsg_version() == 1


def test_static_facts_match_reference_constants():
  """NumDistinctActions / MaxGameLength / ObservationTensorSize."""
  import open_spiel_1_b200 as bsg
  facts = {"tic_tac_toe": (9, 9, 27, [3, 3, 3]), "connect_four": (7, 42, 126, [3, 6, 7]),
           "breakthrough": (768, 209, 192, [3, 8, 8]), "hex": (121, 121, 1089, [9, 11, 11]),
           "pentago": (288, 36, 108, [3, 6, 6]), "othello": (65, 64, 192, [3, 8, 8]),
           "y": (361, 190, 1083, [3, 19, 19]), "y(board_size=9)": (81, 45, 243, [3, 9, 9]),
           "havannah": (225, 169, 675, [3, 15, 15]), "havannah(board_size=4)": (49, 37, 147, [3, 7, 7]),
           "go(board_size=9)": (82, 162, 324, [4, 9, 9]), "go(board_size=7,komi=4.5)": (50, 98, 196, [4, 7, 7])}
  for name, (na, ml, osz, shape) in facts.items():
    g = bsg.load_game(name)
    assert g.num_distinct_actions() == na
    assert g.max_game_length() == ml
    assert g.observation_tensor_size() == osz
    assert g.observation_tensor_shape() == shape


def test_no_cpu_fallback_without_gpu():
  import torch
  import open_spiel_1_b200 as bsg
  if torch.cuda.is_available():
    pytest.skip("GPU present")
  with pytest.raises(RuntimeError):
    bsg.Engine("connect_four")
  with pytest.raises(RuntimeError):
    bsg.GpuRolloutEvaluator("connect_four", n_rollouts=4)
