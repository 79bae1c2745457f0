// TEST INFRASTRUCTURE ONLY.  A reduced Python module named `pyspiel` over the UNMODIFIED
// reference C++ (spiel.cc + the four games), compiled where the sources lie under
// /root/reference by oracle/Makefile into oracle/_ref/.  It exposes exactly the surface the
// reference's python/algorithms/mcts.py touches (last row) so that file —
// byte-compiled from its own location into oracle/_ref/ref_mcts.bytecode — runs unmodified:
//   * on top of the reference's own states with THIS repo's GpuRolloutEvaluator plugged into
//     its `evaluator` argument: the drop-in test proper (tests/test_gpu_dropin.py);
//   * with its own RandomRolloutEvaluator as the Python CPU baseline of BASELINE config 1.
// The full binding (python/pybind11/pyspiel.cc) links every game and algorithm of the
// reference, several of which need un-vendored dependencies; this file is not a copy of it.
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "open_spiel/spiel.h"

namespace py = pybind11;
using open_spiel::Action;
using open_spiel::Game;
using open_spiel::GameType;
using open_spiel::Player;
using open_spiel::State;

namespace {
// pyspiel.Bot is only a base class for mcts.MCTSBot (mcts.py:199,239).
struct BotBase {
  virtual ~BotBase() {}
};
// SpielFatalError exits the process by default (spiel_utils.cc:94-110); from Python it has to
// surface as RuntimeError (python/tests/pyspiel_test.py:203-206).
void ThrowingHandler(const std::string& message) { throw std::runtime_error(message); }
}  // namespace

PYBIND11_MODULE(pyspiel, m) {
  m.doc() = "reduced pyspiel over the unmodified reference (test infrastructure)";
  open_spiel::SetErrorHandler(ThrowingHandler);

  py::class_<BotBase>(m, "Bot").def(py::init<>());

  py::enum_<open_spiel::PlayerId>(m, "PlayerId")
      .value("INVALID", open_spiel::kInvalidPlayer)
      .value("TERMINAL", open_spiel::kTerminalPlayerId)
      .value("CHANCE", open_spiel::kChancePlayerId)
      .value("SIMULTANEOUS", open_spiel::kSimultaneousPlayerId);

  py::class_<GameType> game_type(m, "GameType");
  game_type.def_readonly("short_name", &GameType::short_name)
      .def_readonly("long_name", &GameType::long_name)
      .def_readonly("dynamics", &GameType::dynamics)
      .def_readonly("chance_mode", &GameType::chance_mode)
      .def_readonly("information", &GameType::information)
      .def_readonly("utility", &GameType::utility)
      .def_readonly("reward_model", &GameType::reward_model)
      .def_readonly("max_num_players", &GameType::max_num_players)
      .def_readonly("min_num_players", &GameType::min_num_players);
  py::enum_<GameType::Dynamics>(game_type, "Dynamics")
      .value("SEQUENTIAL", GameType::Dynamics::kSequential)
      .value("SIMULTANEOUS", GameType::Dynamics::kSimultaneous);
  py::enum_<GameType::ChanceMode>(game_type, "ChanceMode")
      .value("DETERMINISTIC", GameType::ChanceMode::kDeterministic)
      .value("EXPLICIT_STOCHASTIC", GameType::ChanceMode::kExplicitStochastic)
      .value("SAMPLED_STOCHASTIC", GameType::ChanceMode::kSampledStochastic);
  py::enum_<GameType::Information>(game_type, "Information")
      .value("ONE_SHOT", GameType::Information::kOneShot)
      .value("PERFECT_INFORMATION", GameType::Information::kPerfectInformation)
      .value("IMPERFECT_INFORMATION", GameType::Information::kImperfectInformation);
  py::enum_<GameType::Utility>(game_type, "Utility")
      .value("ZERO_SUM", GameType::Utility::kZeroSum)
      .value("CONSTANT_SUM", GameType::Utility::kConstantSum)
      .value("GENERAL_SUM", GameType::Utility::kGeneralSum)
      .value("IDENTICAL", GameType::Utility::kIdentical);
  py::enum_<GameType::RewardModel>(game_type, "RewardModel")
      .value("REWARDS", GameType::RewardModel::kRewards)
      .value("TERMINAL", GameType::RewardModel::kTerminal);

  py::class_<State>(m, "State")
      .def("current_player", &State::CurrentPlayer)
      .def("apply_action", &State::ApplyAction)
      .def("legal_actions", (std::vector<Action>(State::*)(Player) const) & State::LegalActions)
      .def("legal_actions", (std::vector<Action>(State::*)() const) & State::LegalActions)
      .def("legal_actions_mask", (std::vector<int>(State::*)() const) & State::LegalActionsMask)
      .def("action_to_string",
           (std::string(State::*)(Player, Action) const) & State::ActionToString)
      .def("is_terminal", &State::IsTerminal)
      .def("is_chance_node", &State::IsChanceNode)
      .def("chance_outcomes", &State::ChanceOutcomes)
      .def("returns", &State::Returns)
      .def("player_return", &State::PlayerReturn)
      .def("history", &State::History)
      .def("clone", &State::Clone)
      .def("child", &State::Child)
      .def("num_players", &State::NumPlayers)
      .def("num_distinct_actions", &State::NumDistinctActions)
      .def("observation_tensor",
           (std::vector<double>(State::*)(Player) const) & State::ObservationTensor)
      .def("observation_tensor", (std::vector<double>(State::*)() const) & State::ObservationTensor)
      .def("serialize", &State::Serialize)
      .def("get_game", &State::GetGame)
      .def("__str__", &State::ToString);

  py::class_<Game, std::shared_ptr<Game>>(m, "Game")
      .def("num_distinct_actions", &Game::NumDistinctActions)
      .def("new_initial_state", &Game::NewInitialState)
      .def("num_players", &Game::NumPlayers)
      .def("min_utility", &Game::MinUtility)
      .def("max_utility", &Game::MaxUtility)
      .def("utility_sum", &Game::UtilitySum)
      .def("get_type", &Game::GetType)
      .def("max_game_length", &Game::MaxGameLength)
      .def("observation_tensor_shape", &Game::ObservationTensorShape)
      .def("observation_tensor_size", &Game::ObservationTensorSize)
      .def("deserialize_state", &Game::DeserializeState)
      .def("__str__", &Game::ToString);

  m.def("load_game", [](const std::string& name) {
    return std::const_pointer_cast<Game>(open_spiel::LoadGame(name));
  });
  m.def("registered_names", []() { return open_spiel::RegisteredGames(); });
}
