// empty stand-in (nothing from absl/strings/match.h is used by the hot-path TUs)
