// empty stand-in (nothing from absl/strings/ascii.h is used by the hot-path TUs)
