"""newmap_b200: B200-native `newmap search` hot path (min-unique-k over a reference-built
.awfmi FM-index).  Host side mirrors hoffmangroup/newmap's Python interface; all counting
runs in hand-written CUDA (sm_100a) behind the C-ABI in include/newmap_b200.h."""

__version__ = "0.1.0"
