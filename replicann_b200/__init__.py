"""replicann_b200 -- B200-native (sm_100a) training hot path behind replicann's nn.Module surface.

Sub-modules mirror the reference's import paths (`replicann.nn.attention`, `replicann.arch.transformer`)
and add the CNN layers / models BASELINE.json's north_star names.  All arithmetic runs in hand-written
CUDA kernels reached through the C ABI in include/rnn_b200.h; there is no CPU fallback.
"""
__version__ = "0.1.0"
