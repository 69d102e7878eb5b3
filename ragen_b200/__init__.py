"""ragen_b200 — B200-native rollout hot path for RAGEN (env reset/step/render, action parse and
reward, trajectory masks/scores, GAE / group reward normalisation) behind RAGEN's own Python API.

All compute runs in hand-written sm_100a CUDA kernels in `lib/libragen_b200.so`, reached through the
C ABI in `include/ragen_b200.h`.  There is no CPU fallback.
"""
__version__ = "0.1.0"
