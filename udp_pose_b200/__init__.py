"""udp_pose_b200 -- B200 (sm_100a) implementation of the UDP heatmap codec hot path of
realphongha/UDP-Pose: unbiased affine crop, UDP / UDPv1 / RSN target encoding, flip-back,
DARK-style / offset-vote decoding, inverse affine.

Layers:
  csrc/ + libudp_codec.so   hand-written CUDA kernels behind the C ABI of include/udp_codec.h
  device_api                batched CUDA-tensor API (one launch per call) -- the fast path
  reference_api             the reference's own function names and signatures (numpy in/out)
  sharding                  person sharding across GPUs + the single keypoint all-gather
There is no CPU fallback: every function needs the CUDA library and a GPU.
"""
from .configs import CONFIGS, CodecConfig  # noqa: F401

__version__ = "0.1.0"
