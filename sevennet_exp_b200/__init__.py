"""sevennet_exp_b200: B200-native (sm_100a) SevenNet energy/force/stress evaluation.

Host side is Python/PyTorch; all arithmetic on the hot path runs in the
hand-written CUDA library ``csrc/`` reached through the C ABI declared in
``include/sevenn_b200.h``.  There is no CPU fallback: importing the kernels
without the built library, or without a CUDA device, raises.
"""
__version__ = '0.1.0'
