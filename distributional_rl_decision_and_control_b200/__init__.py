"""B200-native hot path of pszenher/Distributional_RL_Decision_and_Control's ``rfarl`` package.

The batched ASV env step, the device replay buffer and the fused IQN quantile-Huber update run
as hand-written sm_100a CUDA kernels in ``csrc/`` behind the C-ABI of ``include/asvrl.h``
(``libasvrl.so``, loaded with ctypes by ``_lib``).  There is no CPU fallback: every compute entry
point raises if the library or a CUDA device is missing.
"""
__version__ = "0.1.0"
