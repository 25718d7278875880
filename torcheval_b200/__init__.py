"""torcheval_b200 — B200-native implementation of torcheval's ranking & counting metric hot path.

Layout: ``csrc/`` (hand-written sm_100a kernels behind the C ABI of include/te_b200.h, built into
``lib/libte_b200.so``), ``_native`` / ``ops`` (ctypes binding, tensor-level launchers),
``metrics`` (the drop-in ``torcheval.metrics`` surface) and ``distributed`` (NCCL paths).
"""
__version__ = "0.1.0"
