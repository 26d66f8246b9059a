"""b200step — host side of the B200-native data-parallel training step behind task.py's interface.

The compute lives in libb200step.so (hand-written sm_100a CUDA behind a C ABI, see include/b200step.h);
this package owns device memory through PyTorch, builds the per-step launch schedule and mirrors the
reference script's launcher / sampler / checkpoint behaviour (reference: task.py).
"""
__version__ = "0.1.0"
