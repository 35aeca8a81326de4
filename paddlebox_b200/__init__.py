"""paddlebox_b200 — a B200-native sparse-CTR training framework with the capabilities of
PaddleBox (BoxPS): pass-based SlotRecord datasets, pull/push sparse embedding ops on an
HBM-sharded table, fused_seqpool_cvm, AUC metrics, data-parallel dense training and
base/delta checkpoints.  PyTorch is the framework layer; the hot paths are hand-written
sm_100a CUDA (csrc/cuda) and the host runtime is C++ (csrc/host).
"""
__version__ = "0.1.0"

from . import flags  # noqa: F401
