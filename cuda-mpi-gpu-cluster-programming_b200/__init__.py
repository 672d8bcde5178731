"""b200cnn — Blackwell-native spatially-decomposed CNN inference framework.

Import as ``import b200cnn`` (alias package at the repo root).  Capabilities mirror
mykolas-perevicius/CUDA-MPI-GPU-Cluster-Programming (AlexNet blocks 1-2 forward, five
"versions" of parallelism, harness, log analysis, HW1 distributed matmul) re-designed
for B200: hand-written sm_100a kernels behind a C ABI (``include/b200cnn.h``),
``torch.distributed`` for rank bootstrap, peer memory over NVLink for scatter / halo /
gather.
"""
import os as _os

REPO_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))

from .config import LayerParams, NetConfig, default_config  # noqa: E402,F401
from .shapes import conv_out_dim, pool_out_dim, block12_shapes  # noqa: E402,F401

__version__ = "0.1.0"
