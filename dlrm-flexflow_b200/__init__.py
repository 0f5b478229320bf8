"""dlrm-flexflow_b200 — B200-native embedding hot path for DLRM-FlexFlow.

The product is libffemb.so (hand-written sm_100a CUDA behind the C ABI of include/ffemb.h).
This package is the thin Python host side used by the tests and bench.py: it loads the
library with ctypes and mirrors the reference operator's interface for this path
(/root/reference/include/model.h:1136-1202, src/ops/embedding.cu).  torch is used only for
device memory, streams and torch.distributed plumbing.  There is no CPU fallback: importing
works anywhere, but every compute call requires the built library and a CUDA device.

The directory name contains a hyphen; import it with
    importlib.import_module("dlrm-flexflow_b200")
"""
from .binding import (AGGR_MODE_AVG, AGGR_MODE_MEAN, AGGR_MODE_NONE, AGGR_MODE_SUM, FFEmbError,
                      build_library, declared_symbols, exported_symbols, launch_count, lib,
                      library_path, profile_collect, profile_enable)
from . import sharding, synth
from . import exchange
from .embedding import Embedding, EmbeddingCollection, EmbeddingMeta, StepGraph, load_sparse_inputs
from .hoststep import HostStep

__all__ = ["AGGR_MODE_AVG", "AGGR_MODE_NONE", "AGGR_MODE_SUM", "AGGR_MODE_MEAN", "FFEmbError",
           "build_library", "declared_symbols", "exported_symbols", "lib", "library_path",
           "launch_count", "profile_collect", "profile_enable", "Embedding", "EmbeddingCollection", "EmbeddingMeta", "HostStep", "StepGraph", "load_sparse_inputs", "synth", "sharding", "exchange"]
