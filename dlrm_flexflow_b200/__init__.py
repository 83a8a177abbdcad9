"""B200-native DLRM embedding-bag path (drop-in for TravisDai/DLRM-FlexFlow's
`ff.embedding` operator): hand-written sm_100a kernels behind a C-ABI
(include/ffemb_c.h), plus this thin Python host mirror of the reference's
binding.  Importing the package does not load CUDA; the first call does, and
fails loudly if the extension cannot be built or loaded.
"""
from .core import (AggrMode, DataType, OpType, FFConfig, FFModel, Tensor, Parameter, Op, Embedding,
                   SGDOptimizer, Initializer, UniformInitializer, ZeroInitializer, Strategy,
                   round_robin_device, balance_placement, plan_placement)
from ._lib import FFEmbError, lib, library_path

__all__ = ["AggrMode", "DataType", "OpType", "FFConfig", "FFModel", "Tensor", "Parameter", "Op",
           "Embedding", "SGDOptimizer", "Initializer", "UniformInitializer", "ZeroInitializer",
           "Strategy", "round_robin_device", "balance_placement", "plan_placement", "FFEmbError", "lib",
           "library_path"]
