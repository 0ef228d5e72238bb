"""torch_distributed_sandbox_b200 (short name: b200dp) - a Blackwell-native data-parallel
training framework with the capabilities of AditMeh/torch-distributed-sandbox.

Layer map:
  launch        spawn / find_free_port / sharing-strategy helpers          (L1)
  store         TCP key-value store (native)                               (L2)
  distributed   init_process_group, new_group, all_reduce, barrier, ...    (L2/L3)
  parallel      DistributedDataParallel, DistributedSampler, buckets       (L4/L5)
  data          MNIST idx reader/writer, synthetic data, loader, resize    (L5)
  models        ConvNet (eager fp32) and FusedConvNet (sm_100a engine)     (L6)
  ops           the hand-written CUDA kernels                              (L6)
  utils         timing, clocks, topology, logging                          (L7)
"""
__version__ = "0.1.0"

from . import distributed  # noqa: F401
from .distributed import (ReduceOp, all_gather, all_gather_into_tensor, all_reduce, barrier, broadcast,  # noqa: F401
                          destroy_process_group, get_rank, get_world_size, init_process_group, is_initialized,
                          new_group)
from .launch import find_free_port, set_sharing_strategy, spawn, use_file_system_sharing_strategy  # noqa: F401
