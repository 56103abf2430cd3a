"""Parallelism: overlapped DDP (per-parameter, bucketed), optimizer-state sharding, NVLink collectives."""
from .ddp import (DDPBucketed, DDPIndividualParameters, allreduce_gradients_flat, allreduce_gradients_naive,
                  broadcast_parameters, plan_buckets)
from .process_group import cleanup, find_free_port, setup, world
from .sharded_optimizer import ShardedOptimizer

__all__ = ["DDPIndividualParameters", "DDPBucketed", "ShardedOptimizer", "broadcast_parameters", "plan_buckets",
           "allreduce_gradients_naive", "allreduce_gradients_flat", "setup", "cleanup", "world", "find_free_port"]
