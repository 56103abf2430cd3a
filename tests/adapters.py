"""Public API adapters — the same eight entry points as the reference's ``tests/adapters.py:10-140``."""
from __future__ import annotations

from typing import Type

import torch

from cs336_a2_systems_b200.ops.attention import FlashAttentionSM100, FlashAttentionTorch
from cs336_a2_systems_b200.parallel import DDPBucketed, DDPIndividualParameters, ShardedOptimizer


def get_flashattention_autograd_function_pytorch() -> Type:
    """autograd.Function implementing FlashAttention-2 with PyTorch ops only."""
    return FlashAttentionTorch


def get_flashattention_autograd_function_triton() -> Type:
    """autograd.Function implementing FlashAttention-2 on hand-written GPU kernels.

    The reference returns its Triton implementation here; ours is tcgen05/TMEM/TMA CUDA for sm_100a.
    """
    return FlashAttentionSM100


def get_ddp_individual_parameters(module: torch.nn.Module) -> torch.nn.Module:
    return DDPIndividualParameters(module)


def ddp_individual_parameters_on_after_backward(ddp_model: torch.nn.Module, optimizer: torch.optim.Optimizer):
    ddp_model.finish_gradient_synchronization()


def get_ddp_bucketed(module: torch.nn.Module, bucket_size_mb: float) -> torch.nn.Module:
    return DDPBucketed(module, bucket_size_mb)


def ddp_bucketed_on_after_backward(ddp_model: torch.nn.Module, optimizer: torch.optim.Optimizer):
    ddp_model.finish_gradient_synchronization()


def ddp_bucketed_on_train_batch_start(ddp_model: torch.nn.Module, optimizer: torch.optim.Optimizer):
    ddp_model.on_train_batch_start()


def get_sharded_optimizer(params, optimizer_cls: Type[torch.optim.Optimizer], **kwargs) -> torch.optim.Optimizer:
    return ShardedOptimizer(params, optimizer_cls, **kwargs)
