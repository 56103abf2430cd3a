"""Training: the step loop (eager or CUDA-graph), synthetic data, checkpointing, CLI drivers."""
from .trainer import GradArena, Trainer

__all__ = ["Trainer", "GradArena"]
