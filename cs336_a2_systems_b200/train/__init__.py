"""Training: the step loop (eager or CUDA-graph), drivers for every DDP flavour, synthetic data, checkpoints."""
from .checkpoint import load_checkpoint, save_checkpoint
from .data import SyntheticBatches, generate_random_data, get_batch, set_seed
from .trainer import GradArena, Trainer

__all__ = ["Trainer", "GradArena", "save_checkpoint", "load_checkpoint", "get_batch", "SyntheticBatches",
           "generate_random_data", "set_seed"]
