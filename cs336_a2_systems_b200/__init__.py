"""cs336_a2_systems_b200 — a Blackwell (B200, sm_100a) native training-systems framework.

Capabilities mirror the CS336 "Systems" assignment solution (reference layer map:
SURVEY.md §1): a Transformer LM library, FlashAttention-2 autograd Functions,
overlapped data-parallel wrappers (per-parameter and bucketed), an optimizer-state
sharding wrapper, training drivers and a benchmark/profiling harness.

Layout
------
``ops``       op library: pure-PyTorch oracles + autograd Functions backed by the
              hand-written sm_100a kernels in ``csrc/`` (loaded through a C ABI).
``models``    ``TransformerLM`` (state-dict compatible with ``BasicsTransformerLM``).
``optim``     AdamW (course semantics) with a fused flat-arena kernel, cosine LR.
``parallel``  DDP wrappers, sharded optimizer, NVLink symmetric-memory collectives.
``train``     step loop, CUDA-graph trainer, checkpointing, CLI drivers.
``bench``     device-timed harnesses (LM step, attention sweep, all-reduce, memory).
``utils``     timing, NVTX, clocks sampling, logging.

The hyphenated name asked for by the project brief (``cs336-a2-systems_b200``) is
not importable in Python, so the package directory uses underscores.
"""

__version__ = "0.1.0"

from . import _native  # noqa: F401  (loads libb200sys.so when present)

__all__ = ["__version__"]
