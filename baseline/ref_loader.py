"""Load the UNMODIFIED reference from ``baseline/_ref`` (pip --target install, git-ignored).

Three things keep the reference from importing as shipped: it imports the author's
private assignment-1 modules ``cs336_basics.{transformer,training,optimizers}``, the shipped model
needs ``einx``, and package metadata must be installed.  The metadata is satisfied by the pip
install; the other two are aliased here at run time — no file under ``baseline/_ref`` is edited:

* ``einx``                      → ``baseline/einx_shim.py`` (the five fixed patterns)
* ``cs336_basics.transformer``  → ``TransformerLM`` = the shipped ``model.BasicsTransformerLM``
* ``cs336_basics.training``     → ``cross_entropy_loss`` = the shipped ``nn_utils.cross_entropy``
* ``cs336_basics.optimizers``   → ``AdamW`` = the shipped ``optimizer.AdamW``
"""
import os
import sys
import types

REF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def available() -> tuple[bool, str]:
    if not os.path.isdir(os.path.join(REF_DIR, "cs336_systems")):
        return False, "baseline/_ref missing (pip --target install of /root/reference not present)"
    return True, ""


def load():
    """Returns a namespace with the reference's public classes."""
    ok, why = available()
    if not ok:
        raise ImportError(why)
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    here = os.path.dirname(os.path.abspath(__file__))
    if here not in sys.path:
        sys.path.insert(0, here)
    import einx_shim
    sys.modules.setdefault("einx", einx_shim)
    import cs336_basics
    from cs336_basics import model as ref_model, nn_utils as ref_nn, optimizer as ref_opt

    def alias(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        setattr(cs336_basics, name.rsplit(".", 1)[1], m)

    alias("cs336_basics.transformer", TransformerLM=ref_model.BasicsTransformerLM)
    alias("cs336_basics.training", cross_entropy_loss=ref_nn.cross_entropy)
    alias("cs336_basics.optimizers", AdamW=ref_opt.AdamW)
    from cs336_systems import ddp_bucketed_overlapped_sharded as ref_ddp
    from cs336_systems import flash_attention as ref_fa

    return types.SimpleNamespace(
        BasicsTransformerLM=ref_model.BasicsTransformerLM, AdamW=ref_opt.AdamW, cross_entropy=ref_nn.cross_entropy,
        DDP=ref_ddp.DDP, DDP_Bucketed=ref_ddp.DDP_Bucketed, ShardedStateOptimizer=ref_ddp.ShardedStateOptimizer,
        FlashAttentionTriton=ref_fa.FlashAttentionTriton, FlashAttentionTorch=ref_fa.FlashAttentionTorch,
        scaled_dot_product_attention=ref_model.scaled_dot_product_attention)
