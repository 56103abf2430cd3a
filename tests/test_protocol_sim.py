"""The mbarrier protocols of the warp-specialised FlashAttention variants, replayed in a discrete-event model
(``tools/protocol_sim.py``): no deadlock and no buffer hazard for randomised latencies — and the model does flag
deliberately broken protocols.  Pure Python, no GPU."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import protocol_sim as ps  # noqa: E402


@pytest.mark.parametrize("n_iter", [0, 1, 2, 3, 6])
def test_backward_reordered_protocol(n_iter):
    for warps in (8, 16):
        for seed in range(8):
            ps.sim_bwd_reordered(n_iter, n_warps=warps, seed=seed)


@pytest.mark.parametrize("n_kv", [(1, 1), (2, 1), (1, 2), (5, 4), (4, 5), (9, 9)])
def test_forward_cluster_multicast_protocol(n_kv):
    for slots in (3, 4):
        for seed in range(8):
            ps.sim_fwd_cluster(n_kv, slots=slots, seed=seed)


@pytest.mark.parametrize("n_kv", [0, 1, 2, 7])
def test_forward_o_in_tensor_memory_protocol(n_kv):
    for seed in range(8):
        ps.sim_fwd_o_in_tmem(n_kv, seed=seed)


def _caught(fn, seeds=30):
    n = 0
    for seed in range(seeds):
        try:
            fn(seed)
        except AssertionError:
            n += 1
    return n


def test_model_flags_broken_protocols():
    assert _caught(lambda s: ps.sim_bwd_reordered(5, seed=s, mutate="no_dq_full_wait")) > 0
    assert _caught(lambda s: ps.sim_bwd_reordered(5, seed=s, mutate="no_dq_empty_wait")) > 0
    assert _caught(lambda s: ps.sim_fwd_cluster((6, 6), seed=s, mutate="local_release")) > 0
