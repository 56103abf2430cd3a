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


def test_cross_rank_flag_barrier_protocol():
    """The 1-deep flag channels of the NVLink kernels' cross-rank barrier: back-to-back kernels re-use the same words
    without losing or duplicating an arrival, also when only part of a grid is resident (in-order block dispatch)."""
    for seed in range(4):
        ps.sim_cross_rank_barrier(ws=4, blocks=5, kernels=4, seed=seed)
        ps.sim_cross_rank_barrier(ws=3, blocks=6, kernels=3, resident=2, seed=seed)


def test_model_reproduces_the_partial_residency_hazard():
    """Known limit of per-block cross-rank barriers: a grid that is not fully co-resident deadlocks if the ranks'
    block schedulers dispatch blocks in different orders.  (The kernels rely on in-order dispatch and on grids of at
    most one CTA per SM.)"""
    with pytest.raises(AssertionError):
        ps.sim_cross_rank_barrier(ws=3, blocks=8, kernels=1, resident=3, seed=1, shuffled_launch=True)
