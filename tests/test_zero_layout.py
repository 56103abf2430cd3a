"""Host-side layout helpers of the bucket-fused ZeRO-1 step (parallel/zero.py): slice sizes, fp32-eager ranges and the
bucket <-> per-parameter state conversion used by checkpoints.  Pure CPU."""

import re

import torch

from cs336_a2_systems_b200.parallel import zero


def test_slices_cover_the_bucket_in_whole_vectors():
    for numel in (1, 7, 8, 9, 1000, 1600 * 6400, 26_214_400 + 3):
        for ws in (1, 2, 3, 4, 8):
            per = zero.slice_floats(numel, ws)
            assert per % 8 == 0  # 32-byte vectors
            assert per * ws >= numel  # the slices cover the bucket
            assert (per - 8) * ws < (numel + 7) // 8 * 8 or per == 8  # and are not a whole vector too long


def test_eager_ranges_merge_adjacent_slots():
    ps = [torch.empty(n) for n in (16, 5, 24, 8, 40)]
    offs, off = [], 0
    for p in ps:
        offs.append(off)
        off += (p.numel() + 7) // 8 * 8
    eager = {id(ps[0]), id(ps[1]), id(ps[3])}
    rs = zero.eager_ranges(ps, offs, lambda q: id(q) in eager)
    assert rs == [(0, 24), (48, 56)]  # 16 + round8(5) merged; the 8-element slot after the 24-element one
    assert zero.eager_ranges(ps, offs, lambda q: False) == []
    assert zero.eager_ranges(ps, offs, lambda q: True) == [(0, off)]


def test_bucket_state_round_trip():
    shapes = [(3, 4), (5,), (2, 2, 2)]
    offs = [0, 16, 24]
    total = 40
    full = torch.arange(total, dtype=torch.float32)
    parts = zero.split_bucket_state(full, shapes, offs)
    assert [tuple(p.shape) for p in parts] == shapes
    back = zero.assemble_bucket_state(parts, offs, total)
    for shape, o in zip(shapes, offs):
        n = 1
        for d in shape:
            n *= d
        assert torch.equal(back[o:o + n], full[o:o + n])
    assert float(back[12:16].abs().sum()) == 0.0  # padding stays zero
    assert torch.equal(zero.assemble_bucket_state([parts[0], None, parts[2]], offs, total)[16:21], torch.zeros(5))


def test_eager_range_cap_matches_the_kernel():
    import os

    src = open(os.path.join(os.path.dirname(__file__), "..", "csrc", "comm_adamw.cu")).read()
    assert int(re.search(r"constexpr int kZEager = (\d+);", src).group(1)) == zero.MAX_EAGER_RANGES
