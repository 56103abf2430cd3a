"""ops/gemm_tuned.json (measured tile plans) must only name plans the kernels accept (CPU check of the table)."""

import json
import os

from cs336_a2_systems_b200.ops import kernels

PATH = os.path.join(os.path.dirname(kernels.__file__), "gemm_tuned.json")


def test_tuned_table_is_legal():
    plans = json.load(open(PATH))["plans"]
    assert plans, "empty plan table"
    for key, (use_pair, bn) in plans.items():
        M, N, K, a_mn, b_mn = (int(x) for x in key.split(","))
        assert bn in (128, 192, 256), key
        assert M > 0 and N > 0 and K > 0 and a_mn in (0, 1) and b_mn in (0, 1), key
        if use_pair:
            # CTA pairs: 256-row tiles, no (A MN-major, B K-major) variant (ops/kernels.py: gemm)
            assert M >= 256 and not (a_mn and not b_mn), key


def test_plan_lookup_falls_back_to_the_model():
    key = next(iter(kernels._TUNED)) if kernels._TUNED else None
    if key is not None:
        M, N, K, a_mn, b_mn = (int(x) for x in key.split(","))
        assert kernels.plan_gemm(M, N, K, bool(a_mn), bool(b_mn), True) == kernels._TUNED[key]
        # a tuned pair plan must not be used where pairs are not allowed
        pair, bn = kernels.plan_gemm(M, N, K, bool(a_mn), bool(b_mn), False)
        assert not pair and bn in (128, 192, 256)
    pair, bn = kernels.plan_gemm(777, 333, 4096, False, False, True)  # not in the table
    assert bn in (128, 192, 256)
