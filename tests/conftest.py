"""pytest configuration: the ``gpu`` marker and collection rules.

``python -m pytest tests -m "not gpu"`` must pass on a CPU-only box; ``-m gpu`` runs on a B200.
GPU tests are skipped (not failed) when no CUDA device exists.
"""
import os
import sys
import pathlib

import pytest
import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

os.environ.setdefault("MASTER_ADDR", "127.0.0.1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    has_cuda = torch.cuda.is_available()
    n = torch.cuda.device_count() if has_cuda else 0
    for item in items:
        if "gpu" in item.keywords and not has_cuda:
            item.add_marker(pytest.mark.skip(reason="no CUDA device"))
        if "multigpu" in item.keywords and n < 2:
            item.add_marker(pytest.mark.skip(reason="needs >= 2 GPUs"))


# ---------------------------------------------------------------------------------------------
# Array snapshots (reference: tests/conftest.py:10-190 ships the same facility, unused by its tests; here
# tests/test_model.py pins the LM forward/backward numerics with it).  Stored as .npz under tests/_snapshots;
# regenerate with B200SYS_UPDATE_SNAPSHOTS=1.
# ---------------------------------------------------------------------------------------------
def pytest_addoption(parser):
    parser.addoption("--snapshot-exact", action="store_true", default=False,
                     help="compare array snapshots bit-for-bit instead of within tolerance")


class ArraySnapshot:
    def __init__(self, directory: pathlib.Path, name: str, exact: bool, rtol: float = 1e-4, atol: float = 1e-5):
        self.path = directory / f"{name}.npz"
        self.exact, self.rtol, self.atol = exact, rtol, atol

    @staticmethod
    def _to_numpy(x):
        import numpy as np

        if torch.is_tensor(x):
            return x.detach().to(torch.float32).cpu().numpy()
        return np.asarray(x)

    def assert_match(self, actual, rtol: float | None = None, atol: float | None = None) -> None:
        import numpy as np

        arrays = {k: self._to_numpy(v) for k, v in (actual.items() if isinstance(actual, dict) else [("value", actual)])}
        if os.environ.get("B200SYS_UPDATE_SNAPSHOTS") == "1":
            self.path.parent.mkdir(parents=True, exist_ok=True)
            np.savez_compressed(self.path, **arrays)
            return
        assert self.path.exists(), f"snapshot {self.path} missing (run once with B200SYS_UPDATE_SNAPSHOTS=1)"
        with np.load(self.path) as want:
            assert set(want.files) == set(arrays), f"snapshot keys {sorted(want.files)} != {sorted(arrays)}"
            for k in want.files:
                if self.exact:
                    np.testing.assert_array_equal(arrays[k], want[k], err_msg=k)
                else:
                    np.testing.assert_allclose(arrays[k], want[k], rtol=self.rtol if rtol is None else rtol,
                                               atol=self.atol if atol is None else atol, err_msg=k)


@pytest.fixture
def array_snapshot(request):
    name = request.node.name.replace("[", "_").replace("]", "").replace("/", "_")
    return ArraySnapshot(pathlib.Path(__file__).parent / "_snapshots", name,
                         exact=request.config.getoption("--snapshot-exact"))
