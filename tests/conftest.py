import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))
GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs the upstream reference checkout (build container only)")


def pytest_collection_modifyitems(config, items):
    import torch
    has_gpu = torch.cuda.is_available()
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords and not has_gpu:
            item.add_marker(skip_gpu)


def densify(npz, prefix=""):
    shape = tuple(int(x) for x in npz[f"{prefix}shape"])
    out = np.zeros(int(np.prod(shape)), dtype=np.float32)
    out[npz[f"{prefix}idx"].astype(np.int64)] = npz[f"{prefix}val"]
    return out.reshape(shape)


def bits(a: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="session")
def tables():
    from pokeai_b200.tables import load_tables
    return load_tables()


@pytest.fixture(scope="session")
def synth_golden():
    """name -> (packed records, party_size, reference features, reference masks)"""
    from pokeai_b200.packed import PK_OBS_DTYPE
    z = np.load(GOLDEN / "synth_features.npz")
    out = {}
    for name in ("full6", "ragged6", "ragged3", "ragged1", "full4"):
        rec = np.frombuffer(z[f"{name}_packed"].tobytes(), dtype=PK_OBS_DTYPE).copy()
        out[name] = (rec, int(z[f"{name}_party_size"]), densify(z, f"{name}_"), z[f"{name}_mask"])
    return out


@pytest.fixture(scope="session")
def fixture_golden():
    """The 220 real logged decisions: (list of dicts, reference features (220,617,6), masks)"""
    import gzip
    import json
    with gzip.open(GOLDEN / "fixture_decisions.json.gz", "rt", encoding="utf-8") as f:
        dec = json.load(f)
    z = np.load(GOLDEN / "fixture_features.npz")
    return dec, densify(z), z["mask"]
