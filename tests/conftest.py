import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def rtb():
    return importlib.import_module("unity-raytracer_b200")


@pytest.fixture(scope="session")
def scenes(rtb):
    return rtb.scenes


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def emu():
    from emu import emu as E
    E.lib()
    return E


def compare_renders(ref, got, check_float=True):
    """Parity rule of BASELINE.json north_star: ids/t bit-exact, u8 within +-1 LSB on >= 99.9 % of pixels.
    Returns a dict of mismatch counts; MeshIndex is compared only for MeshTriangle hits (Scene.cs:96-97 keeps a stale
    value for the other types)."""
    rid, gid = ref["hit_id"], got["hit_id"]
    idm = (rid[:, :2] != gid[:, :2]).any(axis=1)
    idm |= (rid[:, 0] == 3) & (rid[:, 2] != gid[:, 2])
    tm = ref["hit_t"] != got["hit_t"]
    d8 = np.abs(ref["rgba8"].astype(np.int32) - got["rgba8"].astype(np.int32)).max(axis=1)
    out = dict(n=len(idm), id=int(idm.sum()), t=int(tm.sum()), u8_gt0=int((d8 > 0).sum()), u8_gt1=int((d8 > 1).sum()))
    if check_float:
        a, b = ref["rgba32f"], got["rgba32f"]
        out["f32"] = int(((a != b) & ~(np.isnan(a) & np.isnan(b))).any(axis=1).sum())
    return out
