"""pytest configuration.  `-m "not gpu"`: oracle vs golden vectors, host logic, ABI surface (runs on CPU in
minutes).  `-m gpu`: the parity tests proper, CUDA path vs oracle through the C ABI (need a B200)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)

import util  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the product library and the oracle restatement exist (build() is idempotent)."""
    from monte_carlo_path_tracing_b200 import api
    import oracle
    if not os.path.exists(api.LIB_PATH):
        api.build()
    if not os.path.exists(oracle.ORACLE_SO):
        oracle.build(ref=False)


@pytest.fixture(scope="session", params=list(util.SCENE_KW))
def scene_name(request):
    return request.param


@pytest.fixture(scope="session")
def golden():
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = dict(np.load(os.path.join(util.GOLDEN_DIR, name + ".npz")))
        return cache[name]
    return get


@pytest.fixture(scope="session")
def scene_files(tmp_path_factory):
    d = tmp_path_factory.mktemp("scenes")
    cache = {}

    def get(name):
        if name not in cache:
            sc = util.scene_text(name)
            cache[name] = (sc, sc.write(str(d)))
        return cache[name]
    return get


@pytest.fixture(scope="session")
def arrays_of(scene_files):
    """SceneArrays of a named small scene through OUR host loaders (no GPU needed)."""
    from monte_carlo_path_tracing_b200 import api
    cache = {}

    def get(name):
        if name not in cache:
            sc, stem = scene_files(name)
            a, cam = api.load_scene_files(stem + ".obj", stem + ".xml")
            cache[name] = (a, sc.camera)
        return cache[name]
    return get


@pytest.fixture(scope="session")
def oracle_of(arrays_of):
    import oracle
    cache = {}

    def get(name):
        if name not in cache:
            a, cam = arrays_of(name)
            o = oracle.OracleLib(a)
            o.build_grid(cam.eye, 100000)
            cache[name] = o
        return cache[name]
    return get


@pytest.fixture(scope="session")
def gpu_scene_of(arrays_of):
    from monte_carlo_path_tracing_b200 import api
    cache = {}

    def get(name):
        if name not in cache:
            a, cam = arrays_of(name)
            cache[name] = api.Scene(a, 0)
        return cache[name]
    yield get
    for s in cache.values():
        s.close()
