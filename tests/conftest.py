import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box: pytest -m gpu)")
    config.addinivalue_line("markers", "ref: needs oracle/_ref/libref_oracle.so (the compiled, unmodified reference)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """librtb.so must exist (built in-tree by __graft_entry__.build()); there is no fallback path."""
    import _util
    if not os.path.exists(_util.rtb.LIB_PATH):
        sys.path.insert(0, _util.ROOT)
        import __graft_entry__
        __graft_entry__.build()
    yield


@pytest.fixture(scope="session")
def ref():
    """The compiled reference (oracle/_ref).  Built where /root/reference exists; travels to the GPU box prebuilt."""
    from oracle import ref_oracle
    if not ref_oracle.available():
        pytest.skip("oracle/_ref/libref_oracle.so not built (needs /root/reference)")
    return ref_oracle


_host_scenes = {}


@pytest.fixture(scope="session")
def host_scene():
    import _util

    def get(name, width=64, height=48, samples=4):
        key = (name, width, height, samples)
        if key not in _host_scenes:
            _host_scenes[key] = _util.rtb.HostScene.load_gltf(_util.scene_path(name), width, height, samples)
        return _host_scenes[key]

    return get
