"""Shared fixtures.  `gpu`-marked tests need a B200; everything else runs on CPU."""
import importlib.util
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_package():
    """Import the product package from its on-disk directory `our-tree_b200/`
    (a hyphen is not importable by name) as module `our_tree_b200`."""
    if "our_tree_b200" in sys.modules:
        return sys.modules["our_tree_b200"]
    path = os.path.join(ROOT, "our-tree_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location(
        "our_tree_b200", path, submodule_search_locations=[os.path.dirname(path)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["our_tree_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session", autouse=True)
def _built_artifacts():
    """The in-tree builds normally travel with the snapshot; on a checkout without them
    (fresh clone) compile once — nvcc cross-compiles without a GPU.  A failed build is not
    hidden: the tests that need the library then fail loudly in `lib()`."""
    lib = os.path.join(ROOT, "our-tree_b200", "libbulkcipher.so")
    harness = os.path.join(ROOT, "our-tree_b200", "build", "aes_test")
    if not (os.path.exists(lib) and os.path.exists(harness)):
        try:
            import __graft_entry__ as g
            g.build()
        except Exception as e:  # noqa: BLE001
            print(f"[conftest] build() failed: {e}")
    yield


@pytest.fixture(scope="session")
def orc():
    import oracle
    return oracle.Oracle()


@pytest.fixture(scope="session")
def ref():
    import oracle
    if not oracle.Ref.available():
        try:
            oracle.build()
        except Exception:
            pass
    if not oracle.Ref.available():
        pytest.skip("oracle/_ref/libref.so not built (no /root/reference here)")
    return oracle.Ref()


@pytest.fixture(scope="session")
def golden():
    g = {}
    for name in ("kat", "digests", "vectors"):
        with open(os.path.join(ROOT, "tests", "golden", name + ".json")) as f:
            g[name] = json.load(f)
    return g


@pytest.fixture(scope="session")
def pkg():
    return load_package()


@pytest.fixture(scope="session")
def harness_input(orc):
    """The 16 MiB BASELINE config-1 input (srand(1337); rand()%255) and its 32-byte key."""
    return orc.harness_input(16 << 20)
