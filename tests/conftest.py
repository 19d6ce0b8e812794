import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")
    config.addinivalue_line("markers", "slow: multi-process / long running")
    config.addinivalue_line("markers", "tensorcores: run with the tcgen05 routing enabled")


def reference_src():
    """Directory holding the UNMODIFIED reference sources, or None (tests then skip)."""
    for cand in (os.path.join(ROOT, "baseline", "_ref", "fedref_src"),
                 os.path.join(ROOT, "baseline", "_ref"), "/root/reference/src"):
        if os.path.isfile(os.path.join(cand, "models", "lenet.py")):
            return cand
    return None


@pytest.fixture(scope="session")
def ref_models():
    """The reference's ``models`` package imported under a private name."""
    src = reference_src()
    if src is None:
        pytest.skip("reference sources not available")
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "ref_models", os.path.join(src, "models", "__init__.py"),
        submodule_search_locations=[os.path.join(src, "models")])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_models"] = mod
    try:
        spec.loader.exec_module(mod)
    except Exception:
        # models/__init__ star-imports shufflenet etc.; all import fine, constructors may not
        raise
    return mod
