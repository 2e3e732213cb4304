import os
import sys
import pathlib

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

REFERENCE_SRC = pathlib.Path("/root/reference/src")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
        n_gpu = torch.cuda.device_count() if has_gpu else 0
    except Exception:
        has_gpu, n_gpu = False, 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and not has_gpu:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and n_gpu < 2:
            item.add_marker(skip_multi)


def load_reference_module(part: str, name: str):
    """Import a file of the read-only reference as a module (the CPU oracle).
    Looks in /root/reference first, then baseline/_ref (which travels to the GPU box)."""
    import importlib.util
    for base in (REFERENCE_SRC, ROOT / "baseline" / "_ref" / "src"):
        path = base / f"Part {part}" / f"{name}.py"
        if path.exists():
            spec = importlib.util.spec_from_file_location(f"_ref_part{part}_{name}", path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod
    pytest.skip("reference sources not available on this machine")


@pytest.fixture(scope="session")
def ref_model_module():
    return load_reference_module("1", "model")
