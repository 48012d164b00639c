import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # a fresh checkout has no built library (built artefacts are git-ignored): build it once, as
    # __graft_entry__.build() does, instead of failing every import of the package
    lib = os.path.join(ROOT, "monogs_dev_b200", "libb2gs.so")
    if not os.path.exists(lib):
        import subprocess
        subprocess.check_call([sys.executable, os.path.join(ROOT, "monogs_dev_b200", "build.py")])


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)
