"""pytest configuration: registers the `gpu` marker and makes the package importable.

`-m "not gpu"` runs here (no GPU): oracle vs reference/golden vectors, host logic, C-ABI symbol checks.
`-m gpu` runs on a B200: CUDA path vs oracle and golden fixtures, always through the C-ABI.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_PARENT = os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200")
for p in (ROOT, PKG_PARENT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
