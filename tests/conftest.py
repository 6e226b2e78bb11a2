"""pytest configuration: the `gpu` marker and import paths.

`-m "not gpu"` runs here (no GPU): oracle vs known answers / golden fixtures, host
logic, and that the C-ABI library loads and exports every declared symbol.
`-m gpu` tests are the parity tests proper and call through the C ABI.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "k8s-gpu-device-plugin_b200")
for p in (ROOT, PKG, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    return oracle_lib.load()


@pytest.fixture(scope="session")
def probe():
    """One gp_ctx for the whole GPU session (arena kept between tests)."""
    import gpuprobe
    p = gpuprobe.Probe(keep_arena=1, hbm_bytes=1 << 30, arena_cap_bytes=100 << 30)
    yield p
    p.close()


@pytest.fixture(scope="session")
def probe_vec():
    import gpuprobe
    p = gpuprobe.Probe(keep_arena=1, hbm_bytes=1 << 30, arena_cap_bytes=100 << 30,
                       stream_variant=gpuprobe.GP_STREAM_VEC)
    yield p
    p.close()
