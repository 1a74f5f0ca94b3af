import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as orc

    orc.build()
    return orc


@pytest.fixture(scope="session")
def tmp_models(tmp_path_factory):
    """Seeded synthetic GGUF files shared by the session: name -> path."""
    from aios_b200 import gguf_synth as gs

    d = tmp_path_factory.mktemp("gguf")
    made = {}

    def get(arch: str, ftype: str = "Q4_K_M", seed: int = 7):
        key = (arch, ftype, seed)
        if key not in made:
            p = str(d / f"{arch}-{ftype}-{seed}.gguf")
            gs.write_synthetic_gguf(p, arch, ftype, seed)
            made[key] = p
        return made[key]

    return get


@pytest.fixture(scope="session")
def engine():
    from aios_b200 import capi

    e = capi.Engine(device=0, max_batch=4)
    yield e
    e.close()


@pytest.fixture(scope="session")
def engine_pfd():
    """Engine whose greedy decode goes through the persistent fused decode kernel (pfd.cu)."""
    from aios_b200 import capi

    e = capi.Engine(device=0, max_batch=2, flags=capi.ENGINE_USE_PFD)
    yield e
    e.close()
