"""pytest configuration: `-m "not gpu"` runs on any CPU box (oracle vs golden fixtures, host logic, C-ABI symbol
check); `-m gpu` needs a B200 and calls the CUDA library through its C ABI (ctypes)."""
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT / "nemotron-asr.cpp_b200", ROOT / "oracle", ROOT / "tools", ROOT / "tests", ROOT):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))

GOLD = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); runs the CUDA library through the C ABI")


@pytest.fixture(scope="session")
def gguf_q8():
    import inputs
    return inputs.gguf_path("q8_0")


@pytest.fixture(scope="session")
def gguf_f32():
    import inputs
    return inputs.gguf_path("f32dq")


@pytest.fixture(scope="session")
def oracle_q8(gguf_q8):
    import nemo_oracle as no
    return no.load_gguf(gguf_q8)


@pytest.fixture(scope="session")
def oracle_f32(gguf_f32):
    import nemo_oracle as no
    return no.load_gguf(gguf_f32)


@pytest.fixture(scope="session")
def lib_built():
    """Builds libnemob200.so in-tree if it is missing (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as ge
    ge.build_library()
    import nb200
    return nb200
