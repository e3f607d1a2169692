import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
# libgomp worker threads of the C oracle must not spin against OpenBLAS's pool
os.environ.setdefault("OMP_WAIT_POLICY", "passive")

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def fixture_index_bytes():
    with open(os.path.join(GOLDEN, "docqa_vector_store.faiss"), "rb") as f:
        return f.read()


@pytest.fixture(scope="session")
def fixture_corpus(fixture_index_bytes):
    from oracle import faiss_io

    d, n, metric, x = faiss_io.read_index_bytes(fixture_index_bytes)
    assert (d, n, metric) == (384, 649, 1)
    return x


def note(line: str) -> None:
    """Measured values a test wants on record (GPU count seen, fallback counts ...): appended to
    gpurun_out/test_notes.log when that directory exists (it travels back from the GPU box)."""
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "test_notes.log"), "a") as f:
            f.write(line.rstrip() + "\n")
