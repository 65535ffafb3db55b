import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port():
    from oracle.binding import Port
    return Port()


@pytest.fixture(scope="session")
def ref():
    from oracle.binding import Ref
    if not Ref.available():
        pytest.skip("oracle/_ref/libsalz_ref.so not built (reference tree absent)")
    return Ref()


@pytest.fixture(scope="session")
def orc():
    """The compiled reference (oracle/_ref/libsalz_ref.so: upstream salz + libsais built by
    oracle/Makefile).  GPU parity is pinned against it and nothing else: when the prebuilt
    library did not travel to this box the tests FAIL - they never fall back to the
    restatement in oracle/salz_oracle.c, which is this repository's own code."""
    from oracle.binding import Ref
    if not Ref.available():
        pytest.fail("oracle/_ref/libsalz_ref.so is absent: build it where /root/reference exists "
                    "(make -C oracle) and let it travel with the tree; refusing to check the CUDA "
                    "path against a self-written oracle")
    print("oracle: compiled reference (oracle/_ref/libsalz_ref.so)")
    return Ref()


@pytest.fixture(scope="session")
def enc():
    """One small CUDA context shared by the GPU tests (1 MiB blocks, batches of 8)."""
    import salz_b200
    e = salz_b200.SalzCuda(device=0, max_block_len=1 << 20, max_batch_blocks=8)
    yield e
    e.close()
