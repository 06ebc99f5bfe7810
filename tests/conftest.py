"""pytest configuration: the `gpu` marker and shared fixtures.

`-m "not gpu"` tests run on a CPU-only box: the oracle against the golden vectors and the
reference's own host build (when present), the host logic, and the C-ABI symbol check.
`-m gpu` tests are the parity tests proper: they call the CUDA path through the C ABI and
compare with the oracle.  Nothing here reads /root/reference at run time.
"""
import os
import sys

import pytest

# the peer-exchange tests put up to 8 ranks on ONE device: each rank's stream needs its own
# hardware queue, or a spinning last block can sit in front of the kernel it is waiting for
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) device")


@pytest.fixture(scope="session")
def oracle():
    import oracle_py
    return oracle_py.Oracle()


@pytest.fixture(scope="session")
def ref():
    import oracle_py
    if not oracle_py.have_ref():
        pytest.skip("oracle/_ref/libref_cpu.so not built (needs /root/reference at build time)")
    return oracle_py.Ref()


@pytest.fixture(scope="session")
def gsm():
    from gsmfretda_b200 import capi
    capi.load()  # raises if the CUDA library is missing: no fallback
    return capi
