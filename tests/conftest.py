import os
import sys
from pathlib import Path

import pytest

# The multi-GPU tests put several replicas (each with its own streams) on ONE GPU and let a spin kernel wait
# for flags that kernels on other streams publish.  With the default 8 hardware work queues two streams can
# share a queue, and a kernel queued behind the spinning one would never start (false dependency): give every
# stream its own queue.  Must be set before the CUDA context exists.  Real multi-GPU runs do not need it:
# the frame owner's three streams never share a queue with the peers' (they are other devices).
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def ctx():
    from simpleraytracer_b200 import Context
    c = Context()
    yield c
    c.release()
