import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)


def pytest_configure(config):
  config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


REFERENCE_SUITE = "/root/reference/dm_control/suite"
has_reference = os.path.isdir(REFERENCE_SUITE)
needs_reference = pytest.mark.skipif(
    not has_reference, reason="/root/reference not present on this machine")
