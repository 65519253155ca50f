"""pytest configuration: the ``gpu`` marker and shared synthetic-input builders."""

import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)


def pytest_configure(config):
  config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def pytest_collection_modifyitems(config, items):
  """GPU tests are selected with ``-m gpu``; without a device they are skipped, never faked."""
  try:
    import torch

    has_gpu = torch.cuda.is_available()
  except Exception:  # pragma: no cover
    has_gpu = False
  if has_gpu:
    return
  skip = pytest.mark.skip(reason="no CUDA device in this container")
  for item in items:
    if "gpu" in item.keywords:
      item.add_marker(skip)


@pytest.fixture(scope="session")
def rng():
  return np.random.default_rng(1234)
