import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)


def pytest_configure(config):
  config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def scene1():
  from foundationpose_b200 import synthetic
  return synthetic.make_scene(1, n_points=600000)


@pytest.fixture(scope='session')
def scene2():
  from foundationpose_b200 import synthetic
  return synthetic.make_scene(2, n_points=600000)
