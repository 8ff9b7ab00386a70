import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


@pytest.fixture(autouse=True)
def _lowest_index_tie_rule():
    """The C oracle's `fps_lowidx` and the committed goldens use the lowest-index FPS tie rule, so that is the host-side
    default INSIDE the test-suite; the package default is 'reference' (tests/test_host.py checks it) and the tests of
    the reference kernel's own order switch to it explicitly (`reference_tie_rule`)."""
    import deepclr_b200
    before = deepclr_b200.get_fps_tie_rule()
    deepclr_b200.set_fps_tie_rule('lowest_index')
    yield
    deepclr_b200.set_fps_tie_rule(before)


@pytest.fixture
def reference_tie_rule(_lowest_index_tie_rule):
    import deepclr_b200
    deepclr_b200.set_fps_tie_rule('reference')
    yield
    deepclr_b200.set_fps_tie_rule('lowest_index')


@pytest.fixture(scope='session')
def golden_dir():
    return os.path.join(ROOT, 'tests', 'golden')
