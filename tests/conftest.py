import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def params():
    import yaml
    from rotors_mpc_controller_b200.params import default_params_path
    return yaml.safe_load(open(default_params_path()))


@pytest.fixture(scope='session')
def built():
    import __graft_entry__ as g
    g.build()
    return True
