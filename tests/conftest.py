import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture
def workdir(tmp_path, monkeypatch):
    """The reference writes CellCenters.csv / FlowData_*.csv into the CWD (Grid.cpp:15, FS.cpp:588)."""
    monkeypatch.chdir(tmp_path)
    return tmp_path
