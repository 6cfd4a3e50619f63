import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_sessionstart(session):
    """A fresh checkout has no built artefacts (they are git-ignored): build them once, the way __graft_entry__.build() does.
    On the GPU box they arrive with the snapshot and nothing happens here."""
    need = [os.path.join(ROOT, "navierstokessolver_b200", "libnsgpu.so"), os.path.join(ROOT, "tests", "emul", "libnsemul.so"),
            os.path.join(ROOT, "bin", "main")]
    if os.path.isdir("/root/reference/SRC"):
        need += [os.path.join(ROOT, "oracle", "_ref", "libnsref.so"), os.path.join(ROOT, "oracle", "_ref", "main_ref")]
    if all(os.path.exists(p) for p in need):
        return
    import subprocess
    subprocess.run(["make", "-C", ROOT, "all", "emul"], check=True, stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/SRC"):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True, stdout=subprocess.DEVNULL)


@pytest.fixture
def workdir(tmp_path, monkeypatch):
    """The reference writes CellCenters.csv / FlowData_*.csv into the CWD (Grid.cpp:15, FS.cpp:588)."""
    monkeypatch.chdir(tmp_path)
    return tmp_path
