"""Test configuration.

There are deliberately NO ``gpu``-marked tests: the reference has no compute
path (SURVEY.md sections 0 and 8), so nothing here runs on a GPU.  The marker is still
registered because the driver invokes ``pytest -m gpu``; an empty selection is
turned into a successful run with an explicit message instead of pytest's
"no tests collected" failure code.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE = os.environ.get("TF_REFERENCE_ROOT", "/root/reference")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (none exist in this repo; see DESIGN.md)")
    config.addinivalue_line("markers", "reference: reads /root/reference (skipped where it is absent)")
    config.addinivalue_line("markers", "terraform: runs the real terraform / terraform-docs binaries beside the "
                                       "stand-ins (skipped until $TF_BIN or PATH supplies them)")
    config.addinivalue_line("filterwarnings", "ignore:Generating overly large repr")


def pytest_sessionfinish(session, exitstatus):
    if exitstatus == pytest.ExitCode.NO_TESTS_COLLECTED and session.config.getoption("-m") == "gpu":
        tr = session.config.pluginmanager.get_plugin("terminalreporter")
        if tr:
            tr.write_line("no gpu tests by design: the reference has no compute path, "
                          "so this repository ships no GPU code (DESIGN.md)")
        session.exitstatus = pytest.ExitCode.OK


@pytest.fixture(scope="session")
def reference_root():
    if not os.path.isdir(os.path.join(REFERENCE, "eks")):
        pytest.skip("reference tree not present on this machine")
    return REFERENCE


def write_module(tmp_path, files: dict) -> str:
    for rel, text in files.items():
        p = tmp_path / rel
        p.parent.mkdir(parents=True, exist_ok=True)
        p.write_text(text)
    return str(tmp_path)


@pytest.fixture
def module_dir(tmp_path):
    def make(files: dict) -> str:
        return write_module(tmp_path, files)
    return make
