"""Does the dormant parity harness itself work?  (Plumbing only -- NOT parity.)

tests/test_against_terraform.py cannot run here: there is no Terraform binary.  A
harness that has never run is a harness with typos in it, so this test points ``$TF_BIN``
at ``tests/support/fake_terraform.py`` -- a shim that answers Terraform's command line
with this repository's own code -- and runs the harness in a child pytest.  Every
comparison in that child is the code against itself, so its being green proves only
that the process calls, JSON shapes, corpus building and comparisons are sound, and
that the harness will produce a real verdict the day ``$TF_BIN`` names a real binary.
"""
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _run_harness(version: str, *extra: str) -> subprocess.CompletedProcess:
    env = dict(os.environ, TF_BIN=os.path.join(HERE, "support", "fake_terraform.py"),
               FAKE_TERRAFORM_VERSION=version, PYTHONPATH=ROOT, TF_HARNESS_FMT_SAMPLE="60")
    env.pop("TF_PLUGIN_MIRROR", None)
    env["TFDOCS_BIN"] = os.path.join(HERE, "support", "fake_terraform_docs.py")
    return subprocess.run([sys.executable, "-m", "pytest", os.path.join(HERE, "test_against_terraform.py"),
                           "-q", "-m", "terraform", "-p", "no:cacheprovider", "-rs", *extra],
                          cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)


def test_harness_runs_end_to_end_against_the_shim():
    p = _run_harness("1.10.0")
    tail = p.stdout[-3000:] + p.stderr[-2000:]
    assert p.returncode == 0, tail
    m = re.search(r"(\d+) passed", p.stdout)
    assert m and int(m.group(1)) >= 95, tail          # fmt corpus + command + questions + validate cases + docs
    assert "failed" not in p.stdout.splitlines()[-1], tail
    assert "$TF_PLUGIN_MIRROR not set" in p.stdout      # provider-backed cases say why they did not run


def test_harness_follows_the_release_the_binary_reports():
    """Under an older release the gated cases (check blocks, optional() defaults, ...)
    must still agree, because the harness hands ``terraform version`` to ``validate``."""
    p = _run_harness("1.2.9", "-k", "validate")
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
