"""Command line entry point, run the way a maintainer's hook would run it."""
import json
import os
import subprocess
import sys

import pytest

from nvidia_terraform_modules_b200.synth import write_tree

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def cli(args, cwd):
    env = dict(os.environ, PYTHONPATH=ROOT)
    return subprocess.run([sys.executable, "-m", "nvidia_terraform_modules_b200", *args], cwd=cwd, env=env,
                          capture_output=True, text=True, timeout=120)


@pytest.fixture(scope="module")
def trees(tmp_path_factory):
    base = tmp_path_factory.mktemp("cli")
    write_tree(str(base / "clean"), seed=11, modules=1)
    write_tree(str(base / "bad"), seed=11, modules=1, defects=8)
    return base


def test_help_lists_commands(trees):
    p = cli(["--help"], trees)
    assert p.returncode == 0 and all(c in p.stdout for c in ("fmt", "validate", "lint", "docs", "check", "report"))
    assert cli(["frobnicate"], trees).returncode == 1


def test_fmt_exit_codes(trees):
    assert cli(["fmt", "-check", "-recursive", "."], trees / "clean").returncode == 0
    p = cli(["fmt", "-check", "-recursive", "."], trees / "bad")
    assert p.returncode == 3
    assert p.stdout.split() == ["cloud0/outputs.tf", "cloud0/variables.tf", "cloud0/versions.tf"]


def test_validate_text_and_json(trees):
    p = cli(["validate", "cloud0"], trees / "clean")
    assert p.returncode == 0 and "Success! The configuration is valid." in p.stdout and "stand-in" in p.stdout
    p = cli(["validate", "-json", "cloud0"], trees / "bad")
    data = json.loads(p.stdout)
    assert p.returncode == 1 and data["valid"] is False and data["error_count"] == 2
    assert {d["summary"] for d in data["diagnostics"]} == {"Reference to undeclared input variable",
                                                          "Call to unknown function"}
    assert data["diagnostics"][0]["range"]["filename"] == "cloud0/outputs.tf"


def test_lint_lock_docs(trees):
    assert cli(["lint", "cloud0"], trees / "clean").returncode == 0
    p = cli(["lint", "--json", "cloud0"], trees / "bad")
    assert p.returncode == 1
    assert sorted(d["rule"] for d in json.loads(p.stdout)) == ["length-of-object", "unused-local", "unused-variable"]
    assert cli(["lint", "--list-rules"], trees).stdout.count("\n") >= 20
    assert cli(["lock", "cloud0", "cloud0/examples/basic"], trees / "clean").returncode == 0
    assert cli(["docs", "--check", "cloud0"], trees / "clean").returncode == 0
    p = cli(["docs", "cloud0"], trees / "clean")
    assert p.stdout.startswith("## Requirements") and "## Outputs" in p.stdout


def test_check_and_report(trees, tmp_path):
    p = cli(["check", "--json", "."], trees / "clean")
    data = json.loads(p.stdout)
    assert p.returncode == 0 and data["ok"] and data["counts"] == {"error": 0, "warning": 0, "notice": 0,
                                                                   "fmt_differing": 0}
    assert [c["status"] for c in data["baseline_configs"]] == ["stand-in", "not run", "unavailable", "unavailable",
                                                               "unavailable"]
    p = cli(["check", "."], trees / "bad")
    assert p.returncode == 1 and "3 unformatted, 2 errors" in p.stdout
    assert cli(["check", "--strict", "."], trees / "clean").returncode == 0
    out = tmp_path / "r.md"
    assert cli(["report", "-o", str(out), "."], trees / "bad").returncode == 0
    text = out.read_text()
    assert text.startswith("# Static-check findings") and "```diff" in text and "| `unused-variable` |" in text
