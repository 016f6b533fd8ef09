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
    rng = data["diagnostics"][0]["range"]
    assert rng["filename"] == "outputs.tf"                        # relative to the directory argument
    assert rng["abs_filename"] == str(trees / "bad" / "cloud0" / "outputs.tf")
    assert data["terraform_version"] == "1.3" and "required_version" in data["terraform_version_source"]


def test_file_names_are_relative_to_the_directory_argument_from_any_cwd(trees, tmp_path):
    """Round-1 verdict: validating /tmp/x from elsewhere printed ``on ../../tmp/x/main.tf:3``."""
    elsewhere = tmp_path / "deep" / "er"
    elsewhere.mkdir(parents=True)
    bad = trees / "bad" / "cloud0"
    p = cli(["validate", str(bad)], elsewhere)
    assert "  on outputs.tf:" in p.stderr and ".." not in p.stderr
    p = cli(["docs", "--check", str(trees / "clean" / "cloud0")], elsewhere)
    assert p.returncode == 0 and ".." not in p.stdout
    p = cli(["lint", "--json", str(bad)], elsewhere)
    assert {d["file"] for d in json.loads(p.stdout)} <= {"main.tf", "variables.tf", "outputs.tf"}
    assert all(d["abs_file"] == str(bad / d["file"]) for d in json.loads(p.stdout))
    # several directories: names stay distinguishable, still no climbing out
    p = cli(["lint", "--json", str(bad), str(trees / "bad" / "cloud0" / "examples" / "basic")], elsewhere)
    assert all(not d["file"].startswith("..") for d in json.loads(p.stdout))
    # a child module outside the directory is the one case that needs "..", as in Terraform
    p = cli(["check", "--json", str(trees / "bad")], elsewhere)
    files = {d["file"] for m in json.loads(p.stdout)["modules"] for d in m["diagnostics"]}
    assert files and all(f.startswith("cloud0/") for f in files)


def test_terraform_version_option(trees):
    mod = trees / "gated"
    mod.mkdir()
    (mod / "main.tf").write_text('moved {\n  from = a_b.c\n  to   = a_b.d\n}\nresource "a_b" "d" {}\n')
    assert cli(["validate", str(mod)], trees).returncode == 0
    p = cli(["validate", "--terraform-version", "1.0", str(mod)], trees)
    assert p.returncode == 1 and 'Blocks of type "moved" are not expected here.' in p.stderr
    assert "needs Terraform 1.1 or later; this run restates Terraform 1.0 (--terraform-version)" in p.stderr
    assert cli(["validate", "-terraform-version=1.1.9", str(mod)], trees).returncode == 0
    p = cli(["validate", "--terraform-version", "latest", str(mod)], trees)
    assert p.returncode == 1 and "not a Terraform version" in p.stderr
    p = cli(["check", "--terraform-version", "1.0", str(mod)], trees)
    assert p.returncode == 1 and "validated as Terraform 1.0 (--terraform-version)" in p.stdout
    p = cli(["validate", "--frobnicate", str(mod)], trees)
    assert p.returncode == 1 and "unknown option" in p.stderr


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
