"""The driver-facing entry points keep their contract and their honesty."""
import json
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(args, **env):
    e = dict(os.environ, **env)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          env=e, timeout=300)


def test_bench_prints_one_json_line_with_the_contract_keys():
    p = run(["--steps", "2", "--warmup", "1", "--e2e-steps", "1", "--gpus", "1"])
    assert p.returncode == 0, p.stderr
    lines = [ln for ln in p.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "roofline",
                "cpu_baseline"):
        assert key in d
    assert d["steps"] == 2 and d["warmup"] >= 3 and d["value"] > 0 and d["data"] == "synthetic"
    # nothing may suggest GPU work was done
    assert d["gpu_launches"] == 0 and d["gpus_used"] == 0 and d["roofline"] is None and d["vs_baseline"] is None
    assert "N/A: no compute path" in d["metric"] and "stand-in" in d["config"]["status"]
    assert d["findings"]["planted"] == d["findings"]["found"] == d["findings"]["cli_found"] > 0
    assert d["findings"]["identical_to_planted"] is True and d["findings"]["mismatch"] is None
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["value"] > 0


def test_reference_arm_is_unavailable_and_says_why():
    p = run(["--impl", "reference", "--gpus", "1", "--steps", "1", "--warmup", "1"])
    d = json.loads(p.stdout)
    assert p.returncode == 0 and d["impl"] == "reference" and "terraform binary" in d["unavailable"]


def test_non_zero_ranks_stay_silent():
    p = run(["--gpus", "2", "--steps", "1"], RANK="1", WORLD_SIZE="2")
    assert p.returncode == 0 and p.stdout == ""


def test_graft_entry_build_and_smoke(capsys, monkeypatch):
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    monkeypatch.setenv(g._NESTED, "1")          # smoke() must not start the suite from inside the suite
    g.build()
    g.smoke()
    out = capsys.readouterr().out
    assert "no CUDA/C++ sources exist by design" in out and "no GPU kernel was launched" in out
    assert "cpu_suite: not run again" in out


def test_smoke_reports_the_cpu_suite_outcome(monkeypatch, tmp_path):
    """The line the driver's record keeps (GPUTEST smoke_tail): counts taken from a real
    child pytest -- here over two small files so the test stays quick."""
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    monkeypatch.delenv(g._NESTED, raising=False)
    line = g.cpu_suite(("tests/test_scanner.py", "tests/test_lockfile.py"))
    assert re.match(r"cpu_suite: \d+ passed, 0 skipped in [0-9.]+ s ", line), line
    assert int(line.split()[1]) > 20
    failing = tmp_path / "test_bad.py"
    failing.write_text("def test_no():\n    assert False\n")
    with pytest.raises(AssertionError, match="the CPU test suite failed"):
        g.cpu_suite((str(failing),))


def test_bench_compares_which_defects_were_found_not_how_many():
    sys.path.insert(0, ROOT)
    import bench
    report = {"modules": [{"diagnostics": [{"rule": "unused-variable", "file": "a/variables.tf"},
                                           {"rule": "cycle", "file": "a/main.tf"}]}],
              "fmt": {"changed": ["a/outputs.tf"]}}
    assert bench.identities(report) == [["cycle", "a/main.tf"], ["fmt", "a/outputs.tf"],
                                        ["unused-variable", "a/variables.tf"]]


def test_repository_contains_no_native_or_gpu_sources():
    bad = []
    for base, dirs, files in os.walk(ROOT):
        dirs[:] = [d for d in dirs if d not in (".git", "gpurun_out", "__pycache__")]
        bad += [f for f in files if f.endswith((".cu", ".cuh", ".cpp", ".so"))]
    assert bad == []
