"""REGRESSION ONLY: the checker compared with its own earlier output.

``tests/snapshots/regression_snapshot.json`` is written by the checker
(``tests/make_reference_records.py``).  A green run here says the code still behaves
as it did when the snapshot was taken -- nothing more.  It is NOT evidence of parity
with Terraform; that evidence, such as it is, lives in tests/test_reference_evidence.py.
"""
import importlib.util
import json
import os

import pytest

HERE = os.path.dirname(__file__)
SNAPSHOT = os.path.join(HERE, "snapshots", "regression_snapshot.json")


@pytest.fixture(scope="module")
def snapshot():
    with open(SNAPSHOT) as fh:
        return json.load(fh)


def test_snapshot_declares_itself_self_generated(snapshot):
    assert snapshot["_about"].startswith("SELF-GENERATED")
    assert set(snapshot) == {"_about", "fmt_checked", "fmt_changed", "fmt_output_sha256", "modules"}


@pytest.mark.reference
def test_live_run_equals_snapshot(reference_root, snapshot):
    spec = importlib.util.spec_from_file_location("make_reference_records",
                                                  os.path.join(HERE, "make_reference_records.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    live = json.loads(json.dumps(mod.snapshot(reference_root)))
    assert live == snapshot


@pytest.mark.reference
def test_formatting_the_flagged_files_is_idempotent_and_white_space_only(reference_root, snapshot):
    from nvidia_terraform_modules_b200.fmt.command import format_source
    from nvidia_terraform_modules_b200.hcl.scanner import scan
    for rel in snapshot["fmt_changed"]:
        with open(os.path.join(reference_root, rel), newline="") as fh:
            src = fh.read()
        once = format_source(src, rel)[0]
        assert format_source(once, rel)[0] == once
        assert [(t.type, t.text) for t in scan(once)] == [(t.type, t.text) for t in scan(src)]
