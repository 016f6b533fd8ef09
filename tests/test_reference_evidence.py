"""What is known about the reference from OUTSIDE this code base, held against what
the checker finds.  Nothing in this file compares the code with itself (that is
tests/test_reference_regression.py).

External sources, in the order used below:

1. ``tests/golden/survey_expectations.json`` -- hand-written from SURVEY.md, which was
   produced by reading the reference before any of this code existed: the files
   ``terraform fmt`` is predicted to rewrite, the defects found by reading (file:line),
   the block counts.
2. ``tests/golden/reference_measurements.json`` -- counts measured against bytes the
   reference itself holds: README rows written by the real terraform-docs, and the
   exact text of the 33 files its authors left ``terraform fmt``-clean.
3. Where /root/reference is mounted (marker ``reference``): the same measurements
   taken again, live.

The checker's findings come from ``tests/snapshots/regression_snapshot.json`` so that
layers 1 and 2 also run on machines without the reference; the regression test ties
that snapshot to a live run.
"""
import json
import os
import random

import pytest

HERE = os.path.dirname(__file__)


def _load(*parts):
    with open(os.path.join(HERE, *parts)) as fh:
        return json.load(fh)


@pytest.fixture(scope="module")
def survey():
    return _load("golden", "survey_expectations.json")


@pytest.fixture(scope="module")
def measured():
    return _load("golden", "reference_measurements.json")


@pytest.fixture(scope="module")
def snapshot():
    return _load("snapshots", "regression_snapshot.json")


SURVEY = _load("golden", "survey_expectations.json")
DEFECTS = [(e["module"], e["rule"], e["file"], e["line"]) for e in SURVEY["defects_found_by_reading"]["entries"]]


# ------------------------------------------- 1. the survey's reading vs. the findings
def test_fmt_flags_exactly_the_files_the_survey_predicted(survey, snapshot):
    assert snapshot["fmt_changed"] == survey["fmt_files_predicted_to_differ"]["files"]
    assert len(snapshot["fmt_checked"]) == survey["formattable_files"]["count"]


@pytest.mark.parametrize("module,rule,file,line", DEFECTS)
def test_every_defect_the_survey_found_by_reading_is_reported(snapshot, module, rule, file, line):
    assert [rule, file, line] in snapshot["modules"][module]["findings"]


@pytest.mark.parametrize("module", sorted(k for k in SURVEY["block_counts"] if k != "source"))
def test_block_counts_match_the_survey(survey, snapshot, module):
    got = snapshot["modules"][module]["counts"]
    want = survey["block_counts"][module]
    assert {k: got[k] for k in want} == want


def test_modules_known_to_deploy_have_no_structural_or_reference_errors(snapshot):
    """The six modules are published as working; a checker that reports an undeclared
    reference or a malformed block in them is wrong."""
    rules = {f[0] for m in snapshot["modules"].values() for f in m["findings"]}
    assert not rules & {"syntax", "structure", "undeclared-variable", "undeclared-resource", "undeclared-local",
                        "undeclared-module", "instance-key", "function", "module-arguments", "module-output",
                        "cycle", "self-reference", "type"}


def test_each_module_validates_under_its_own_required_version_floor(snapshot):
    """eks (>= 1.2.4) and aks (>= 1.3.4) are clean under the oldest release they admit.
    gke says ">= 0.14" (gke/versions.tf:15) but calls one() three times
    (gke/main.tf:40,63,109), a function Terraform 0.15 introduced: under 0.14 those are
    the only errors, and they are the reference's, not the checker's."""
    versions = {m: v["terraform_version"] for m, v in snapshot["modules"].items()}
    assert versions == {"eks": "1.2", "gke": "0.14", "aks": "1.3", "eks/examples/cnpack": "1.10",
                        "gke/examples/cnpack": "1.10", "aks/examples/cnpack": "1.10"}
    gated = {m: [f for f in v["findings"] if f[0] == "terraform-version"] for m, v in snapshot["modules"].items()}
    assert gated.pop("gke") == [["terraform-version", "gke/main.tf", n] for n in (40, 63, 109)]
    assert not any(gated.values())


# ------------------------------ 2. measured against what the reference itself holds
def test_readme_rows_reproduce_real_terraform_docs_output(survey, measured):
    rows = measured["docs_rows"]
    for module in survey["readmes_the_survey_found_up_to_date"]["modules"]:
        assert rows[module]["identical_in_readme"] == rows[module]["generated"] > 0
        assert measured["docs_write_is_a_no_op"][module] is True
    # the two READMEs the survey found stale: every generated row that the README lacks is
    # accounted for by name
    assert rows["eks/examples/cnpack"] == {
        "generated": 35, "identical_in_readme": 34,
        "generated_rows_absent_from_readme": [["Inputs", "metrics_server_enabled"]]}    # Readme.md:229 vs variables.tf:58
    gke = rows["gke/examples/cnpack"]
    assert (gke["generated"], gke["identical_in_readme"]) == (15, 9)
    assert gke["generated_rows_absent_from_readme"] == [
        # the README's Providers table predates the lock file (google "n/a") and random_string
        ["Providers", "google"], ["Providers", "random"],
        # README.md:87-91 documents one of the resources/data sources declared
        ["Resources", "google_project_iam_binding.prometheus_service_account"],
        ["Resources", "google_service_account_iam_binding.prometheus_service_account"],
        ["Resources", "random_string.gke"], ["Resources", "google_client_config.gke"]]
    total = sum(r["generated"] for r in rows.values())
    same = sum(r["identical_in_readme"] for r in rows.values())
    assert (total, same) == (226, 219)


def test_version_constraints_are_respelt_the_way_terraform_init_records_them(measured):
    """eks/versions.tf writes "~>4.45.0"; `terraform init` recorded "~> 4.45.0" in
    eks/.terraform.lock.hcl.  All eight constrained providers across the six modules agree."""
    rec = measured["lock_constraints"]
    assert sum(r["required_providers_with_a_constraint"] for r in rec.values()) == 8
    for module, r in rec.items():
        assert r["canonical_spelling_found_in_lock_file"] == r["required_providers_with_a_constraint"], module


def test_the_files_the_authors_left_canonical_are_fixed_points(measured):
    assert measured["fmt_fixed_points"] == {"files_left_canonical_by_the_authors": 33, "unchanged_by_fmt": 33}


def test_damaged_white_space_formats_back_to_the_authors_bytes(measured):
    """The strongest offline evidence for fmt: 33 files x 25 seeds x 2 kinds of damage,
    every damaged copy different from the original, every one restored byte for byte."""
    rec = measured["fmt_scramble_recovery"]
    n = rec["seeds_per_file"]
    assert n >= 20 and len(rec["files"]) == 33
    for rel, f in rec["files"].items():
        assert (f["lines_recovered"], f["tokens_recovered"]) == (n, n), rel
        assert f["damaged_copies_that_differed_from_the_original"] == 2 * n, rel


# --------------------------------------------------------- 3. the same, taken live
def _clean_files(snapshot):
    return [p for p in snapshot["fmt_checked"] if p not in snapshot["fmt_changed"]]


@pytest.mark.reference
def test_live_files_the_authors_formatted_are_fixed_points(reference_root, snapshot):
    from nvidia_terraform_modules_b200.fmt.command import format_source
    clean = _clean_files(snapshot)
    assert len(clean) == 33
    for rel in clean:
        with open(os.path.join(reference_root, rel), newline="") as fh:
            src = fh.read()
        assert format_source(src, rel)[0] == src, rel


@pytest.mark.reference
def test_live_scramble_recovery_line_damage(reference_root, snapshot):
    """Random indent (tabs included), 0-3 blanks round every ``=``, trailing-comment gaps
    collapsed or stretched, trailing blanks: the authors' bytes must come back."""
    from scramble import damage_lines, token_texts
    from nvidia_terraform_modules_b200.fmt.command import format_source
    for rel in _clean_files(snapshot):
        with open(os.path.join(reference_root, rel), newline="") as fh:
            src = fh.read()
        differed = 0
        for seed in range(20):
            bad = damage_lines(src, random.Random(f"{rel}/live-lines/{seed}"))
            assert token_texts(bad) == token_texts(src)
            differed += bad != src
            got, diags = format_source(bad, rel)
            assert not diags and got == src, (rel, seed)
        assert differed >= 15, rel      # (a file of four comment lines offers few gaps to damage)


@pytest.mark.reference
def test_live_scramble_recovery_token_damage(reference_root, snapshot):
    """Every gap between two tokens outside templates re-drawn at random."""
    from scramble import damage_tokens, token_texts
    from nvidia_terraform_modules_b200.fmt.command import format_source
    for rel in _clean_files(snapshot):
        with open(os.path.join(reference_root, rel), newline="") as fh:
            src = fh.read()
        differed = 0
        for seed in range(20):
            bad = damage_tokens(src, random.Random(f"{rel}/live-tokens/{seed}"))
            assert token_texts(bad) == token_texts(src)
            differed += bad != src
            got, diags = format_source(bad, rel)
            assert not diags and got == src, (rel, seed)
        assert differed >= 15, rel      # (a file of four comment lines offers few gaps to damage)


@pytest.mark.reference
def test_live_scrambling_the_five_reformatted_files_gives_one_answer(reference_root, snapshot):
    """For the files fmt does rewrite there is no author-held answer; what can be said is
    that the answer does not depend on the layout it started from."""
    from scramble import damage_lines, damage_tokens
    from nvidia_terraform_modules_b200.fmt.command import format_source
    for rel in snapshot["fmt_changed"]:
        with open(os.path.join(reference_root, rel), newline="") as fh:
            src = fh.read()
        want = format_source(src, rel)[0]
        for seed in range(10):
            for fn in (damage_lines, damage_tokens):
                bad = fn(want, random.Random(f"{rel}/{fn.__name__}/{seed}"))
                assert format_source(bad, rel)[0] == want, (rel, fn.__name__, seed)


@pytest.mark.reference
def test_live_measurements_equal_the_committed_ones(reference_root, snapshot, measured):
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_reference_records",
                                                  os.path.join(HERE, "make_reference_records.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert json.loads(json.dumps(mod.measurements(reference_root, snapshot))) == measured


@pytest.mark.reference
def test_live_every_reference_hcl_file_round_trips_through_the_scanner(reference_root, survey):
    from nvidia_terraform_modules_b200.hcl.parser import parse
    from nvidia_terraform_modules_b200.hcl.tokens import render
    seen = 0
    for base, dirs, files in os.walk(reference_root):
        dirs[:] = [d for d in dirs if d != ".git"]
        for name in files:
            if name.endswith((".tf", ".tfvars", ".hcl")):
                with open(os.path.join(base, name), newline="") as fh:
                    src = fh.read()
                body, diags, toks = parse(src, name)
                assert not diags, (name, diags)
                assert render(toks, keep_raw=True) == src
                seen += 1
    assert seen == survey["hcl_files"]["count"]


@pytest.mark.reference
@pytest.mark.parametrize("module", SURVEY["readmes_the_survey_found_up_to_date"]["modules"])
def test_live_regenerating_an_up_to_date_readme_changes_nothing(reference_root, module):
    """The whole generated region -- headings, table headers, row order, blank
    lines -- equals what real terraform-docs wrote into these four READMEs."""
    from nvidia_terraform_modules_b200.config.lockfile import load_lockfile
    from nvidia_terraform_modules_b200.config.model import load_module
    from nvidia_terraform_modules_b200.docs.tfdocs import find_readme, inject, render
    d = os.path.join(reference_root, module)
    with open(find_readme(d), encoding="utf-8") as fh:
        text = fh.read()
    assert inject(text, render(load_module(d, reference_root), load_lockfile(d, reference_root))) == text
