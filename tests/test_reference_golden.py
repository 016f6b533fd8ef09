"""Findings on the reference tree, pinned.

Two layers:

* everywhere: ``tests/golden/reference_findings.json`` (results only; made by
  ``tests/golden/make_reference_golden.py``) must contain the acceptance list
  SURVEY.md section 4 wrote down by reading the reference -- written out by hand below;
* where /root/reference is mounted: re-run the checks and compare with the
  golden file, and use what the reference itself offers as ground truth for the
  two restated third-party tools:
    - the 33 files its authors left ``terraform fmt``-clean must be fixed points
      of ``fmt`` here;
    - README table rows committed from real ``terraform-docs`` runs must be
      reproduced byte for byte, except rows the survey already knows are stale.
"""
import json
import os

import pytest

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "reference_findings.json")


@pytest.fixture(scope="module")
def golden():
    with open(GOLDEN) as fh:
        return json.load(fh)


def findings(golden, module):
    return {tuple(f) for f in golden["modules"][module]["findings"]}


# files `terraform fmt -check` is predicted to flag
FMT_EXPECTED = ["aks/main.tf", "aks/variables.tf", "eks/main.tf", "gke/examples/cnpack/main.tf", "gke/main.tf"]

# SURVEY.md sections 2 and 4: defects found by reading, as (module, rule, file, line)
ACCEPTANCE = [
    ("eks", "unused-variable", "eks/variables.tf", 8),          # aws_profile
    ("eks", "unused-variable", "eks/variables.tf", 14),         # region
    ("eks", "unused-variable", "eks/variables.tf", 144),        # cpu_node_pool_additional_user_data
    ("eks", "unused-variable", "eks/variables.tf", 188),        # additional_user_data
    ("eks", "unused-variable", "eks/variables.tf", 223),        # enable_dns_support
    ("eks", "unused-local", "eks/main.tf", 179),                # local.ami_id never referenced
    ("eks", "length-of-object", "eks/main.tf", 186),            # length(object) always > 0
    ("eks", "module-unpinned", "eks/main.tf", 145),             # iam module without version
    ("eks", "required-provider-missing", "eks/main.tf", 185),   # helm used, not declared
    ("eks", "required-provider-unused", "eks/versions.tf", 10),  # kubernetes declared, unused
    ("eks", "legacy-splat", "eks/output.tf", 5),
    ("eks", "helper-script", "eks/setup-kube-config.sh", 1),    # orphaned (CHANGELOG.md:11)
    ("eks", "helper-script", "eks/setup-kube-config.sh", 4),    # shebang after the licence comments
    ("eks", "helper-script", "eks/setup-kube-config.sh", 6),    # eval of jq output
    ("aks", "unused-variable", "aks/variables.tf", 54),         # cpu_os_sku
    ("aks", "unused-variable", "aks/variables.tf", 82),         # gpu_os_sku
    ("aks", "required-provider-unused", "aks/versions.tf", 10),
    ("gke", "unused-data-source", "gke/data.tf", 4),
    ("gke", "counterpart-mismatch", "gke/main.tf", 83),         # CPU pool tagged with gpu_instance_tags
    ("eks/examples/cnpack", "counterpart-mismatch", "eks/examples/cnpack/aws-fluentbit.tf", 24),
    ("gke", "required-provider-missing", "gke/main.tf", 195),   # helm
    ("gke", "required-provider-missing", "gke/main.tf", 156),   # kubernetes
    ("eks/examples/cnpack", "duplicate-resource-arguments", "eks/examples/cnpack/aws-fluentbit.tf", 22),
    ("eks/examples/cnpack", "guard-mismatch", "eks/examples/cnpack/aws-amp.tf", 111),
    ("eks/examples/cnpack", "docs-drift", "eks/examples/cnpack/Readme.md", 229),
    ("gke/examples/cnpack", "docs-unknown-input", "gke/examples/cnpack/README.md", 101),
    ("gke/examples/cnpack", "module-unpinned", "gke/examples/cnpack/gcp-prometheus.tf", 7),
    ("aks/examples/cnpack", "unused-variable", "aks/examples/cnpack/variables.tf", 12),
    ("aks/examples/cnpack", "unused-variable", "aks/examples/cnpack/variables.tf", 40),
    ("aks/examples/cnpack", "unused-variable", "aks/examples/cnpack/variables.tf", 64),
    ("aks/examples/cnpack", "required-provider-unused", "aks/examples/cnpack/providers.tf", 9),   # azuread
    ("aks/examples/cnpack", "secret-in-command", "aks/examples/cnpack/azure-fluentbit.tf", 29),
]

# block counts from SURVEY.md section 2 headings
COUNTS = {
    "eks": {"variables": 36, "outputs": 11, "module_calls": 3, "resources": 1, "data_sources": 5},
    "gke": {"variables": 25, "outputs": 10, "resources": 8, "data_sources": 4},
    "aks": {"variables": 18, "outputs": 5, "resources": 3, "data_sources": 2},
    "eks/examples/cnpack": {"variables": 7, "outputs": 4, "resources": 18, "data_sources": 3, "module_calls": 1},
}


def test_golden_fmt_matches_survey_prediction(golden):
    assert golden["fmt_changed"] == FMT_EXPECTED
    assert len(golden["fmt_checked"]) == 38


@pytest.mark.parametrize("module,rule,file,line", ACCEPTANCE)
def test_golden_contains_survey_acceptance_list(golden, module, rule, file, line):
    assert (rule, file, line) in findings(golden, module)


@pytest.mark.parametrize("module", sorted(COUNTS))
def test_golden_block_counts_match_survey(golden, module):
    got = golden["modules"][module]["counts"]
    assert {k: got[k] for k in COUNTS[module]} == COUNTS[module]


def test_golden_has_no_validate_errors_on_the_reference(golden):
    rules = {f[0] for m in golden["modules"].values() for f in m["findings"]}
    assert not rules & {"syntax", "structure", "undeclared-variable", "undeclared-resource", "instance-key",
                        "function", "module-arguments", "module-output"}


def test_golden_docs_rows_reproduce_terraform_docs_output(golden):
    rows = golden["docs_rows"]
    for module in ("eks", "gke", "aks", "aks/examples/cnpack"):
        assert rows[module]["identical_in_readme"] == rows[module]["generated"] > 0
    assert rows["eks/examples/cnpack"] == {"generated": 35, "identical_in_readme": 34}   # the one stale default
    total = sum(r["generated"] for r in rows.values())
    same = sum(r["identical_in_readme"] for r in rows.values())
    assert (total, same) == (226, 219)


# --------------------------------------------------------------- live checks
@pytest.mark.reference
def test_live_run_equals_golden(reference_root, golden):
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "make_reference_golden", os.path.join(os.path.dirname(GOLDEN), "make_reference_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    live = json.loads(json.dumps(mod.build(reference_root)))
    assert live == golden


@pytest.mark.reference
def test_files_the_authors_formatted_are_fixed_points(reference_root, golden):
    from nvidia_terraform_modules_b200.fmt.command import format_source
    clean = [p for p in golden["fmt_checked"] if p not in golden["fmt_changed"]]
    assert len(clean) == 33
    for rel in clean:
        with open(os.path.join(reference_root, rel), newline="") as fh:
            src = fh.read()
        assert format_source(src, rel)[0] == src, rel


@pytest.mark.reference
def test_formatting_the_rest_is_idempotent_and_white_space_only(reference_root, golden):
    from nvidia_terraform_modules_b200.fmt.command import format_source
    from nvidia_terraform_modules_b200.hcl.scanner import scan
    for rel in golden["fmt_changed"]:
        with open(os.path.join(reference_root, rel), newline="") as fh:
            src = fh.read()
        once = format_source(src, rel)[0]
        assert format_source(once, rel)[0] == once
        assert [(t.type, t.text) for t in scan(once)] == [(t.type, t.text) for t in scan(src)]


@pytest.mark.reference
def test_every_reference_hcl_file_round_trips_through_the_scanner(reference_root):
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
    assert seen == 44       # 32 .tf + 6 .tfvars + 6 .hcl


@pytest.mark.reference
@pytest.mark.parametrize("module", ["eks", "gke", "aks", "aks/examples/cnpack"])
def test_regenerating_an_up_to_date_readme_changes_nothing(reference_root, module):
    """The whole generated region -- headings, table headers, row order, blank
    lines -- equals what real terraform-docs wrote into these four READMEs."""
    from nvidia_terraform_modules_b200.config.lockfile import load_lockfile
    from nvidia_terraform_modules_b200.config.model import load_module
    from nvidia_terraform_modules_b200.docs.tfdocs import find_readme, inject, render
    d = os.path.join(reference_root, module)
    with open(find_readme(d), encoding="utf-8") as fh:
        text = fh.read()
    assert inject(text, render(load_module(d, reference_root), load_lockfile(d, reference_root))) == text
