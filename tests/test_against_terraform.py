"""The real tools, run beside the stand-ins.  DORMANT until a binary exists.

``fmt`` and ``validate`` here restate Terraform from its published behaviour; nothing
in this container can confirm them (no ``terraform``, no network: SURVEY.md section 8(c),
8(f) "Next 1").  This file is the confirmation, ready to run: the moment a Terraform
binary is on ``PATH`` (or named by ``$TF_BIN``),

    python -m pytest tests -m terraform

pins BASELINE.json configs 1 and 2 against it with no further code:

(a) ``terraform fmt -`` on every hand-written case of tests/test_fmt_canonical.py, every
    file of three seeded synthetic trees with planted defects, white-space-damaged copies
    of them, and -- where /root/reference is mounted -- every reference ``.tf``/``.tfvars``:
    BYTE EQUALITY with ``format_source``;
(b) ``terraform fmt -check -recursive`` (and ``-diff``) over a tree: same exit code,
    same file list, same bytes on stdout as ``fmt/command.py``;
(c) the two open questions answered explicitly: does ``fmt`` add a missing final
    newline (DESIGN.md section 2.2 infers that it does not), and the ``--- old/`` /
    ``+++ new/`` diff header;
(d) ``terraform init -backend=false && terraform validate -json`` on small modules that
    need no provider plugin (built-in ``terraform_data`` only), compared with ``validate``
    run as the SAME release (``--terraform-version`` taken from ``terraform version``):
    the multiset of ``(severity, summary, file, line)`` must agree.  With
    ``$TF_PLUGIN_MIRROR`` pointing at a directory of provider plugins
    (``terraform providers mirror`` layout) the cases that use ``null_resource`` /
    ``random_string`` run too;
(e) ``terraform-docs markdown table`` (``$TFDOCS_BIN`` or ``PATH``) against ``docs``.

What the driver has to supply: see DESIGN.md section 4 and BASELINE.md section 2.
Nothing here reads ``oracle/`` or needs a GPU; without the binaries every test skips
with the reason.
"""
from __future__ import annotations

import difflib
import json
import os
import random
import shutil
import subprocess

import pytest

from nvidia_terraform_modules_b200.config.validate import validate_module
from nvidia_terraform_modules_b200.fmt import command as fmt_command
from nvidia_terraform_modules_b200.fmt.command import discover, format_source
from nvidia_terraform_modules_b200.synth import write_tree

TERRAFORM = os.environ.get("TF_BIN") or shutil.which("terraform")
TFDOCS = os.environ.get("TFDOCS_BIN") or shutil.which("terraform-docs")
MIRROR = os.environ.get("TF_PLUGIN_MIRROR")
REFERENCE = os.environ.get("TF_REFERENCE_ROOT", "/root/reference")

NO_BINARY = ("no Terraform binary: put `terraform` (>= 1.3.4, /root/reference/aks/versions.tf:16) on PATH or set "
             "$TF_BIN; this harness then pins fmt/validate parity (DESIGN.md section 4)")

pytestmark = [pytest.mark.terraform, pytest.mark.skipif(not TERRAFORM, reason=NO_BINARY)]


# ----------------------------------------------------------------------- plumbing
def tf(args: list[str], cwd: str, stdin: str | None = None, timeout: int = 300) -> subprocess.CompletedProcess:
    env = dict(os.environ, TF_IN_AUTOMATION="1", CHECKPOINT_DISABLE="1", TF_INPUT="0",
               TF_CLI_ARGS="", NO_COLOR="1")
    env.pop("TF_WORKSPACE", None)
    return subprocess.run([TERRAFORM, *args], cwd=cwd, input=stdin, capture_output=True, text=True,
                          env=env, timeout=timeout)


def ours_fmt(argv: list[str], cwd: str) -> tuple[int, str, str]:
    import io
    out, err = io.StringIO(), io.StringIO()
    old = os.getcwd()
    os.chdir(cwd)
    try:
        code = fmt_command.main(argv, stdout=out, stderr=err, stdin=io.StringIO(""))
    finally:
        os.chdir(old)
    return code, out.getvalue(), err.getvalue()


def explain(want: str, got: str, name: str) -> str:
    return "".join(difflib.unified_diff(want.splitlines(True), got.splitlines(True),
                                        f"terraform/{name}", f"stand-in/{name}"))


@pytest.fixture(scope="session")
def tf_release(tmp_path_factory) -> str:
    p = tf(["version", "-json"], str(tmp_path_factory.mktemp("v")))
    assert p.returncode == 0, p.stderr
    full = json.loads(p.stdout)["terraform_version"]
    major, minor = full.split(".")[:2]
    return f"{major}.{minor}"


# ------------------------------------------------------------------- (a) fmt corpus
def _hand_written_cases() -> list[tuple[str, str]]:
    from test_fmt_canonical import CASES
    return [(f"case/{name}", src) for name, (src, _) in sorted(CASES.items())]


def _synthetic_files(tmp_root: str) -> list[tuple[str, str]]:
    from scramble import damage_lines, damage_tokens
    found = []
    for seed in (11, 12, 13):
        root = os.path.join(tmp_root, f"synth{seed}")
        os.makedirs(root)
        write_tree(root, seed=seed, modules=2, n_vars=20, n_resources=5, defects=8)
        for path in discover(root, True):
            with open(path, newline="") as fh:
                src = fh.read()
            rel = os.path.relpath(path, tmp_root)
            found.append((f"synth/{rel}", src))
            found.append((f"synth-lines/{rel}", damage_lines(src, random.Random(rel + "l"))))
            found.append((f"synth-tokens/{rel}", damage_tokens(src, random.Random(rel + "t"))))
    return found


def _reference_files() -> list[tuple[str, str]]:
    from scramble import damage_lines, damage_tokens
    found = []
    if not os.path.isdir(os.path.join(REFERENCE, "eks")):
        return found
    for top in ("aks", "eks", "gke"):
        for path in discover(os.path.join(REFERENCE, top), True):
            with open(path, newline="") as fh:
                src = fh.read()
            rel = os.path.relpath(path, REFERENCE)
            found.append((f"reference/{rel}", src))
            found.append((f"reference-lines/{rel}", damage_lines(src, random.Random(rel + "l"))))
            found.append((f"reference-tokens/{rel}", damage_tokens(src, random.Random(rel + "t"))))
    return found


@pytest.fixture(scope="session")
def fmt_corpus(tmp_path_factory) -> list[tuple[str, str]]:
    return (_hand_written_cases() + _synthetic_files(str(tmp_path_factory.mktemp("corpus")))
            + _reference_files())


def test_fmt_stdin_is_byte_identical_on_the_whole_corpus(fmt_corpus, tmp_path):
    """One test over the whole corpus so that a single run lists EVERY disagreement.
    (``$TF_HARNESS_FMT_SAMPLE=N`` thins it to about N inputs: the plumbing self-test uses
    that; a parity run must not.)"""
    assert len(fmt_corpus) > 100
    sample = int(os.environ.get("TF_HARNESS_FMT_SAMPLE", "0"))
    if sample:
        fmt_corpus = fmt_corpus[::max(1, len(fmt_corpus) // sample)]
    wrong = []
    for name, src in fmt_corpus:
        real = tf(["fmt", "-no-color", "-"], str(tmp_path), stdin=src)
        mine, diags = format_source(src, name)
        if real.returncode != 0:
            if not diags:
                wrong.append(f"{name}: terraform rejects the input, the stand-in accepts it\n{real.stderr}")
            continue
        if diags:
            wrong.append(f"{name}: the stand-in rejects input terraform formats: {diags[0].format()}")
        elif real.stdout != mine:
            wrong.append(explain(real.stdout, mine, name))
    assert not wrong, f"{len(wrong)} of {len(fmt_corpus)} inputs differ:\n" + "\n".join(wrong[:20])


# ---------------------------------------------------------- (b) the command itself
@pytest.fixture()
def defect_tree(tmp_path) -> str:
    root = str(tmp_path / "tree")
    os.makedirs(root)
    write_tree(root, seed=21, modules=2, n_vars=12, n_resources=4, defects=8)
    os.makedirs(os.path.join(root, ".hidden"))
    with open(os.path.join(root, ".hidden", "skipped.tf"), "w") as fh:
        fh.write("a   = 1\n")
    return root


@pytest.mark.parametrize("args", [
    ["-check", "-recursive"],
    ["-check", "-recursive", "-diff"],
    ["-check", "-recursive", "-list=false"],
    ["-check"],
    ["-write=false", "-recursive"],
    ["-write=false", "-list=false", "-diff", "-recursive"],
], ids=lambda a: " ".join(a))
def test_fmt_command_same_exit_code_and_stdout(defect_tree, args):
    real = tf(["fmt", "-no-color", *args], defect_tree)
    code, out, _ = ours_fmt(args, defect_tree)
    assert (real.returncode, real.stdout) == (code, out), explain(real.stdout, out, "stdout")


def test_fmt_write_leaves_the_same_bytes_on_disk(defect_tree, tmp_path):
    twin = str(tmp_path / "twin")
    shutil.copytree(defect_tree, twin)
    real = tf(["fmt", "-no-color", "-recursive"], defect_tree)
    code, out, _ = ours_fmt(["-recursive"], twin)
    assert (real.returncode, real.stdout) == (code, out)
    for path in discover(defect_tree, True):
        rel = os.path.relpath(path, defect_tree)
        with open(path, "rb") as a, open(os.path.join(twin, rel), "rb") as b:
            assert a.read() == b.read(), rel


@pytest.mark.reference
def test_fmt_check_recursive_on_the_reference_is_baseline_config_1():
    """BASELINE.json config 1, verbatim: ``terraform fmt -check -recursive`` over eks/, aks/,
    gke/.  -check never writes, so the read-only mount is safe.  SURVEY.md section 4 predicts
    exit 3 and five files (tests/golden/survey_expectations.json)."""
    if not os.path.isdir(os.path.join(REFERENCE, "eks")):
        pytest.skip("reference tree not present on this machine")
    args = ["-check", "-recursive", "eks", "aks", "gke"]
    real = tf(["fmt", "-no-color", *args], REFERENCE)
    code, out, _ = ours_fmt(args, REFERENCE)
    assert (real.returncode, real.stdout) == (code, out), explain(real.stdout, out, "stdout")
    with open(os.path.join(os.path.dirname(__file__), "golden", "survey_expectations.json")) as fh:
        predicted = json.load(fh)["fmt_files_predicted_to_differ"]["files"]
    assert real.returncode == 3 and sorted(real.stdout.split()) == predicted


def test_fmt_syntax_error_exit_code(tmp_path):
    (tmp_path / "broken.tf").write_text("a = (\n")
    (tmp_path / "fine.tf").write_text("a   = 1\n")
    real = tf(["fmt", "-no-color", "-check"], str(tmp_path))
    code, out, _ = ours_fmt(["-check"], str(tmp_path))
    assert (real.returncode, real.stdout) == (code, out)


# -------------------------------------------------- (c) the two open questions
def test_open_question_missing_final_newline(tmp_path):
    """DESIGN.md section 2.2 infers, from nine reference files that lack a final newline
    yet are otherwise canonical, that ``terraform fmt`` leaves the end of file alone.
    SURVEY.md section 4 calls this its least certain prediction.  This settles it."""
    for src in ('a = 1', 'a   = 1', 'b "x" {\n  a = 1\n}', '# only a comment'):
        real = tf(["fmt", "-no-color", "-"], str(tmp_path), stdin=src)
        assert real.returncode == 0, real.stderr
        assert real.stdout == format_source(src)[0], (src, real.stdout)
    (tmp_path / "noeol.tf").write_text("a = 1")
    real = tf(["fmt", "-no-color", "-check"], str(tmp_path))
    code, out, _ = ours_fmt(["-check"], str(tmp_path))
    assert (real.returncode, real.stdout) == (code, out)


def test_open_question_diff_header_and_hunks(tmp_path):
    (tmp_path / "main.tf").write_text('a   = 1\nbb = 2\n\n\n\n\n\n\n\n\nc "d" {\n    e = 1\n}')
    real = tf(["fmt", "-no-color", "-write=false", "-list=false", "-diff"], str(tmp_path))
    code, out, _ = ours_fmt(["-write=false", "-list=false", "-diff"], str(tmp_path))
    assert real.stdout.startswith("--- old/main.tf\n+++ new/main.tf\n")
    assert (real.returncode, real.stdout) == (code, out), explain(real.stdout, out, "diff")
    # and the in-process fallback used where no `diff` program exists
    src = (tmp_path / "main.tf").read_text()
    assert fmt_command.unified_diff("main.tf", src, format_source(src)[0], external=False) == real.stdout


# ------------------------------------------------------------------ (d) validate
CHILD = {
    "child/main.tf": 'variable "label" {\n  type = string\n}\nvariable "size" {\n  type    = number\n  default = 1\n}\n'
                     'output "id" {\n  value = "${var.label}-${var.size}"\n}\n',
}

# name -> files.  One defect per case: Terraform stops after configuration-loading
# errors, so a structural error and a reference error never appear in one run.
VALIDATE_CASES: dict[str, dict[str, str]] = {
    "clean": {"main.tf": 'variable "a" {\n  type    = string\n  default = "x"\n}\nlocals {\n  b = upper(var.a)\n}\n'
                         'output "o" {\n  value = local.b\n}\nmodule "c" {\n  source = "./child"\n  label  = var.a\n}\n'
                         'output "p" {\n  value = module.c.id\n}\n', **CHILD},
    "undeclared_variable": {"main.tf": 'output "o" {\n  value = var.nope\n}\n'},
    "undeclared_local": {"main.tf": 'output "o" {\n  value = local.nope\n}\n'},
    "undeclared_module": {"main.tf": 'output "o" {\n  value = module.nope.id\n}\n'},
    "undeclared_resource": {"main.tf": 'output "o" {\n  value = terraform_data.nope.id\n}\n'},
    "undeclared_data": {"main.tf": 'output "o" {\n  value = data.terraform_remote_state.nope.outputs\n}\n'},
    "var_object_itself": {"main.tf": 'output "o" {\n  value = var\n}\n'},
    "count_outside": {"main.tf": 'output "o" {\n  value = count.index\n}\n'},
    "each_outside": {"main.tf": 'locals {\n  a = each.key\n}\n'},
    "self_outside": {"main.tf": 'locals {\n  a = self.id\n}\n'},
    "path_attr": {"main.tf": 'locals {\n  a = path.nope\n}\n'},
    "terraform_attr": {"main.tf": 'locals {\n  a = terraform.nope\n}\n'},
    "unknown_function": {"main.tf": 'locals {\n  a = frobnicate(1)\n}\n'},
    "too_few_arguments": {"main.tf": 'locals {\n  a = substr("abc", 1)\n}\n'},
    "too_many_arguments": {"main.tf": 'locals {\n  a = upper("a", "b")\n}\n'},
    "removed_function_list": {"main.tf": 'locals {\n  a = list("a")\n}\n'},
    "self_referencing_local": {"main.tf": 'locals {\n  a = local.a\n}\n'},
    "local_cycle": {"main.tf": 'locals {\n  a = local.b\n  b = local.a\n}\n'},
    "duplicate_variable": {"main.tf": 'variable "a" {}\nvariable "a" {}\n'},
    "duplicate_output": {"main.tf": 'output "a" {\n  value = 1\n}\noutput "a" {\n  value = 2\n}\n'},
    "duplicate_local": {"main.tf": 'locals {\n  a = 1\n}\nlocals {\n  a = 2\n}\n'},
    "duplicate_module": {"main.tf": 'module "c" {\n  source = "./child"\n  label  = "a"\n}\nmodule "c" {\n'
                                    '  source = "./child"\n  label  = "b"\n}\n', **CHILD},
    "output_without_value": {"main.tf": 'output "a" {\n  description = "d"\n}\n'},
    "variable_unsupported_argument": {"main.tf": 'variable "a" {\n  colour = "red"\n}\n'},
    "variable_bad_type_keyword": {"main.tf": 'variable "a" {\n  type = strin\n}\n'},
    "variable_quoted_type": {"main.tf": 'variable "a" {\n  type = "string"\n}\n'},
    "variable_default_wrong_type": {"main.tf": 'variable "a" {\n  type    = number\n  default = "abc"\n}\n'},
    "variable_default_refers": {"main.tf": 'variable "a" {\n  default = var.b\n}\nvariable "b" {}\n'},
    "variable_reserved_name": {"main.tf": 'variable "count" {}\n'},
    "variable_bad_name": {"main.tf": 'variable "1a" {}\n'},
    "validation_without_own_variable": {"main.tf": 'variable "a" {\n  validation {\n    condition     = true\n'
                                                   '    error_message = "No."\n  }\n}\n'},
    "validation_missing_message": {"main.tf": 'variable "a" {\n  validation {\n    condition = var.a != ""\n  }\n}\n'},
    "validation_cross_variable": {"main.tf": 'variable "a" {}\nvariable "b" {\n  validation {\n'
                                             '    condition     = var.b != var.a\n    error_message = "Differ."\n  }\n}\n'},
    "unknown_top_level_block": {"main.tf": 'resorce "a" "b" {}\n'},
    "top_level_argument": {"main.tf": 'a = 1\n'},
    "missing_label": {"main.tf": 'variable {}\n'},
    "extra_label": {"main.tf": 'output "a" "b" {\n  value = 1\n}\n'},
    "module_without_source": {"main.tf": 'module "c" {\n  label = "a"\n}\n'},
    "module_unsupported_argument": {"main.tf": 'module "c" {\n  source = "./child"\n  label  = "a"\n  colour = 1\n}\n',
                                    **CHILD},
    "module_missing_argument": {"main.tf": 'module "c" {\n  source = "./child"\n}\n', **CHILD},
    "module_unknown_output": {"main.tf": 'module "c" {\n  source = "./child"\n  label  = "a"\n}\n'
                                         'output "o" {\n  value = module.c.nope\n}\n', **CHILD},
    "module_instance_key_without_count": {"main.tf": 'module "c" {\n  source = "./child"\n  label  = "a"\n}\n'
                                                     'output "o" {\n  value = module.c[0].id\n}\n', **CHILD},
    "module_missing_instance_key": {"main.tf": 'module "c" {\n  source = "./child"\n  count  = 2\n  label  = "a"\n}\n'
                                               'output "o" {\n  value = module.c.id\n}\n', **CHILD},
    "module_count_and_for_each": {"main.tf": 'module "c" {\n  source   = "./child"\n  count    = 1\n'
                                             '  for_each = toset(["a"])\n  label    = "a"\n}\n', **CHILD},
    "module_version_on_local_source": {"main.tf": 'module "c" {\n  source  = "./child"\n  version = "1.0.0"\n'
                                                  '  label   = "a"\n}\n', **CHILD},
    "required_version_syntax": {"main.tf": 'terraform {\n  required_version = "~~> 1"\n}\n'},
    "required_version_excludes_this_binary": {"main.tf": 'terraform {\n  required_version = "< 0.12.0"\n}\n'},
    "terraform_unsupported_argument": {"main.tf": 'terraform {\n  colour = "red"\n}\n'},
    "inconsistent_conditional": {"main.tf": 'locals {\n  a = true ? "x" : [1]\n}\n'},
    "deprecated_interpolation_only": {"main.tf": 'variable "a" {\n  default = "x"\n}\noutput "o" {\n  value = "${var.a}"\n}\n'},
    "depends_on_expression": {"main.tf": 'variable "a" {\n  default = "x"\n}\noutput "o" {\n  value      = 1\n'
                                         '  depends_on = [upper(var.a)]\n}\n'},
    "moved_block": {"main.tf": 'moved {\n  from = terraform_data.a\n  to   = terraform_data.b\n}\n'
                               'resource "terraform_data" "b" {}\n'},
    "check_block": {"main.tf": 'variable "a" {\n  default = 1\n}\ncheck "c" {\n  assert {\n    condition     = var.a > 0\n'
                               '    error_message = "Positive."\n  }\n}\n'},
    "check_without_assert": {"main.tf": 'check "c" {\n}\n'},
    "optional_attribute_default": {"main.tf": 'variable "o" {\n  type = object({\n    a = optional(string, "x")\n  })\n'
                                              '  default = {}\n}\n'},
    # the built-in provider (Terraform >= 1.4) gives real resources without a plugin
    "resource_count_index": {"main.tf": 'resource "terraform_data" "a" {\n  count = 2\n  input = count.index\n}\n'
                                        'output "o" {\n  value = terraform_data.a[0].output\n}\n'},
    "resource_missing_instance_key": {"main.tf": 'resource "terraform_data" "a" {\n  count = 2\n}\n'
                                                 'output "o" {\n  value = terraform_data.a.output\n}\n'},
    "resource_unexpected_instance_key": {"main.tf": 'resource "terraform_data" "a" {}\n'
                                                    'output "o" {\n  value = terraform_data.a[0].output\n}\n'},
    "resource_self_reference": {"main.tf": 'resource "terraform_data" "a" {\n  input = terraform_data.a.id\n}\n'},
    "resource_cycle": {"main.tf": 'resource "terraform_data" "a" {\n  input = terraform_data.b.id\n}\n'
                                  'resource "terraform_data" "b" {\n  input = terraform_data.a.id\n}\n'},
    "resource_count_and_for_each": {"main.tf": 'resource "terraform_data" "a" {\n  count    = 1\n'
                                               '  for_each = toset(["x"])\n}\n'},
    "resource_for_each_tuple": {"main.tf": 'resource "terraform_data" "a" {\n  for_each = ["x", "y"]\n}\n'},
    "resource_each_value": {"main.tf": 'resource "terraform_data" "a" {\n  for_each = toset(["x"])\n  input    = each.value\n}\n'},
    "resource_duplicate": {"main.tf": 'resource "terraform_data" "a" {}\nresource "terraform_data" "a" {}\n'},
    "resource_ignore_changes_keyword": {"main.tf": 'resource "terraform_data" "a" {\n  lifecycle {\n'
                                                   '    ignore_changes = "all"\n  }\n}\n'},
    "resource_lifecycle_unknown_argument": {"main.tf": 'resource "terraform_data" "a" {\n  lifecycle {\n'
                                                       '    colour = "red"\n  }\n}\n'},
    "resource_precondition": {"main.tf": 'variable "a" {\n  default = 1\n}\nresource "terraform_data" "a" {\n  lifecycle {\n'
                                         '    precondition {\n      condition     = var.a > 0\n'
                                         '      error_message = "Positive."\n    }\n  }\n}\n'},
    "resource_postcondition_self": {"main.tf": 'resource "terraform_data" "a" {\n  lifecycle {\n    postcondition {\n'
                                               '      condition     = self.id != ""\n      error_message = "Id."\n'
                                               '    }\n  }\n}\n'},
    "provisioner_keywords": {"main.tf": 'resource "terraform_data" "a" {\n  provisioner "local-exec" {\n'
                                        '    when       = destroy\n    on_failure = continue\n'
                                        '    command    = "echo ${self.id}"\n  }\n}\n'},
    "provisioner_bad_when": {"main.tf": 'resource "terraform_data" "a" {\n  provisioner "local-exec" {\n'
                                        '    when    = later\n    command = "x"\n  }\n}\n'},
    "provisioner_destroy_reference": {"main.tf": 'variable "h" {\n  default = "x"\n}\nresource "terraform_data" "a" {\n'
                                                 '  provisioner "local-exec" {\n    when    = destroy\n'
                                                 '    command = "echo ${var.h}"\n  }\n}\n'},
    "depends_on_quoted": {"main.tf": 'resource "terraform_data" "a" {}\nresource "terraform_data" "b" {\n'
                                     '  depends_on = ["terraform_data.a"]\n}\n'},
    "nullable_false_null_default": {"main.tf": 'variable "a" {\n  type     = string\n  nullable = false\n  default  = null\n}\n'},
    "sensitive_not_bool": {"main.tf": 'variable "a" {\n  sensitive = "maybe"\n}\n'},
    "locals_nested_block": {"main.tf": 'locals {\n  a = 1\n  nested {\n  }\n}\n'},
    "count_negative": {"main.tf": 'resource "terraform_data" "a" {\n  count = -1\n}\n'},
    "count_null": {"main.tf": 'resource "terraform_data" "a" {\n  count = null\n}\n'},
    "count_string": {"main.tf": 'resource "terraform_data" "a" {\n  count = "many"\n}\n'},
    "for_each_string": {"main.tf": 'resource "terraform_data" "a" {\n  for_each = "x"\n}\n'},
    "count_wrong_attr": {"main.tf": 'resource "terraform_data" "a" {\n  count = 1\n  input = count.number\n}\n'},
    "each_wrong_attr": {"main.tf": 'resource "terraform_data" "a" {\n  for_each = toset(["x"])\n  input    = each.index\n}\n'},
    "depends_on_not_a_list": {"main.tf": 'resource "terraform_data" "a" {}\nresource "terraform_data" "b" {\n'
                                         '  depends_on = terraform_data.a\n}\n'},
    "module_lifecycle_block": {"main.tf": 'module "c" {\n  source = "./child"\n  label  = "a"\n  lifecycle {\n  }\n}\n',
                               **CHILD},
    "moved_missing_to": {"main.tf": 'moved {\n  from = terraform_data.a\n}\n'},
    "import_missing_id": {"main.tf": 'import {\n  to = terraform_data.a\n}\nresource "terraform_data" "a" {}\n'},
    "check_bad_name": {"main.tf": 'check "1x" {\n  assert {\n    condition     = true\n    error_message = "m"\n  }\n}\n'},
    "duplicate_required_providers_entry": {"main.tf": 'terraform {\n  required_providers {\n    null = {\n'
                                                      '      source = "hashicorp/null"\n    }\n  }\n'
                                                      '  required_providers {\n    null = {\n'
                                                      '      source = "hashicorp/null"\n    }\n  }\n}\n'},
    "required_providers_bad_key": {"main.tf": 'terraform {\n  required_providers {\n    null = {\n'
                                              '      source = "hashicorp/null"\n      colour = "red"\n    }\n  }\n}\n'},
    "output_sensitive_not_bool": {"main.tf": 'output "o" {\n  value     = 1\n  sensitive = "yes"\n}\n'},
    "attribute_redefined": {"main.tf": 'output "o" {\n  value = 1\n  value = 2\n}\n'},
    "legacy_index_warning_free": {"main.tf": 'locals {\n  l = [1, 2]\n  a = local.l.0\n}\n'},
    "for_expression_scopes": {"main.tf": 'locals {\n  l = ["a"]\n  a = [for i, v in local.l : "${i}${v}" if v != ""]\n'
                                         '  b = { for v in local.l : v => i }\n}\n'},
    "template_directive_scopes": {"main.tf": 'locals {\n  l = ["a"]\n  a = "%{for v in local.l}${v}%{endfor}${v}"\n}\n'},
    "provisioner_on_failure_bad": {"main.tf": 'resource "terraform_data" "a" {\n  provisioner "local-exec" {\n'
                                              '    on_failure = retry\n    command    = "x"\n  }\n}\n'},
    "provisioner_quoted_keyword": {"main.tf": 'resource "terraform_data" "a" {\n  provisioner "local-exec" {\n'
                                              '    when    = "destroy"\n    command = "x"\n  }\n}\n'},
    "self_in_precondition": {"main.tf": 'resource "terraform_data" "a" {\n  lifecycle {\n    precondition {\n'
                                        '      condition     = self.id != ""\n      error_message = "m"\n    }\n  }\n}\n'},
    "quoted_type_keyword": {"main.tf": 'variable "a" {\n  type = "string"\n}\n'},
    "legacy_type_hint": {"main.tf": 'variable "a" {\n  type = "list(string)"\n}\n'},
    "optional_outside_object": {"main.tf": 'variable "a" {\n  type = optional(string)\n}\n'},
    "type_constructor_arity": {"main.tf": 'variable "a" {\n  type = list(string, number)\n}\n'},
    "scoped_data_source": {"main.tf": 'check "c" {\n  data "terraform_remote_state" "s" {\n    backend = "local"\n  }\n'
                                      '  assert {\n    condition     = data.terraform_remote_state.s.outputs != null\n'
                                      '    error_message = "State."\n  }\n}\n'
                                      'output "o" {\n  value = data.terraform_remote_state.s.outputs\n}\n'},
}

# Cases where this checker's wording or placement is a guess that no published text
# settles; they are compared all the same, and reported as xfail/xpass rather than
# failing the run, so that the first run against a real binary says which way each went.
UNCERTAIN = {
    "module_instance_key_without_count": "Terraform may report HCL's own \"Invalid index\" on the module object",
    "module_missing_instance_key": "Terraform may report HCL's \"Unsupported attribute\" on the list of instances",
    "resource_ignore_changes_keyword": "wording of a non-list, non-keyword ignore_changes",
    "removed_function_list": "summary of the error a removed function raises",
    "scoped_data_source": "whether validate or only plan reports a reference to a scoped data source",
    "count_null": "wording for a null count",
    "count_string": "wording for a count that is not a number",
    "for_each_string": "wording for a for_each that is a string literal",
    "count_negative": "whether validate or only plan rejects a negative literal count",
    "sensitive_not_bool": "wording for a non-boolean sensitive argument",
    "output_sensitive_not_bool": "wording for a non-boolean sensitive argument",
    "legacy_index_warning_free": "whether `.0` still parses silently",
    "template_directive_scopes": "placement of the error for a name used after its %{for} has ended",
    "module_lifecycle_block": "wording for a reserved block type inside a module call",
    "check_bad_name": "wording for an invalid check block name",
    "optional_outside_object": "wording for optional() outside an object type",
    "type_constructor_arity": "wording for a type constructor with too many arguments",
    "duplicate_required_providers_entry": "two required_providers blocks in one terraform block",
}

# with $TF_PLUGIN_MIRROR: providers the reference's example lock files pin
# (random 3.5.1, null 3.2.1 -- eks/examples/cnpack/.terraform.lock.hcl)
MIRROR_CASES: dict[str, dict[str, str]] = {
    "null_resource_triggers": {"main.tf": 'resource "null_resource" "n" {\n  count    = 2\n  triggers = {\n'
                                          '    i = count.index\n  }\n}\noutput "o" {\n  value = null_resource.n[*].id\n}\n'},
    "random_string_missing_key": {"main.tf": 'resource "random_string" "r" {\n  count  = 1\n  length = 4\n}\n'
                                             'output "o" {\n  value = random_string.r.result\n}\n'},
}


# `terraform init` refuses a configuration that does not load, so the child module of such
# a case is never installed and validate adds this on top of the error under test
INIT_ARTEFACTS = {"Module not installed"}


def _write(root: str, files: dict[str, str]) -> None:
    for rel, text in files.items():
        path = os.path.join(root, rel)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "w", newline="") as fh:
            fh.write(text)


def _real_validate(root: str, plugin_dir: str | None = None) -> list[tuple]:
    init = ["init", "-backend=false", "-input=false", "-no-color"]
    if plugin_dir:
        init.append(f"-plugin-dir={plugin_dir}")
    tf(init, root)           # its errors (if any) come back from validate as well
    p = tf(["validate", "-json", "-no-color"], root)
    data = json.loads(p.stdout)
    out = []
    for d in data["diagnostics"]:
        if d["summary"] in INIT_ARTEFACTS:
            continue
        rng = d.get("range") or {}
        out.append((d["severity"], d["summary"], rng.get("filename", ""), (rng.get("start") or {}).get("line", 0)))
    return sorted(_normal(d) for d in out)


def _normal(d: tuple) -> tuple:
    """Cycle errors come from Terraform's graph, which names its nodes ("local.a (expand)"),
    lists them in walk order and attaches no source range: compare the members only."""
    severity, summary, file, line = d
    if summary.startswith("Cycle: "):
        members = sorted(m.strip().split(" ")[0] for m in summary[len("Cycle: "):].split(","))
        return severity, "Cycle: " + ", ".join(members), "", 0
    return d


def _our_validate(root: str, release: str) -> list[tuple]:
    res = validate_module(root, root, release)
    return sorted(_normal((d.severity, d.summary, d.file, d.line)) for d in res.diags)


@pytest.mark.parametrize("name", [
    pytest.param(n, marks=pytest.mark.xfail(reason=UNCERTAIN[n], strict=False)) if n in UNCERTAIN else n
    for n in sorted(VALIDATE_CASES)])
def test_validate_diagnostics_agree(name, tmp_path, tf_release):
    root = str(tmp_path)
    _write(root, VALIDATE_CASES[name])
    assert _our_validate(root, tf_release) == _real_validate(root)


@pytest.mark.skipif(not MIRROR, reason="$TF_PLUGIN_MIRROR not set: provider-backed validate cases need plugin binaries "
                                       "(hashicorp/null 3.2.1, hashicorp/random 3.5.1)")
@pytest.mark.parametrize("name", sorted(MIRROR_CASES))
def test_validate_diagnostics_agree_with_providers(name, tmp_path, tf_release):
    root = str(tmp_path)
    _write(root, MIRROR_CASES[name])
    real = _real_validate(root, MIRROR)
    if any("provider" in s.lower() and ("install" in s.lower() or "query" in s.lower()) for _, s, _, _ in real):
        pytest.skip(f"the mirror at {MIRROR} does not hold the providers this case needs: {real}")
    assert _our_validate(root, tf_release) == real


def test_validate_synthetic_trees_are_clean_for_both(tmp_path, tf_release):
    """The synthetic modules use invented providers, which no mirror can supply; what both
    tools can still be asked is whether the provider-free example modules load."""
    root = str(tmp_path)
    _write(root, VALIDATE_CASES["clean"])
    assert _real_validate(root) == [] == _our_validate(root, tf_release)


# --------------------------------------------------------------- (e) terraform-docs
@pytest.mark.skipif(not TFDOCS, reason="no terraform-docs binary: put it on PATH or set $TFDOCS_BIN "
                                       "(/root/reference/CONTRIBUTING.md:14)")
def test_docs_tables_are_byte_identical(tmp_path):
    from nvidia_terraform_modules_b200.config.lockfile import load_lockfile
    from nvidia_terraform_modules_b200.config.model import load_module
    from nvidia_terraform_modules_b200.docs.tfdocs import render
    root = str(tmp_path)
    tree = write_tree(root, seed=31, modules=2, n_vars=15, n_resources=4)
    dirs = [os.path.join(root, m) for m in tree.modules]
    if os.path.isdir(os.path.join(REFERENCE, "eks")):
        dirs += [os.path.join(REFERENCE, m) for m in ("eks", "gke", "aks", "eks/examples/cnpack",
                                                      "gke/examples/cnpack", "aks/examples/cnpack")]
    for d in dirs:
        p = subprocess.run([TFDOCS, "markdown", "table", d], capture_output=True, text=True, timeout=120)
        assert p.returncode == 0, p.stderr
        real = p.stdout[p.stdout.index("## Requirements"):].rstrip("\n") + "\n"
        mine = render(load_module(d, d), load_lockfile(d, d)).rstrip("\n") + "\n"
        assert real == mine, explain(real, mine, d)
