"""One Terraform release per run: which release is chosen, and one test per feature
gate (config/tfversion.py).  Each gate is tried one release before and at the release
that introduced the feature; before it, the message must start with the words that
release of Terraform uses for something it has never heard of."""
import pytest

from nvidia_terraform_modules_b200.config import tfversion
from nvidia_terraform_modules_b200.config.validate import validate_module
from nvidia_terraform_modules_b200.config.versions import Constraints
from nvidia_terraform_modules_b200.hcl.diagnostics import ERROR, WARNING


def run(module_dir, text, version=None, **more):
    files = {"main.tf": text}
    files.update(more)
    d = module_dir(files)
    return validate_module(d, d, version)


def errors(res):
    return [(d.summary, d.line) for d in res.diags if d.severity == ERROR]


# ----------------------------------------------------------------- choosing the release
@pytest.mark.parametrize("constraint,want", [
    (">= 1.2.4", (1, 2)),           # eks/versions.tf:16
    (">= 0.14", (0, 14)),           # gke/versions.tf:15
    (">= 1.3.4", (1, 3)),           # aks/versions.tf:16
    ("~> 1.5", (1, 5)),
    ("~> 1.5.2", (1, 5)),
    ("= 1.7.3", (1, 7)),
    ("1.7.3", (1, 7)),
    ("> 1.4.9, < 2.0", (1, 4)),     # 1.4.10 is the first release allowed
    (">= 1.0, != 1.0.0", (1, 0)),
    (">= 1.3, < 1.2", None),        # contradictory
    ("< 2.0", None),                # no floor
    ("!= 1.4.0", None),
])
def test_lowest_release_a_constraint_admits(constraint, want):
    assert tfversion.lowest_admitted([Constraints.parse(constraint)]) == want


def test_several_required_version_constraints_combine():
    cons = [Constraints.parse(">= 1.1"), Constraints.parse(">= 1.4.2")]
    assert tfversion.lowest_admitted(cons) == (1, 4)


def test_release_comes_from_flag_then_required_version_then_default(module_dir):
    text = 'terraform {\n  required_version = ">= 1.2.4"\n}\n'
    r = run(module_dir, text).module.restated
    assert r.release == (1, 2) and 'required_version ">= 1.2.4" at main.tf:2' in r.source
    r = run(module_dir, text, "1.6.2").module.restated
    assert r.release == (1, 6) and r.source == "--terraform-version"
    r = run(module_dir, 'variable "a" {}\n').module.restated
    assert r.release == tfversion.NEWEST and "default" in r.source
    r = run(module_dir, 'terraform {\n  required_version = "< 2.0"\n}\n').module.restated
    assert r.release == tfversion.NEWEST and "sets no floor" in r.source


def test_a_named_release_the_module_excludes_is_refused_like_terraform_would(module_dir):
    text = 'terraform {\n  required_version = ">= 1.2.4, < 1.6"\n}\n'
    for asked, ok in (("1.1", False), ("1.2", True), ("1.5.7", True), ("1.6", False), ("1.14", False)):
        res = run(module_dir, text, asked)
        assert (errors(res) == []) is ok, asked
        if not ok:
            assert errors(res) == [("Unsupported Terraform Core version", 2)]
            want = ".".join(asked.split(".")[:2])
            assert res.diags[0].detail.startswith(f"This configuration does not support Terraform version {want}.")
    assert errors(run(module_dir, text)) == []          # its own floor is always acceptable


def test_releases_outside_the_restated_range_are_clamped_and_say_so(module_dir):
    r = run(module_dir, 'terraform {\n  required_version = ">= 0.12"\n}\n').module.restated
    assert r.release == tfversion.OLDEST and "read as 0.13" in r.describe()
    r = run(module_dir, 'terraform {\n  required_version = ">= 1.12"\n}\n').module.restated
    assert r.release == tfversion.NEWEST and "nothing newer than 1.10 is restated" in r.describe()
    with pytest.raises(tfversion.TerraformVersionError):
        tfversion.from_flag("newest")


def test_a_child_module_is_read_under_its_callers_release(module_dir):
    files = {
        "child/main.tf": 'terraform {\n  required_version = ">= 1.9"\n}\nvariable "a" {\n  type     = string\n'
                         '  nullable = false\n}\n',
    }
    root = 'terraform {\n  required_version = ">= 1.0"\n}\nmodule "c" {\n  source = "./child"\n  a      = "x"\n}\n'
    res = run(module_dir, root, **files)
    child = res.validator.children["c"]
    assert child.restated.release == (1, 0)
    assert [(d.summary, d.file.replace("\\", "/").split("/")[-2:]) for d in child.diags] == [
        ("Unsupported argument", ["child", "main.tf"])]


# ------------------------------------------------------------------------ block gates
BLOCKS = [
    ("block:moved", 'moved {\n  from = a_b.c\n  to   = a_b.d\n}\nresource "a_b" "d" {}\n'),
    ("block:import", 'import {\n  to = a_b.d\n  id = "x"\n}\nresource "a_b" "d" {}\n'),
    ("block:check", 'check "c" {\n  assert {\n    condition     = true\n    error_message = "m"\n  }\n}\n'),
    ("block:removed", 'removed {\n  from = a_b.c\n  lifecycle {\n    destroy = false\n  }\n}\n'),
    ("block:ephemeral", 'ephemeral "a_b" "c" {}\n'),
]


def before(feature):
    """The release just before the one that introduced ``feature``."""
    since = tfversion.SINCE[feature]
    return tfversion.text(tfversion.SERIES[tfversion.SERIES.index(since) - 1])


def at(feature):
    return tfversion.text(tfversion.SINCE[feature])


@pytest.mark.parametrize("feature,text", BLOCKS, ids=[b[0] for b in BLOCKS])
def test_top_level_block_gate(module_dir, feature, text):
    kind = feature.split(":")[1]
    old = run(module_dir, text, before(feature))
    assert errors(old) == [("Unsupported block type", 1)]
    d = old.diags[0]
    assert d.detail.startswith(f'Blocks of type "{kind}" are not expected here.')
    assert f"needs Terraform {at(feature)} or later" in d.detail and d.rule == "terraform-version"
    assert errors(run(module_dir, text, at(feature))) == []


# --------------------------------------------------------------------- argument gates
ARGS = [
    ("variable:sensitive", 'variable "a" {\n  sensitive = true\n}\n', 2, "Unsupported argument"),
    ("variable:nullable", 'variable "a" {\n  nullable = false\n}\n', 2, "Unsupported argument"),
    ("variable:ephemeral", 'variable "a" {\n  ephemeral = true\n}\n', 2, "Unsupported argument"),
    ("output:ephemeral", 'output "a" {\n  value     = 1\n  ephemeral = true\n}\n', 3, "Unsupported argument"),
    ("output:precondition", 'output "a" {\n  value = 1\n  precondition {\n    condition     = true\n'
                            '    error_message = "m"\n  }\n}\n', 3, "Unsupported block type"),
    ("lifecycle:precondition", 'resource "a_b" "c" {\n  lifecycle {\n    precondition {\n      condition     = true\n'
                               '      error_message = "m"\n    }\n  }\n}\n', 3, "Unsupported block type"),
    ("lifecycle:postcondition", 'resource "a_b" "c" {\n  lifecycle {\n    postcondition {\n'
                                '      condition     = self.id != ""\n      error_message = "m"\n    }\n  }\n}\n',
     3, "Unsupported block type"),
    ("lifecycle:replace_triggered_by", 'resource "a_b" "d" {}\nresource "a_b" "c" {\n  lifecycle {\n'
                                       '    replace_triggered_by = [a_b.d]\n  }\n}\n', 4, "Unsupported argument"),
    ("terraform:cloud", 'terraform {\n  cloud {\n    organization = "o"\n  }\n}\n', 2, "Unsupported block type"),
    ("import:for_each", 'import {\n  for_each = toset(["x"])\n  to       = a_b.d[each.key]\n  id       = each.key\n}\n'
                        'resource "a_b" "d" {\n  for_each = toset(["x"])\n}\n', 2, "Unsupported argument"),
    ("removed:provisioner", 'removed {\n  from = a_b.c\n  lifecycle {\n    destroy = true\n  }\n'
                            '  provisioner "local-exec" {\n    when    = destroy\n    command = "x"\n  }\n}\n',
     6, "Unsupported block type"),
]


@pytest.mark.parametrize("feature,text,line,summary", ARGS, ids=[a[0] for a in ARGS])
def test_argument_and_nested_block_gate(module_dir, feature, text, line, summary):
    old = run(module_dir, text, before(feature))
    assert (summary, line) in errors(old)
    d = next(d for d in old.diags if d.line == line and d.severity == ERROR)
    assert f"needs Terraform {at(feature)} or later" in d.detail and d.rule == "terraform-version"
    assert errors(run(module_dir, text, at(feature))) == []


# ---------------------------------------------------------------- language behaviour
OPTIONAL = 'variable "o" {\n  type = object({\n    a = string\n    b = optional(string, "x")\n  })\n}\n'


def test_optional_attributes_need_the_experiment_before_1_3(module_dir):
    old = run(module_dir, OPTIONAL, "1.2")
    assert errors(old) == [("Optional object type attributes are experimental", 1)]
    assert old.diags[0].detail.startswith("This feature is currently an opt-in experiment")
    assert errors(run(module_dir, OPTIONAL, "1.3")) == []


def test_the_experimental_form_of_optional_took_no_default(module_dir):
    exp = 'terraform {\n  experiments = [module_variable_optional_attrs]\n}\n'
    res = run(module_dir, OPTIONAL, "1.2", **{"versions.tf": exp})
    assert errors(res) == [("Invalid type specification", 4)]
    assert [(d.severity, d.summary) for d in res.diags if d.file.endswith("versions.tf")] == [
        (WARNING, 'Experimental feature "module_variable_optional_attrs" is active')]
    one_arg = OPTIONAL.replace(', "x"', "")
    assert errors(run(module_dir, one_arg, "1.2", **{"versions.tf": exp})) == []
    # and from 1.3 on the experiment keyword itself is refused
    res = run(module_dir, one_arg, "1.3", **{"versions.tf": exp})
    assert errors(res) == [("Experiment has concluded", 2)]


CROSS = ('variable "lo" {\n  type = number\n}\nlocals {\n  cap = 10\n}\nvariable "hi" {\n  type = number\n'
         '  validation {\n    condition     = var.hi > var.lo && var.hi < local.cap\n'
         '    error_message = "hi must exceed ${var.lo}."\n  }\n}\n')


def test_validation_rules_see_only_their_own_variable_before_1_9(module_dir):
    old = run(module_dir, CROSS, "1.8")
    assert errors(old) == [("Invalid reference in variable validation", 10),
                           ("Invalid reference in variable validation", 10),
                           ("Invalid reference in variable validation", 11)]
    assert old.diags[0].detail.startswith(
        'The condition for variable "hi" can only refer to the variable itself, using var.hi.')
    assert old.diags[2].detail.startswith('The error message for variable "hi" can only refer to')
    assert errors(run(module_dir, CROSS, "1.9")) == []


def test_a_validation_rule_must_mention_its_own_variable_in_every_release(module_dir):
    text = 'variable "a" {}\nvariable "b" {\n  validation {\n    condition     = var.a != ""\n    error_message = "m"\n  }\n}\n'
    assert ("Invalid variable validation condition", 4) in errors(run(module_dir, text, "1.8"))
    assert errors(run(module_dir, text, "1.10")) == [("Invalid variable validation condition", 4)]


def test_for_expression_names_in_a_validation_rule_are_not_references(module_dir):
    text = ('variable "xs" {\n  type = list(string)\n  validation {\n'
            '    condition     = alltrue([for x in var.xs : length(x) > 0])\n    error_message = "m"\n  }\n}\n')
    assert errors(run(module_dir, text, "1.2")) == []


def test_provider_defined_functions_gate(module_dir):
    text = 'output "o" {\n  value = provider::aws::arn_parse("arn:aws:s3:::b")\n}\n'
    old = run(module_dir, text, "1.7")
    assert errors(old) == [("Provider-defined functions are not available", 2)]
    assert errors(run(module_dir, text, "1.8")) == []
    core = 'output "o" {\n  value = core::upper("a")\n}\n'
    assert errors(run(module_dir, core, "1.7")) == [("Call to unknown function", 2)]
    assert errors(run(module_dir, core, "1.8")) == []


def test_configuration_aliases_gate(module_dir):
    text = ('terraform {\n  required_providers {\n    aws = {\n      source                = "hashicorp/aws"\n'
            '      configuration_aliases = [aws.west]\n    }\n  }\n}\n')
    old = run(module_dir, text, "0.14")
    assert errors(old) == [("Invalid required_providers object", 3)]
    assert "needs Terraform 0.15 or later" in old.diags[0].detail
    assert errors(run(module_dir, text, "0.15")) == []


def test_terraform_applying_gate(module_dir):
    text = 'output "o" {\n  value = terraform.applying\n}\n'
    assert errors(run(module_dir, text, "1.9")) == [('Invalid "terraform" attribute', 2)]
    assert errors(run(module_dir, text, "1.10")) == []


@pytest.mark.parametrize("name,call", [
    ("anytrue", "anytrue([true])"), ("one", "one([1])"), ("nonsensitive", 'nonsensitive("a")'),
    ("startswith", 'startswith("ab", "a")'), ("strcontains", 'strcontains("ab", "a")'),
    ("issensitive", 'issensitive("a")'), ("templatestring", 'templatestring("a", {})'),
    ("ephemeralasnull", 'ephemeralasnull("a")'),
])
def test_function_availability_gate(module_dir, name, call):
    text = f'output "o" {{\n  value = {call}\n}}\n'
    since = tfversion.FUNCTION_SINCE[name]
    prev = tfversion.text(tfversion.SERIES[tfversion.SERIES.index(since) - 1])
    old = run(module_dir, text, prev)
    assert errors(old) == [("Call to unknown function", 2)]
    assert old.diags[0].detail.startswith(f'There is no function named "{name}".')
    assert f'"{name}" was added in Terraform {tfversion.text(since)}' in old.diags[0].detail
    assert errors(run(module_dir, text, tfversion.text(since))) == []


def test_list_and_map_functions_were_only_deprecated_before_0_15(module_dir):
    text = 'output "o" {\n  value = list("a", "b")\n}\n'
    old = run(module_dir, text, "0.14")
    assert errors(old) == [] and [(d.severity, d.line) for d in old.diags] == [(WARNING, 2)]
    new = run(module_dir, text, "0.15")
    assert errors(new) == [("Error in function call", 2)]


def test_quoted_type_keywords_were_only_deprecated_before_0_15(module_dir):
    text = 'variable "a" {\n  type = "list"\n}\n'
    old = run(module_dir, text, "0.14")
    assert errors(old) == [] and [(d.severity, d.summary) for d in old.diags] == [
        (WARNING, "Quoted type constraints are deprecated")]
    assert str(old.module.variables["a"].type) == "list(string)"
    assert errors(run(module_dir, text, "0.15")) == [("Invalid quoted type constraints", 2)]


def test_every_gated_function_is_a_known_function():
    from nvidia_terraform_modules_b200.config.functions import FUNCTIONS
    assert set(tfversion.FUNCTION_SINCE) <= set(FUNCTIONS)
    assert all(v in tfversion.SERIES for v in list(tfversion.SINCE.values()) + list(tfversion.FUNCTION_SINCE.values()))
