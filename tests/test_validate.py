"""validate: structural rules of the configuration loader and reference checks.
Each case is a tiny module written for the test; messages are compared with
Terraform's wording."""
import pytest

from nvidia_terraform_modules_b200.config.validate import validate_module


def diag_set(module_dir, files):
    res = validate_module(module_dir(files))
    return res, [(d.severity, d.summary) for d in res.diags]


def summaries(module_dir, text, **more):
    files = {"main.tf": text}
    files.update(more)
    return [s for _, s in diag_set(module_dir, files)[1]]


GOOD = {
    "versions.tf": 'terraform {\n  required_version = ">= 1.3"\n  required_providers {\n'
                   '    aws = {\n      source  = "hashicorp/aws"\n      version = "~> 4.0"\n    }\n  }\n}\n',
    "variables.tf": 'variable "name" {\n  type = string\n}\nvariable "on" {\n  type    = bool\n  default = true\n}\n'
                    'variable "zones" {\n  type    = list(string)\n  default = []\n}\n',
    "main.tf": '''
locals {
  tags = { for z in var.zones : z => upper(z) }
}
resource "aws_thing" "many" {
  count = var.on ? 1 : 0
  name  = "${var.name}-${count.index}"
  tags  = local.tags
}
resource "aws_thing" "each" {
  for_each = toset(var.zones)
  name     = each.key
  peer     = aws_thing.many[0].id
  dynamic "rule" {
    for_each = var.zones
    iterator = it
    content {
      zone = it.value
      all  = aws_thing.many[*].id
    }
  }
  provisioner "local-exec" {
    command = "echo ${self.id} ${path.module} ${terraform.workspace}"
  }
  lifecycle {
    ignore_changes = [tags]
  }
  depends_on = [aws_thing.many, data.aws_zone.z]
}
data "aws_zone" "z" {
  name = one(aws_thing.many[*].name)
}
module "child" {
  source = "./child"
  label  = var.name
}
output "ids" {
  value = [for k, t in aws_thing.each : t.id if k != ""]
}
output "child" {
  value = module.child.out
}
''',
    "child/main.tf": 'variable "label" {\n  type = string\n}\nvariable "opt" {\n  default = 1\n}\n'
                     'output "out" {\n  value = var.label\n}\n',
}


def test_well_formed_module_is_valid(module_dir):
    res, diags = diag_set(module_dir, GOOD)
    assert diags == [] and res.valid
    assert res.validator.used_vars == {"name", "on", "zones"}
    assert res.validator.used_locals == {"tags"}


# ------------------------------------------------------------- references
@pytest.mark.parametrize("expr,summary", [
    ("var.nope", "Reference to undeclared input variable"),
    ("local.nope", "Reference to undeclared local value"),
    ("module.nope.x", "Reference to undeclared module"),
    ("aws_x.nope.id", "Reference to undeclared resource"),
    ("data.aws_x.nope.id", "Reference to undeclared resource"),
    ("var", "Invalid reference"),
    ("aws_x", "Invalid reference"),
    ("data.aws_x", "Invalid reference"),
    ("count.index", 'Reference to "count" in non-counted context'),
    ("each.key", 'Reference to "each" in context without for_each'),
    ("self.id", 'Invalid "self" reference'),
    ("path.nowhere", 'Invalid "path" attribute'),
    ("terraform.version", 'Invalid "terraform" attribute'),
    ("uppercase(\"a\")", "Call to unknown function"),
    ("length()", "Not enough function arguments"),
    ("list(1, 2)", "Error in function call"),
    ("[for x in [1] : y]", "Invalid reference"),
])
def test_reference_errors(module_dir, expr, summary):
    got = summaries(module_dir, f'output "o" {{\n  value = {expr}\n}}\n')
    assert got == [summary]


def test_for_and_dynamic_names_are_scoped(module_dir):
    text = '''
resource "aws_a" "b" {
  dynamic "rule" {
    for_each = [1]
    content {
      v = rule.value
    }
  }
  leak = rule.value
}
'''
    # outside the content block "rule" is just an (undeclared) resource type
    assert summaries(module_dir, text) == ["Reference to undeclared resource"]


def test_count_and_each_contexts(module_dir):
    text = '''
resource "aws_a" "counted" {
  count = 2
  i     = count.index
  bad   = count.total
  worse = each.key
}
resource "aws_a" "plain" {
  i = count.index
}
resource "aws_a" "self_count" {
  count = count.index
}
'''
    assert summaries(module_dir, text) == ['Invalid "count" attribute',
                                           'Reference to "each" in context without for_each',
                                           'Reference to "count" in non-counted context',
                                           'Reference to "count" in non-counted context']


def test_instance_keys_must_match_count(module_dir):
    text = '''
resource "aws_a" "one" {}
resource "aws_a" "many" {
  count = 2
}
output "a" {
  value = aws_a.one[0].id
}
output "b" {
  value = aws_a.many.id
}
output "c" {
  value = [aws_a.one.id, aws_a.many[0].id, aws_a.many[*].id, aws_a.many, length(aws_a.many)]
}
output "d" {
  value = aws_a.one[count.index]
}
'''
    assert summaries(module_dir, text) == ["Unexpected resource instance key", "Missing resource instance key",
                                           "Unexpected resource instance key",
                                           'Reference to "count" in non-counted context']


def test_module_instance_keys_and_outputs(module_dir):
    files = {
        "main.tf": '''
module "one" {
  source = "./c"
  need   = 1
}
module "many" {
  source = "./c"
  count  = 2
  need   = count.index
}
output "a" {
  value = [module.one.out, module.many[0].out, module.many.*.out, module.many[*].out]
}
output "b" {
  value = module.one[0].out
}
output "c" {
  value = module.many.out
}
output "d" {
  value = module.one.missing
}
''',
        "c/main.tf": 'variable "need" {}\noutput "out" {\n  value = var.need\n}\n',
    }
    _, diags = diag_set(module_dir, files)
    assert [s for _, s in diags] == ["Unexpected module instance key", "Missing module instance key",
                                     "Unsupported attribute"]


def test_local_module_call_arguments(module_dir):
    files = {
        "main.tf": 'module "m" {\n  source = "./c"\n  extra  = 1\n}\n'
                   'module "gone" {\n  source = "./missing"\n}\n'
                   'module "remote" {\n  source  = "ns/name/aws"\n  version = "1.0.0"\n  anything = 1\n}\n',
        "c/main.tf": 'variable "need" {}\nvariable "opt" {\n  default = 1\n}\n',
    }
    _, diags = diag_set(module_dir, files)
    assert sorted(s for _, s in diags) == ["Missing required argument", "Unreadable module directory",
                                           "Unsupported argument"]


def test_depends_on_must_be_static_references(module_dir):
    text = '''
resource "aws_a" "b" {}
resource "aws_a" "c" {
  depends_on = [aws_a.b, "aws_a.b", upper("x"), aws_a.nope]
}
resource "aws_a" "d" {
  depends_on = aws_a.b
}
'''
    assert summaries(module_dir, text) == ["Quoted references are deprecated", "Invalid expression",
                                           "Reference to undeclared resource", "Invalid expression"]


def test_deprecation_warnings_do_not_make_a_module_invalid(module_dir):
    res, diags = diag_set(module_dir, {"main.tf": 'variable "a" {}\noutput "o" {\n  value = "${var.a}"\n}\n'
                                                  'provider "aws" {\n  version = "~> 4.0"\n}\n'})
    assert diags == [("warning", "Interpolation-only expressions are deprecated"),
                     ("warning", "Version constraints inside provider configuration blocks are deprecated")]
    assert res.valid


# --------------------------------------------------------------- structure
@pytest.mark.parametrize("text,summary", [
    ("x = 1\n", "Unsupported argument"),
    ("frob {}\n", "Unsupported block type"),
    ('resource "a" {}\n', "Missing name for resource"),
    ("resource {}\n", "Missing type for resource"),
    ('variable "a" "b" {}\n', "Extraneous label for variable"),
    ('locals "x" {}\n', "Extraneous label for locals"),
    ('resource "9a" "b" {}\n', "Invalid resource type name"),
    ('variable "count" {}\n', "Invalid variable name"),
    ('variable "a" {\n  typ = string\n}\n', "Unsupported argument"),
    ('variable "a" {\n  type = strng\n}\n', "Invalid type specification"),
    ('variable "a" {\n  type = "string"\n}\n', "Invalid quoted type constraints"),
    ('variable "a" {\n  type = "list(string)"\n}\n', "Invalid legacy variable type hint"),
    ('variable "a" {\n  type    = number\n  default = "x"\n}\n', "Invalid default value for variable"),
    ('variable "a" {\n  type    = string\n  default = var.b\n}\n', "Variables not allowed"),
    ('variable "a" {\n  default  = null\n  nullable = false\n}\n', "Invalid default value for variable"),
    ('variable "a" {}\nvariable "a" {}\n', "Duplicate variable declaration"),
    ('output "a" {}\n', "Missing required argument"),
    ('output "a" {\n  value = 1\n}\noutput "a" {\n  value = 2\n}\n', "Duplicate output definition"),
    ('locals {\n  a = 1\n}\nlocals {\n  a = 2\n}\n', "Duplicate local value definition"),
    ('resource "a_b" "c" {}\nresource "a_b" "c" {}\n', 'Duplicate resource "a_b" configuration'),
    ('data "a_b" "c" {}\ndata "a_b" "c" {}\n', 'Duplicate data "a_b" configuration'),
    ('module "m" {}\n', "Missing required argument"),
    ('module "m" {\n  source = "a/b/c"\n  version = "one"\n}\n', "Invalid version constraint"),
    ('module "m" {\n  source = var.s\n}\nvariable "s" {}\n', "Variables not allowed"),
    ('resource "a_b" "c" {\n  count    = 1\n  for_each = {}\n}\n', 'Invalid combination of "count" and "for_each"'),
    ('resource "a_b" "c" {\n  provider = 1 + 1\n}\n', "Invalid provider configuration reference"),
    ('resource "a_b" "c" {\n  lifecycle {\n    frob = 1\n  }\n}\n', "Unsupported argument"),
    ('data "a_b" "c" {\n  provisioner "x" {}\n}\n', "Data resources do not have provisioners"),
    ('resource "a_b" "c" {\n  dynamic "x" {\n    content {}\n  }\n}\n', "Missing required argument"),
    ('resource "a_b" "c" {\n  dynamic "x" {\n    for_each = []\n  }\n}\n', "Missing dynamic content block"),
    ('provider "aws" {}\nprovider "aws" {}\n', "Duplicate provider configuration"),
    ('provider "aws" {\n  alias = "9x"\n}\n', "Invalid provider configuration alias"),
    ('terraform {\n  required_version = ">> 1"\n}\n', "Invalid version constraint"),
    ('terraform {\n  frob = 1\n}\n', "Unsupported argument"),
    ('terraform {\n  required_providers {\n    aws = {\n      sauce = "x"\n    }\n  }\n}\n',
     "Invalid required_providers object"),
    ('terraform {\n  required_providers {\n    aws = {\n      source = "a/b/c/d"\n    }\n  }\n}\n',
     "Invalid provider source string"),
    ('terraform {\n  required_providers {\n    aws = "~> 4.0"\n    aws = "~> 5.0"\n  }\n}\n',
     "Duplicate required providers configuration"),
    ('resource "a_b" "c" {\n  x = 1\n  x = 2\n}\n', None),    # body arguments need the provider schema
])
def test_structural_errors(module_dir, text, summary):
    got = summaries(module_dir, text)
    assert got == ([summary] if summary else [])


def test_version_is_only_for_registry_sources(module_dir):
    _, diags = diag_set(module_dir, {"main.tf": 'module "m" {\n  source  = "./x"\n  version = "1.0"\n}\n',
                                     "x/main.tf": ""})
    assert [s for _, s in diags] == ["Invalid registry module source address"]


def test_self_references_and_cycles(module_dir):
    text = '''
locals {
  a = local.b
  b = local.c
  c = [local.a]
  me = local.me
  fine = local.a
}
resource "aws_a" "x" {
  name = aws_a.x.id
  peer = aws_a.y.id
  provisioner "local-exec" {
    command = "echo ${aws_a.x.id} ${self.id}"
  }
}
resource "aws_a" "y" {
  peer = data.aws_d.z.id
}
data "aws_d" "z" {
  v = aws_a.x.id
}
'''
    assert summaries(module_dir, text) == ["Cycle: local.a, local.b, local.c", "Self-referencing local value",
                                           "Cycle: aws_a.x, aws_a.y, data.aws_d.z", "Self-referential block"]


@pytest.mark.parametrize("text,summary", [
    ('variable "a" {\n  validation {\n    condition     = true\n    error_message = "m"\n  }\n}\n',
     "Invalid variable validation condition"),
    ('variable "a" {\n  validation {\n    condition = var.a != ""\n  }\n}\n', "Missing required argument"),
    ('output "o" {\n  value = 1\n  precondition {\n    error_message = "m"\n  }\n}\n', "Missing required argument"),
    ('resource "a_b" "c" {\n  lifecycle {\n    ignore_changes = "tags"\n  }\n}\n', "Invalid expression"),
    ('resource "a_b" "c" {\n  lifecycle {\n    ignore_changes = all\n  }\n}\n', None),
    ('resource "a_b" "c" {\n  lifecycle {\n    postcondition {\n      condition = self.id != ""\n'
     '      error_message = "m"\n    }\n  }\n}\n', None),
    ('variable "b" {}\nterraform {\n  backend "s3" {\n    bucket = var.b\n  }\n}\n', "Variables not allowed"),
    ('terraform {\n  backend "s3" {}\n  backend "gcs" {}\n}\n', "Duplicate backend configuration"),
    ('moved {\n  from = aws_a.b\n}\n', "Missing required argument"),
    ('import {\n  to = aws_a.b\n  id = "x"\n}\nresource "aws_a" "b" {}\n', None),
    ('check "health" {\n}\n', "Invalid check block"),
    ('check "health" {\n  assert {\n    condition     = 1 == 1\n    error_message = "m"\n  }\n}\n', None),
])
def test_more_structural_rules(module_dir, text, summary):
    assert summaries(module_dir, text) == ([summary] if summary else [])


@pytest.mark.parametrize("arg,summary", [
    ('count = "two"', "Incorrect value type"), ("count = null", "Invalid count argument"),
    ("count = -1", "Invalid count argument"), ("count = 1.5", "Invalid count argument"),
    ('count = "2"', None), ("count = 3", None), ("count = var.n", None),
    ('for_each = ["a", "b"]', "Invalid for_each argument"), ("for_each = [for x in var.n : x]", "Invalid for_each argument"),
    ('for_each = "a"', "Invalid for_each argument"), ('for_each = toset(["a"])', None),
    ("for_each = { a = 1 }", None), ("for_each = { for x in var.n : x => x }", None),
])
def test_literal_count_and_for_each(module_dir, arg, summary):
    text = f'variable "n" {{}}\nresource "a_b" "c" {{\n  {arg}\n}}\n'
    assert summaries(module_dir, text) == ([summary] if summary else [])


@pytest.mark.parametrize("expr,summary", [
    ('var.n ? "a" : ["b"]', "Inconsistent conditional result types"),
    ("var.n ? {} : []", "Inconsistent conditional result types"),
    ("var.n ? 1 : false", "Inconsistent conditional result types"),
    ("var.n ? length(var.n) : keys(var.n)", "Inconsistent conditional result types"),
    ('var.n ? "a" : 1', None), ("var.n ? [] : var.n", None), ("var.n ? null : {}", None),
    ("var.n ? { a = 1 } : {}", None), ("var.n ? [1] : tolist([])", None),
    ('var.n ? "${var.n}" : []', "Interpolation-only expressions are deprecated"),   # passes var.n through: no type clash
    ("[1] + 1", "Invalid operand"), ('"abc" * 2', "Invalid operand"), ('"12" * 2', None),
    ("{} < 1", "Invalid operand"), ("1 && true", "Invalid operand"), ("!1", "Invalid operand"),
    ('!"true"', None), ("-[1]", "Invalid operand"), ("var.n + 1", None), ("[1] == [1]", None),
    ("length(var.n) > 0 && var.n != null", None), ("upper(var.n) + 1", None),
])
def test_static_type_checks(module_dir, expr, summary):
    text = f'variable "n" {{}}\noutput "o" {{\n  value = {expr}\n}}\n'
    assert summaries(module_dir, text) == ([summary] if summary else [])


def test_aliased_providers_are_distinct(module_dir):
    assert summaries(module_dir, 'provider "aws" {}\nprovider "aws" {\n  alias = "west"\n}\n'
                                 'resource "aws_a" "b" {\n  provider = aws.west\n}\n') == []


def test_syntax_errors_are_reported_with_file_names(module_dir):
    res, _ = diag_set(module_dir, {"main.tf": 'output "a" {\n  value = (\n}\n', "ok.tf": 'variable "v" {}\n'})
    assert not res.valid
    assert [(d.rule, d.summary) for d in res.diags] == [("syntax", "Invalid expression")]
    assert res.diags[0].file.endswith("main.tf")
    assert "v" in res.module.variables          # the healthy file was still loaded
    assert "a" not in res.module.outputs        # nothing is salvaged from the broken one


# ------------------------------------------------------------------ tfvars
def test_tfvars_are_checked_against_declarations(module_dir):
    files = {
        "main.tf": 'variable "n" {\n  type = number\n}\nvariable "s" {\n  type = string\n}\n',
        "terraform.tfvars": 'n = "many"\ns = var.n\nghost = 1\n',
        "extra.auto.tfvars": 'blk {}\n',
        "ignored.tfvars": 'ghost2 = 1\n',
    }
    res, diags = diag_set(module_dir, files)
    assert sorted(s for _, s in diags) == ["Invalid value for input variable", "Unexpected block",
                                           "Value for undeclared variable", "Variables not allowed"]


def test_module_providers_map_takes_dotted_keys(module_dir):
    """`providers = { aws.src = aws.a }` is the documented way to hand aliased provider
    configurations to a child; it used to be rejected as a syntax error (and fmt refused
    the file)."""
    from nvidia_terraform_modules_b200.fmt.command import format_source
    child = ('terraform {\n  required_providers {\n    aws = {\n      source                = "hashicorp/aws"\n'
             '      configuration_aliases = [aws.src, aws.dst]\n    }\n  }\n}\n')
    root = ('provider "aws" {\n  alias = "a"\n}\nprovider "aws" {\n  alias = "b"\n}\n'
            'module "copy" {\n  source = "./child"\n  providers = {\n    aws.src = aws.a\n    aws.dst   = aws.b\n  }\n}\n')
    assert summaries(module_dir, root, **{"child/main.tf": child}) == []
    out, diags = format_source(root, "main.tf")
    assert not diags and "    aws.src = aws.a\n    aws.dst = aws.b\n" in out


def test_dotted_key_in_an_ordinary_object_is_ambiguous(module_dir):
    text = 'locals {\n  a = { x.y = 1 }\n  b = { (local.k) = 1, "x.y" = 2 }\n  k = "k"\n}\n'
    res, diags = diag_set(module_dir, {"main.tf": text})
    assert diags == [("error", "Ambiguous attribute key")]
    assert (res.diags[0].line, res.diags[0].col) == (2, 9)


def test_child_configuration_aliases_must_be_handed_over(module_dir):
    child = ('terraform {\n  required_providers {\n    aws = {\n      source                = "hashicorp/aws"\n'
             '      configuration_aliases = [aws.src, aws.dst]\n    }\n  }\n}\n')
    root = ('provider "aws" {\n  alias = "a"\n}\nmodule "copy" {\n  source = "./child"\n  providers = {\n'
            '    aws.src = aws.a\n  }\n}\nmodule "none" {\n  source = "./child"\n}\n')
    res, diags = diag_set(module_dir, {"main.tf": root, "child/main.tf": child})
    assert diags == [("error", "Missing required provider configuration")] * 3
    assert [d.line for d in res.diags] == [4, 10, 10]
    assert res.diags[0].detail.startswith('The child module requires an additional configuration for provider '
                                          'hashicorp/aws, with the local name "aws.dst".')
