"""Lint rules: one positive and one negative case per rule."""
import pytest

from nvidia_terraform_modules_b200.config.validate import validate_module
from nvidia_terraform_modules_b200.lint.rules import RULES, run_rules

HEADER = ('terraform {\n  required_version = ">= 1.3"\n  required_providers {\n    aws = {\n'
          '      source  = "hashicorp/aws"\n      version = "~> 4.0"\n    }\n  }\n}\n')


def lint(module_dir, files, only=None):
    res = validate_module(module_dir(files))
    return [(d.rule, d.line) for d in run_rules(res, only={only} if only else None)]


def rules_hit(module_dir, text, only, **more):
    files = {"main.tf": text}
    files.update({k.replace("__", "/").replace("_tf", ".tf"): v for k, v in more.items()})
    return lint(module_dir, files, only)


def test_rule_ids_are_unique_and_documented():
    ids = [r.id for r in RULES]
    assert len(ids) == len(set(ids)) and len(ids) >= 20
    assert all(r.summary and r.severity in ("warning", "notice") for r in RULES)


def test_tidy_module_has_no_findings(module_dir):
    files = {
        "versions.tf": HEADER,
        "variables.tf": 'variable "name" {\n  type        = string\n  description = "n"\n}\n',
        "main.tf": 'resource "aws_thing" "a" {\n  name = var.name\n}\n',
        "outputs.tf": 'output "id" {\n  description = "id"\n  value       = aws_thing.a.id\n}\n',
    }
    assert lint(module_dir, files) == []


@pytest.mark.parametrize("rule,text,lines", [
    ("unused-variable", 'variable "a" {}\nvariable "b" {}\noutput "o" {\n  value = var.a\n}\n', [2]),
    ("unused-local", 'locals {\n  a = 1\n  b = local.a\n}\n', [3]),
    ("unused-data-source", 'data "aws_a" "x" {}\ndata "aws_a" "y" {}\noutput "o" {\n  value = data.aws_a.x.id\n}\n', [2]),
    ("module-unpinned", 'module "a" {\n  source = "ns/n/aws"\n}\nmodule "b" {\n  source  = "ns/n/aws"\n'
                        '  version = "1.0.0"\n}\nmodule "c" {\n  source = "git::https://x/y.git"\n}\n'
                        'module "d" {\n  source = "git::https://x/y.git?ref=v1"\n}\n', [1, 8]),
    ("required-provider-missing", HEADER + 'resource "aws_a" "x" {}\nresource "helm_release" "y" {}\n'
                                           'provider "kubernetes" {}\n', [11, 12]),
    ("required-provider-unused", HEADER + 'resource "null_resource" "x" {}\n', [4]),
    ("provider-version-missing", 'terraform {\n  required_providers {\n    aws = {\n      source = "hashicorp/aws"\n'
                                 '    }\n  }\n}\n', [3]),
    ("required-version-missing", 'resource "aws_a" "x" {}\n', [1]),
    ("undocumented", 'variable "a" {}\nvariable "b" {\n  description = "d"\n}\noutput "o" {\n  value = 1\n}\n', [1, 5]),
    ("untyped-variable", 'variable "a" {}\nvariable "b" {\n  type = string\n}\n', [1]),
    ("naming-convention", 'resource "aws_a" "my-thing" {}\nresource "aws_a" "my_thing" {}\n'
                          'variable "Upper" {}\n', [1, 3]),
    ("deprecated-index", 'output "o" {\n  value = [aws_a.x.list.0.id, aws_a.x.list[0].id]\n}\nresource "aws_a" "x" {}\n',
     [2]),
    ("legacy-splat", 'resource "aws_a" "x" {\n  count = 1\n}\noutput "o" {\n  value = [aws_a.x.*.id, aws_a.x[*].id]\n}\n',
     [5]),
    ("comment-syntax", '// old style\n# new style\n/* block */\nlocals {}\n', [1]),
    ("length-of-object", 'data "aws_a" "x" {}\ndata "aws_a" "n" {\n  count = 2\n}\nlocals {\n'
                         '  bad  = length(data.aws_a.x)\n  ok1  = length(data.aws_a.x.ids)\n'
                         '  ok2  = length(data.aws_a.n)\n}\n', [6]),
    ("duplicate-resource-arguments",
     'variable "r" {}\nresource "aws_att" "gpu" {\n  role = var.r\n  arn  = "a"\n}\n'
     'resource "aws_att" "cpu" {\n  role = var.r\n  arn  = "a"\n}\n'
     'resource "aws_att" "other" {\n  role = var.r\n  arn  = "b"\n}\n', [6]),
    ("guard-mismatch",
     'variable "a" {}\nvariable "b" {}\nresource "aws_r" "x" {\n  count = var.a ? 1 : 0\n}\n'
     'output "bad" {\n  value = var.b == false ? null : aws_r.x[0].arn\n}\n'
     'output "good" {\n  value = var.a ? aws_r.x[0].arn : null\n}\n', [7]),
    ("secret-in-command",
     'data "az_ws" "w" {}\nresource "null_resource" "n" {\n  provisioner "local-exec" {\n'
     '    command = "kubectl create secret --from-literal=K=${data.az_ws.w.primary_shared_key} '
     '--id=${data.az_ws.w.workspace_id}"\n  }\n}\n', [4]),
    ("counterpart-mismatch",
     'variable "gpu_tags" {}\nvariable "cpu_tags" {}\nresource "x_pool" "cpu_nodes" {\n  tags = var.gpu_tags\n}\n'
     'resource "x_pool" "gpu_nodes" {\n  tags = var.gpu_tags\n}\nresource "x_pool" "both_cpu_gpu" {\n'
     '  tags = var.gpu_tags\n}\nlocals {\n  pools = {\n    cpu_pool = { t = var.cpu_tags }\n'
     '    private-a = { t = var.public_x }\n  }\n}\nvariable "public_x" {}\n', [4, 15]),
    ("duplicate-object-key", 'locals {\n  a = {\n    k = 1\n    "k" = 2\n    j = 3\n  }\n}\n', [4]),
    ("empty-list-equality", 'variable "l" {}\nlocals {\n  a = var.l == []\n  b = length(var.l) == 0\n'
                            '  c = [] != var.l\n}\n', [3, 5]),
    ("deprecated-lookup", 'variable "m" {}\nlocals {\n  a = lookup(var.m, "k")\n  b = lookup(var.m, "k", 1)\n}\n', [3]),
    ("sensitive-output", 'variable "pw" {\n  sensitive = true\n}\nvariable "n" {}\n'
                         'output "a" {\n  value = "x-${var.pw}"\n}\n'
                         'output "b" {\n  value     = var.pw\n  sensitive = true\n}\n'
                         'output "c" {\n  value = var.n\n}\n', [6]),
    ("local-exec", 'resource "null_resource" "n" {\n  provisioner "local-exec" {\n    command = "helm repo add x y"\n'
                   '  }\n}\n', [2]),
])
def test_rule(module_dir, rule, text, lines):
    assert rules_hit(module_dir, text, rule) == [(rule, ln) for ln in lines]


def test_standard_module_structure(module_dir):
    hits = lint(module_dir, {"main.tf": 'variable "a" {}\n', "output.tf": 'output "o" {\n  value = 1\n}\n'},
                "standard-module-structure")
    # variables.tf and outputs.tf missing; one variable and one output misplaced
    assert len(hits) == 4


def test_eof_newline_includes_tfvars(module_dir):
    hits = lint(module_dir, {"main.tf": "locals {}", "ok.tf": "locals {}\n", "terraform.tfvars": "# c"},
                "eof-newline")
    assert hits == [("eof-newline", 1), ("eof-newline", 1)]


def test_provider_in_called_module(module_dir):
    files = {
        "main.tf": 'module "c" {\n  source = "./c"\n}\nmodule "d" {\n  source = "./d"\n  count  = 1\n}\n'
                   'module "e" {\n  source = "./e"\n}\n',
        "c/main.tf": 'provider "helm" {}\n',
        "d/main.tf": 'provider "helm" {}\n',
        "e/main.tf": 'locals {}\n',
    }
    res = validate_module(module_dir(files))
    found = [(d.severity, d.line) for d in run_rules(res, only={"provider-in-module"})]
    assert found == [("notice", 1)]          # the counted call is validate's error, not repeated here
    assert [d.summary for d in res.diags] == ["Module is incompatible with count, for_each, and depends_on"]


def test_unused_provider_passed_to_child_is_used(module_dir):
    text = HEADER + 'module "c" {\n  source = "./c"\n  providers = {\n    aws = aws\n  }\n}\n'
    assert rules_hit(module_dir, text, "required-provider-unused", c__main_tf="") == []


def test_only_and_skip(module_dir):
    res = validate_module(module_dir({"main.tf": '// c\nvariable "a" {}\n'}))
    everything = {d.rule for d in run_rules(res)}
    assert {"comment-syntax", "unused-variable"} <= everything
    assert {d.rule for d in run_rules(res, only={"comment-syntax"})} == {"comment-syntax"}
    assert "comment-syntax" not in {d.rule for d in run_rules(res, skip={"comment-syntax"})}


def test_helper_script_rule(module_dir):
    files = {
        "main.tf": 'data "external" "x" {\n  program = ["bash", "${path.module}/used.sh"]\n}\n',
        "used.sh": "#!/bin/bash\necho '{}'\n",
        "stale.sh": "# licence header\n\n#!/bin/bash\nset -e\neval \"$(jq -r '@sh \"A=\\(.a)\"')\"\n",
    }
    res = validate_module(module_dir(files))
    got = [(d.file.rsplit("/", 1)[-1], d.line) for d in run_rules(res, only={"helper-script"})]
    assert got == [("stale.sh", 1), ("stale.sh", 3), ("stale.sh", 5)]


def test_ignore_comments(module_dir):
    text = ('# tfcheck-ignore: unused-variable\nvariable "a" {}\n'
            'variable "b" {} # tfcheck-ignore: all\n'
            '# tfcheck-ignore: untyped-variable\nvariable "c" {}\n'
            '# an ordinary comment\nvariable "d" {}\n')
    res = validate_module(module_dir({"main.tf": text}))
    got = [(d.rule, d.line) for d in run_rules(res, only={"unused-variable", "untyped-variable"})]
    assert got == [("untyped-variable", 2), ("unused-variable", 5), ("untyped-variable", 7), ("unused-variable", 7)]
