"""Canonical layout: hand-written input/expected pairs.

The expected texts were written by hand from Terraform's documented style
conventions ("Style Conventions": two-space indent, aligned equals signs for
consecutive single-line arguments, ...) and from how files in the wild look
after ``terraform fmt`` -- NOT by running the code under test.  Parity with the
real binary is unpinned (none is available offline); tests/test_reference_evidence.py
adds the evidence the reference tree itself provides.
"""
import pytest

from nvidia_terraform_modules_b200.fmt.command import format_source
from nvidia_terraform_modules_b200.hcl.scanner import scan
from nvidia_terraform_modules_b200.hcl.tokens import T


def fmt(src):
    out, diags = format_source(src)
    assert not diags, [d.format() for d in diags]
    return out


CASES = {
    "align_equals_in_a_group": (
        'resource "a" "b" {\n  ami = "abc"\n  instance_type   = "t2.micro"\n}\n',
        'resource "a" "b" {\n  ami           = "abc"\n  instance_type = "t2.micro"\n}\n',
    ),
    "blank_line_starts_a_new_group": (
        'r "a" "b" {\n  a = 1\n  long_name = 2\n\n  b = 3\n  cc = 4\n}\n',
        'r "a" "b" {\n  a         = 1\n  long_name = 2\n\n  b  = 3\n  cc = 4\n}\n',
    ),
    "comment_line_starts_a_new_group": (
        'm "x" {\n  source = "s"\n  # note\n  count = 1\n  name_long = 2\n}\n',
        'm "x" {\n  source = "s"\n  # note\n  count     = 1\n  name_long = 2\n}\n',
    ),
    "multi_line_value_is_not_aligned_and_splits_groups": (
        'r "a" "b" {\n  name = "n"\n  location = "l"\n  tags = {\n    a = 1\n    bbb = 2\n  }\n  x = 1\n}\n',
        'r "a" "b" {\n  name     = "n"\n  location = "l"\n  tags = {\n    a   = 1\n    bbb = 2\n  }\n  x = 1\n}\n',
    ),
    "tabs_and_wrong_indent_become_two_spaces": (
        'set {\n\tname   = "driver.version"\n        value  = var.v\n  }\n',
        'set {\n  name  = "driver.version"\n  value = var.v\n}\n',
    ),
    "spacing_inside_expressions": (
        'a=1\nb =  [ 1,2 , 3 ]\nc = foo( 1,2 )\nd = { x=1 }\ne = {}\nf = var . x [ 0 ]\n',
        'a = 1\nb = [1, 2, 3]\nc = foo(1, 2)\nd = { x = 1 }\ne = {}\nf = var.x[0]\n',
    ),
    "unary_minus_versus_subtraction": (
        'a = -1\nb = x - 1\nc = x -1\nd = (-y)\ne = [-1, - 2]\nf = !  ok\ng = a ? -1 : - 2\n',
        'a = -1\nb = x - 1\nc = x - 1\nd = (-y)\ne = [-1, -2]\nf = !ok\ng = a ? -1 : -2\n',
    ),
    "operators_and_conditionals": (
        'a = x==1?"y":"n"\nb = p&&q||!r\n',
        'a = x == 1 ? "y" : "n"\nb = p && q || !r\n',
    ),
    "trailing_whitespace_is_removed": (
        'a = 1   \n\nb = 2\t\n   \n',
        'a = 1\n\nb = 2\n\n',
    ),
    "trailing_comments_are_aligned": (
        'm "x" {\n  a = var.a # one\n  bbbb = var.bbbbbb # two\n  c = 1\n}\n',
        'm "x" {\n  a    = var.a      # one\n  bbbb = var.bbbbbb # two\n  c    = 1\n}\n',
    ),
    "heredoc_body_is_left_alone": (
        'v "x" {\n  type = string\n  description = <<EOH\n  (Required)   text \n\tkept\n  EOH\n}\n',
        'v "x" {\n  type        = string\n  description = <<EOH\n  (Required)   text \n\tkept\n  EOH\n}\n',
    ),
    "template_interpolation_is_tightened": (
        'a = "x-${ var.a }-${b [0]}"\n',
        'a = "x-${var.a}-${b[0]}"\n',
    ),
    "interpolation_only_value_is_unwrapped": (
        'a = "${var.a}"\nb = "${var.a}-${var.b}"\nc = ["${var.c}"]\nd = "${ f("${x}") }"\n',
        'a = var.a\nb = "${var.a}-${var.b}"\nc = ["${var.c}"]\nd = f("${x}")\n',
    ),
    "legacy_variable_types": (
        'variable "a" {\n  type = "string"\n}\nvariable "b" {\n  type = "list"\n}\nvariable "c" {\n  type = map\n}\n'
        'variable "d" {\n  type = list(string)\n}\n',
        'variable "a" {\n  type = string\n}\nvariable "b" {\n  type = list(string)\n}\nvariable "c" {\n  type = map(any)\n}\n'
        'variable "d" {\n  type = list(string)\n}\n',
    ),
    "quoted_type_outside_variable_blocks_is_untouched": (
        'output "a" {\n  type = "string"\n}\n',
        'output "a" {\n  type = "string"\n}\n',
    ),
    "labels_are_quoted": (
        'resource aws_instance web {\n  provisioner local-exec {\n  }\n}\n',
        'resource "aws_instance" "web" {\n  provisioner "local-exec" {\n  }\n}\n',
    ),
    "for_expression_spacing": (
        'a = [for x in var.l: upper(x) if x!=""]\nb = {for k,v in m: k=>v...}\nc = [for x in [1,2]: x]\n',
        'a = [for x in var.l : upper(x) if x != ""]\nb = { for k, v in m : k => v... }\nc = [for x in [1, 2] : x]\n',
    ),
    "json_style_object_keys_keep_a_space_before_the_colon": (
        'p = jsonencode({\n"Version":"2012-10-17",\n  "Statement" : [\n{ "Effect":"Allow",\n"Resource" : "*"\n}\n]\n})\n',
        'p = jsonencode({\n  "Version" : "2012-10-17",\n  "Statement" : [\n    { "Effect" : "Allow",\n      "Resource" : "*"\n    }\n  ]\n})\n',
    ),
    "two_closers_on_one_line_dedent_twice": (
        'locals {\n  r = var.on ? {\n    rule = {\n      port = 1\n  } } : {}\n  s = 1\n}\n',
        'locals {\n  r = var.on ? {\n    rule = {\n      port = 1\n  } } : {}\n  s = 1\n}\n',
    ),
    "multi_line_call_arguments_are_indented": (
        'provider "k" {\n  host = "h"\n  cert = base64decode(\n  x.y.0.z,\n  )\n}\n',
        'provider "k" {\n  host = "h"\n  cert = base64decode(\n    x.y.0.z,\n  )\n}\n',
    ),
    "one_line_blocks": (
        'provider "azurerm" {\n  features {   }\n}\nidentity {type="x"}\n',
        'provider "azurerm" {\n  features {}\n}\nidentity { type = "x" }\n',
    ),
    "block_comment_keeps_its_inner_lines": (
        '  /****\n   text\n  ****/\nlocals {\n/* in */ a = 1\n}\n',
        '/****\n   text\n  ****/\nlocals {\n  /* in */ a = 1\n}\n',
    ),
    "splat_and_legacy_index": (
        'a = module . vpc .* . ids\nb = x.y.0.z\nc = x [*] . id\n',
        'a = module.vpc.*.ids\nb = x.y.0.z\nc = x[*].id\n',
    ),
    "no_trailing_newline_is_preserved": (
        'a   = 1',
        'a = 1',
    ),
    "namespaced_function": (
        'a = provider :: aws :: f ( 1 )\n',
        'a = provider::aws::f(1)\n',
    ),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_canonical_form(name):
    src, want = CASES[name]
    assert fmt(src) == want


@pytest.mark.parametrize("name", sorted(CASES))
def test_idempotent(name):
    once = fmt(CASES[name][0])
    assert fmt(once) == once


@pytest.mark.parametrize("name", sorted(CASES))
def test_layout_pass_changes_white_space_only(name):
    """Apart from the three Terraform rewrites, formatting never touches a token."""
    if name in ("interpolation_only_value_is_unwrapped", "legacy_variable_types", "labels_are_quoted"):
        pytest.skip("exercises a token rewrite")
    src = CASES[name][0]

    def sig(text):
        return [(t.type, t.text) for t in scan(text) if t.type is not T.EOF]
    assert sig(fmt(src)) == sig(src)


def test_empty_and_comment_only_files():
    assert fmt("") == ""
    assert fmt("# just a comment\n") == "# just a comment\n"
    assert fmt("   # indented comment\n") == "# indented comment\n"


def test_syntax_error_leaves_text_alone():
    src = "a = (\n"
    out, diags = format_source(src, "x.tf")
    assert out == src and diags and diags[0].file == "x.tf"


def test_multi_line_interpolation_only_value_gets_parentheses():
    # newlines just inside the braces are trimmed; no parentheses needed then
    assert fmt('a = "${\n  x ? 1 : 2\n}"\n') == 'a = x ? 1 : 2\n'
    out = fmt('a = "${x ?\n 1 : 2}"\n')
    assert out.startswith("a = (x ?") and out.rstrip().endswith(")")


def test_byte_order_mark_survives():
    assert fmt("\ufeffa  = 1\n") == "\ufeffa = 1\n"
