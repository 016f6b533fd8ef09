"""Parser: tree shapes for every expression form, Terraform-worded errors with
positions, and recovery after an error."""
import pytest

from nvidia_terraform_modules_b200.hcl import ast as A
from nvidia_terraform_modules_b200.hcl.parser import parse, unescape_quoted


def expr(text):
    body, diags, _ = parse(f"x = {text}\n")
    assert not diags, [d.format() for d in diags]
    return body.items[0].expr


def errors(src):
    _, diags, _ = parse(src, "t.tf")
    return [(d.summary, d.line, d.col) for d in diags]


# ---------------------------------------------------------------- structure
def test_attributes_and_blocks():
    body, diags, _ = parse('a = 1\nresource "t" "n" {\n  b = 2\n  nested {\n    c = 3\n  }\n}\n')
    assert not diags
    attr, block = body.items
    assert isinstance(attr, A.Attribute) and attr.name == "a"
    assert block.type == "resource" and block.label_values == ["t", "n"]
    assert [l.quoted for l in block.labels] == [True, True]
    assert block.body.attributes["b"].expr.value == 2
    assert block.body.blocks[0].type == "nested"
    assert (block.pos.line, block.pos.col) == (2, 1)


def test_one_line_blocks():
    body, diags, _ = parse('features {}\nidentity { type = "x" }\n')
    assert not diags
    assert body.items[0].one_line and body.items[0].body.items == []
    assert body.items[1].body.attributes["type"].expr.literal_value() == "x"


def test_unquoted_labels_are_recorded_as_such():
    body, _, _ = parse("resource aws_x y {\n}\n")
    assert [l.quoted for l in body.items[0].labels] == [False, False]


def test_comment_at_end_of_attribute_acts_as_newline():
    body, diags, _ = parse("a = 1 # c\nb = 2 // d\n")
    assert not diags and [i.name for i in body.items] == ["a", "b"]


def test_file_without_trailing_newline():
    body, diags, _ = parse('output "o" {\n  value = 1\n}')
    assert not diags and body.items[0].type == "output"


# -------------------------------------------------------------- expressions
def test_literals():
    assert expr("null").value is None
    assert expr("true").value is True
    assert expr("12").value == 12
    assert expr("1.5").value == 1.5


def test_precedence():
    e = expr("1 + 2 * 3 == 7 && !false || false")
    assert isinstance(e, A.Binary) and e.op == "||"
    assert e.lhs.op == "&&"
    assert e.lhs.lhs.op == "=="
    assert e.lhs.lhs.lhs.op == "+" and e.lhs.lhs.lhs.rhs.op == "*"
    assert isinstance(e.lhs.rhs, A.Unary) and e.lhs.rhs.op == "!"


def test_conditional_is_right_associative_and_lowest():
    e = expr("a ? b : c ? d : e")
    assert isinstance(e, A.Conditional) and isinstance(e.false, A.Conditional)


def test_scope_traversal_with_static_steps():
    e = expr('module.eks.groups["gpu"][0].name')
    assert isinstance(e, A.ScopeTraversal) and e.root == "module"
    assert [(s.kind, s.value) for s in e.steps] == [("attr", "eks"), ("attr", "groups"), ("index", "gpu"),
                                                    ("index", 0), ("attr", "name")]


def test_dynamic_index_and_relative_traversal():
    e = expr("random_string.amp[count.index].result")
    assert isinstance(e, A.RelativeTraversal) and e.steps[0].value == "result"
    assert isinstance(e.source, A.Index)
    assert e.source.coll.dotted() == "random_string.amp"
    assert e.source.key.dotted() == "count.index"


def test_legacy_index_and_splats():
    e = expr("a.b.0.c")
    assert e.steps[1].legacy and e.steps[1].value == 0
    s = expr("module.vpc.*.private_subnets")
    assert isinstance(s, A.Splat) and not s.full and s.steps[0].value == "private_subnets"
    f = expr("aws_x.y[*].z[0]")
    assert isinstance(f, A.Splat) and f.full and [st.kind for st in f.steps] == ["attr", "index"]


def test_function_calls():
    e = expr("merge(local.a, var.b...)")
    assert isinstance(e, A.FunctionCall) and e.name == "merge" and e.expand_final and len(e.args) == 2
    e = expr("base64decode(\n  x.y,\n)")
    assert len(e.args) == 1
    assert expr("provider::aws::arn_parse(x)").name == "provider::aws::arn_parse"


def test_tuple_and_object_constructors():
    t = expr('[1, "a", [2],]')
    assert isinstance(t, A.TupleCons) and len(t.items) == 3
    o = expr('{\n  a = 1\n  "b" : 2, (var.k) = 3\n}')
    assert isinstance(o, A.ObjectCons)
    assert [k.literal_key() for k, _ in o.items] == ["a", "b", None]
    assert o.colon_keys == [False, True, False]
    assert o.items[2][0].forced_expr
    assert o.get("a").value == 1 and o.get("zz") is None


def test_dotted_object_key_parses_and_is_judged_only_when_evaluated():
    """HCL raises "Ambiguous attribute key" at evaluation time, precisely so that a static
    reader -- Terraform's ``providers = { aws.west = aws.usw2 }`` -- may accept it.  The
    parser therefore takes it; config/validate.py objects where the object is a value."""
    assert errors("x = { a.b = 1 }\n") == []
    o = expr("{ aws.west = aws.usw2 }")
    key, value = o.items[0]
    assert key.bare_name is None and key.literal_key() is None
    assert isinstance(key.wrapped, A.ScopeTraversal) and key.wrapped.dotted() == "aws.west"
    assert value.dotted() == "aws.usw2"


def test_for_expressions():
    e = expr('[for k, v in var.m : "${k}=${v}" if v != null]')
    assert isinstance(e, A.ForExpr) and (e.key_var, e.val_var) == ("k", "v")
    assert e.key_expr is None and e.cond is not None
    o = expr("{ for s in var.l : s => upper(s)... }")
    assert o.key_expr is not None and o.group and o.key_var is None


def test_for_is_only_a_keyword_in_position():
    t = expr("[for, x]")            # 'for' followed by ',' is an ordinary name
    assert isinstance(t, A.TupleCons) and t.items[0].root == "for"
    t = expr("{ for = 1 }")
    assert isinstance(t, A.ObjectCons) and t.items[0][0].literal_key() == "for"


def test_templates():
    t = expr('"tf-${var.name}-vpc"')
    assert isinstance(t, A.Template) and t.parts[0] == "tf-" and t.parts[2] == "-vpc"
    assert isinstance(t.parts[1], A.ScopeTraversal)
    assert expr('"plain"').literal_value() == "plain"
    assert expr('"${var.a}"').is_single_interpolation()
    assert not expr('"${var.a}b"').is_single_interpolation()


def test_template_directives():
    t = expr('"%{ if a }yes%{ else }no%{ endif } %{ for x in l }${x}%{ endfor }"')
    assert isinstance(t.parts[0], A.TemplateIf) and t.parts[0].then == ["yes"] and t.parts[0].otherwise == ["no"]
    assert isinstance(t.parts[2], A.TemplateFor) and t.parts[2].val_var == "x"


def test_heredoc_and_indent_stripping():
    body, diags, _ = parse("a = <<EOT\n  keep\n    more\nEOT\nb = <<-EOT\n    one\n      two\n    EOT\n")
    assert not diags
    assert body.items[0].expr.literal_value() == "  keep\n    more\n"
    assert body.items[1].expr.literal_value() == "one\n  two\n"
    assert body.items[1].expr.heredoc and body.items[1].expr.indent_strip


def test_unescape():
    assert unescape_quoted('a\\n\\t\\"\\\\ $${x} %%{y} \\u00e9') == 'a\n\t"\\ ${x} %{y} é'


def test_dash_belongs_to_the_identifier_before_it():
    # HCL names may contain dashes, so "a-1" is a single name and "a - 1" a subtraction
    assert expr("var.a-1").steps[0].value == "a-1"
    assert isinstance(expr("var.a - 1"), A.Binary)
    assert errors('x = true-"s"\n') == [("Missing newline after argument", 1, 10)]


def test_parentheses_allow_newlines():
    e = expr("(\n  a\n  + b\n)")
    assert isinstance(e, A.Parens) and e.inner.op == "+"


def test_token_ranges_cover_the_expression():
    body, _, toks = parse("x = foo(1, [2])  # c\n")
    a = body.items[0]
    text = "".join(t.raw_before + t.text for t in toks[a.expr_tok_start:a.expr_tok_end]).strip()
    assert text == "foo(1, [2])"


# ------------------------------------------------------------------ errors
@pytest.mark.parametrize("src,want", [
    ("a = \n", [("Invalid expression", 1, 5)]),
    ("a = 1 b = 2\n", [("Missing newline after argument", 1, 7)]),
    ("a { b = 1, c = 2 }\n", [("Invalid single-argument block definition", 1, 10)]),
    ('resource "a" {\n  x = 1\n', [("Unclosed configuration block", 1, 14)]),
    ("a = [1, 2\n", [("Unterminated tuple constructor expression", 1, 5)]),
    ("a = {\n  b = 1\n", [("Unterminated object constructor expression", 1, 5)]),
    ('v = "abc\n', [("Unterminated template string", 1, 5)]),
    ("a = c ? 1\n", [("Missing false expression in conditional", 1, 10)]),
    ("a = f(1 2)\n", [("Missing argument separator", 1, 9)]),
    ("a = x.\n", [("Invalid attribute name", 1, 7)]),
    ("a = x[0\n", [("Missing close bracket on index", 2, 1)]),
    ("block\n", [("Invalid block definition", 1, 6)]),
    ("= 1\n", [("Argument or block definition required", 1, 1)]),
    ('b "l" = 1\n', [("Invalid block definition", 1, 7)]),
    ('b "${x}" {\n}\n', [("Invalid string literal", 1, 4)]),
    ("a = 1 @\n", [("Invalid character", 1, 7), ("Missing newline after argument", 1, 7)]),
])
def test_error_messages_and_positions(src, want):
    assert errors(src) == want


def test_recovery_continues_after_a_broken_item():
    body, diags, _ = parse('r "x" {\n  a = )\n  b = 2\n}\nc = 3\n}\nd = 4\n')
    assert [d.summary for d in diags] == ["Invalid expression", "Argument or block definition required"]
    assert [getattr(i, "name", getattr(i, "type", None)) for i in body.items] == ["r", "c", "d"]
    assert "b" in body.items[0].body.attributes


def test_error_inside_nested_brackets_does_not_corrupt_newline_state():
    body, diags, _ = parse('r {\n  a = f(1,\n    [2, }\n  b = 1\n}\nc = 2\n')
    assert diags
    # parsing went on and later, well-formed items were still seen
    assert body.items[-1].name == "c"


def test_absurd_nesting_is_a_diagnostic_not_a_crash():
    from nvidia_terraform_modules_b200.fmt.command import format_source
    src = "x = " + "[" * 3000 + "1" + "]" * 3000 + "\n"
    out, diags = format_source(src, "deep.tf")
    assert out == src and [d.summary for d in diags] == ["Expression nested too deeply"]
    ok = "x = " + "[" * 60 + "1" + "]" * 60 + "\n"
    assert format_source(ok, "ok.tf") == (ok, [])
