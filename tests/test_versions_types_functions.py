"""Version constraints, type constraints and the built-in function table."""
import pytest

from nvidia_terraform_modules_b200.config import functions, typeexpr
from nvidia_terraform_modules_b200.config.versions import Constraints, Version, VersionError
from nvidia_terraform_modules_b200.hcl.parser import parse


def V(s):
    return Version.parse(s)


# ------------------------------------------------------------------ versions
def test_version_parse_and_order():
    assert V("1.2.3") < V("1.10.0") < V("2.0")
    assert V("1.0.0-beta1") < V("1.0.0")
    assert V("v1.2") == V("1.2.0")
    assert str(V("4.45")) == "4.45.0"
    with pytest.raises(VersionError):
        V("one.two")


@pytest.mark.parametrize("constraint,version,ok", [
    ("~>4.45.0", "4.45.0", True), ("~>4.45.0", "4.45.9", True), ("~>4.45.0", "4.46.0", False),
    ("~> 2.19.0", "2.18.9", False),
    ("~> 1.2", "1.9.0", True), ("~> 1.2", "2.0.0", False), ("~> 1.2", "1.1.0", False),
    (">= 1.2.4", "1.2.4", True), (">= 1.2.4", "1.2.3", False), (">= 0.14", "1.5.0", True),
    ("4.27.0", "4.27.0", True), ("= 4.27.0", "4.27.1", False),
    (">= 3.72.0, >= 4.0.0, ~> 4.45.0", "4.45.0", True), (">= 3.72.0, < 4.0.0", "4.45.0", False),
    ("!= 1.0.0", "1.0.0", False), ("> 1.0", "1.0.1", True), ("<= 1.0", "1.0.0", True),
    (">= 1.0", "1.1.0-rc1", False),          # pre-releases only match exactly
    ("1.1.0-rc1", "1.1.0-rc1", True),
])
def test_constraints_allow(constraint, version, ok):
    assert Constraints.parse(constraint).allows(V(version)) is ok


def test_constraint_spelling_as_in_lock_files():
    assert Constraints.parse("~>4.45.0").canonical() == "~> 4.45.0"
    assert Constraints.parse("4.27.0").canonical() == "4.27.0"
    assert Constraints.parse("~> 2.19.0,>=2.10.0").canonical() == ">= 2.10.0, ~> 2.19.0"
    assert Constraints.parse("4.27.0").is_exact() and not Constraints.parse("~> 4.0").is_exact()


@pytest.mark.parametrize("bad", ["", "  ", "~> x", "=> 1.0", "1.0 2.0", ">= 1.0,"])
def test_malformed_constraints(bad):
    with pytest.raises(VersionError):
        Constraints.parse(bad)


# --------------------------------------------------------------------- types
def ty(text):
    body, diags, _ = parse(f"type = {text}\n")
    assert not diags
    return typeexpr.parse_type(body.items[0].expr)


def val(text):
    body, diags, _ = parse(f"v = {text}\n")
    assert not diags
    return typeexpr.literal_value(body.items[0].expr)


def test_type_expressions():
    assert str(ty("string")) == "string"
    assert str(ty("list(any)")) == "list(any)"
    assert str(ty("map(list(number))")) == "map(list(number))"
    assert str(ty("list")) == "list(any)"
    t = ty("object({\n  vpc_id = string\n  subnet_ids = list(string)\n  note = optional(string, \"x\")\n})")
    assert set(t.attrs) == {"vpc_id", "subnet_ids", "note"} and t.optional == {"note"}
    assert str(ty("tuple([string, number])")) == "tuple([string, number])"


@pytest.mark.parametrize("text", ['"string"', "strng", "list()", "object(string)", "tuple(string)",
                                  "optional(string)", "frob(string)", "1"])
def test_invalid_type_expressions(text):
    body, _, _ = parse(f"type = {text}\n")
    with pytest.raises(typeexpr.TypeError_):
        typeexpr.parse_type(body.items[0].expr)


def test_literal_values():
    assert val('{ a = [1, "x", true, null], "b" : -2 }') == {"a": [1, "x", True, None], "b": -2}
    assert val("var.x") is typeexpr.UNKNOWN
    assert val('"a${b}"') is typeexpr.UNKNOWN
    assert val("[var.x]") == [typeexpr.UNKNOWN]


@pytest.mark.parametrize("type_text,value_text,ok", [
    ("string", '"a"', True), ("string", "1", True), ("string", "true", True), ("string", "[]", False),
    ("number", "1.5", True), ("number", '"12"', True), ("number", '"x"', False), ("number", "true", False),
    ("bool", "false", True), ("bool", '"true"', True), ("bool", '"yes"', False), ("bool", "1", False),
    ("list(string)", '["a", 1]', True), ("list(number)", '["a"]', False), ("list(any)", "{}", False),
    ("set(string)", "[]", True),
    ("map(string)", '{ a = "b" }', True), ("map(number)", '{ a = "b" }', False), ("map(any)", "[]", False),
    ("object({ a = string, b = optional(number) })", '{ a = "x" }', True),
    ("object({ a = string })", "{}", False),
    ("object({ a = string })", '{ a = "x", extra = 1 }', True),
    ("tuple([string, number])", '["a", 1]', True), ("tuple([string])", '["a", 1]', False),
    ("any", "[1, { a = 2 }]", True),
    ("string", "null", True), ("list(string)", "null", True),
    ("number", "var.x", True),          # not a literal: cannot be judged, so no complaint
])
def test_default_conformance(type_text, value_text, ok):
    err = typeexpr.conformance_error(val(value_text), ty(type_text))
    assert (err is None) is ok, err


def test_conformance_error_names_the_path():
    err = typeexpr.conformance_error(val('{ a = { b = ["x", []] } }'), ty("object({ a = map(list(string)) })"))
    assert err.startswith('attribute "a"["b"][1]: string required')


# ----------------------------------------------------------------- functions
def test_reference_functions_are_known():
    for name in ("merge", "concat", "length", "one", "jsonencode", "jsondecode", "base64decode"):
        assert functions.lookup(name) is not None
    assert functions.lookup("core::length") == functions.lookup("length")
    assert functions.lookup("uppercase") is None


@pytest.mark.parametrize("name,n,expanded,summary", [
    ("length", 1, False, None), ("length", 0, False, "Not enough function arguments"),
    ("length", 2, False, "Too many function arguments"), ("merge", 0, False, None), ("merge", 9, False, None),
    ("lookup", 2, False, None), ("lookup", 3, False, None), ("lookup", 1, False, "Not enough function arguments"),
    ("format", 0, False, "Not enough function arguments"), ("replace", 1, True, None),
    ("timestamp", 1, False, "Too many function arguments"),
])
def test_arity(name, n, expanded, summary):
    err = functions.arity_error(name, n, expanded)
    assert (err[0] if err else None) == summary
