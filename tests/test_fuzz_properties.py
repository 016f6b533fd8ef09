"""Property-based tests (hypothesis): the front end never crashes and the
formatter's invariants hold on generated input, not only on hand-picked cases."""
import pytest

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import HealthCheck, given, settings, strategies as st  # noqa: E402

from nvidia_terraform_modules_b200.fmt.command import format_source  # noqa: E402
from nvidia_terraform_modules_b200.hcl.diagnostics import has_errors  # noqa: E402
from nvidia_terraform_modules_b200.hcl.parser import parse  # noqa: E402
from nvidia_terraform_modules_b200.hcl.scanner import scan  # noqa: E402
from nvidia_terraform_modules_b200.hcl.tokens import T, render  # noqa: E402

import os  # noqa: E402

# The suite the driver runs must give the same answer every time, so by default the examples
# are derived from the test itself (derandomize).  Hunting for new failures is a separate,
# deliberate act:   FUZZ_EXAMPLES=5000 python -m pytest tests/test_fuzz_properties.py
# draws fresh random examples on every run.
_LONG = "FUZZ_EXAMPLES" in os.environ
COMMON = settings(max_examples=int(os.environ.get("FUZZ_EXAMPLES", "300")), deadline=None, derandomize=not _LONG,
                  database=None,
                  suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])

# text biased towards characters that mean something to the scanner
ALPHABET = st.sampled_from(list("abxyz_-019 \t\n\r\"'\\${}%[](),.=<>!&|?:*/+#~") + ["${", "%{", "<<EOT\n", "EOT\n", "/*",
                                                                                  "*/", "//", "é", "for", "in", "if"])
NOISE = st.lists(ALPHABET, max_size=60).map("".join)


@COMMON
@given(NOISE)
def test_scanner_accepts_anything_and_loses_nothing(src):
    toks = scan(src)
    assert toks[-1].type is T.EOF
    assert render(toks, keep_raw=True) == src.lstrip("\ufeff")


@COMMON
@given(NOISE)
def test_parser_reports_instead_of_raising(src):
    body, diags, toks = parse(src, "fuzz.tf")
    assert body is not None
    for d in diags:
        assert d.severity == "error" and d.line >= 1


@COMMON
@given(NOISE)
def test_fmt_never_corrupts_unparseable_input(src):
    out, diags = format_source(src, "fuzz.tf")
    if has_errors(diags):
        assert out == src


# ------------------------------------------------------------ valid programs
IDENT = st.from_regex(r"[a-z][a-z0-9_]{0,7}", fullmatch=True).filter(lambda s: s not in ("for", "in", "if", "null",
                                                                                         "true", "false"))
WS = st.sampled_from(["", " ", "  ", "\t"])


def exprs():
    leaf = st.one_of(
        st.integers(0, 999).map(str),
        st.sampled_from(["true", "false", "null"]),
        IDENT.map(lambda n: f"var.{n}"),
        st.tuples(IDENT, IDENT).map(lambda t: f"{t[0]}.{t[1]}[0].id"),
        IDENT.map(lambda n: f'"lit-{n}"'),
        st.tuples(IDENT, IDENT).map(lambda t: f'"a${{var.{t[0]}}}b${{ local.{t[1]} }}"'),
    )

    def grow(inner):
        return st.one_of(
            # operators keep at least one blank: "a-1" is ONE identifier in HCL (names may hold dashes)
            st.tuples(inner, st.sampled_from(["+", "-", "*", "==", "&&", "||", "<="]), inner,
                      st.sampled_from([" ", "  ", "\t"])).map(lambda t: f"{t[0]}{t[3]}{t[1]}{t[3]}{t[2]}"),
            st.tuples(inner, inner, inner).map(lambda t: f"{t[0]} ? {t[1]} : {t[2]}"),
            inner.map(lambda e: f"({e})"),
            inner.map(lambda e: f"!{e}"),
            inner.map(lambda e: f"-{e}"),
            st.lists(inner, max_size=3).map(lambda xs: "[" + ",".join(xs) + "]"),
            st.tuples(IDENT, st.lists(inner, max_size=3), WS).map(
                lambda t: f"{t[0]}({t[2]}" + f",{t[2]}".join(t[1]) + ")"),
            st.lists(st.tuples(IDENT, inner), max_size=3, unique_by=lambda kv: kv[0]).map(
                lambda kvs: "{" + ", ".join(f"{k} = {v}" for k, v in kvs) + "}"),
            st.tuples(IDENT, inner, inner).map(lambda t: f"[for {t[0]} in {t[1]} : {t[2]}]"),
        )
    return st.recursive(leaf, grow, max_leaves=8)


def bodies(depth=2):
    attr = st.tuples(IDENT, WS, WS, exprs(), st.sampled_from(["", " # c", " // c"])).map(
        lambda t: f"{t[1]}{t[0]}{t[2]}={t[2]}{t[3]}{t[4]}\n")
    blank = st.just("\n")
    comment = st.sampled_from(["# note\n", "  // note\n", "/* block */\n"])
    items = [attr, attr, blank, comment]
    if depth:
        block = st.tuples(IDENT, st.lists(IDENT, max_size=2), WS, bodies(depth - 1)).map(
            lambda t: f"{t[2]}{t[0]} " + "".join(f'"{l}" ' for l in t[1]) + "{\n" + t[3] + f"{t[2]}}}\n")
        items.append(block)
    return st.lists(st.one_of(*items), max_size=6).map("".join)


PROGRAM = bodies()


def signature(text):
    return [(t.type, t.text) for t in scan(text) if t.type is not T.EOF]


@COMMON
@given(PROGRAM)
def test_generated_programs_parse(src):
    _, diags, _ = parse(src, "gen.tf")
    assert not diags, (src, [d.format() for d in diags])


@COMMON
@given(PROGRAM)
def test_fmt_is_idempotent_and_only_moves_white_space(src):
    once, diags = format_source(src, "gen.tf")
    assert not diags
    assert signature(once) == signature(src)
    twice, _ = format_source(once, "gen.tf")
    assert twice == once


@COMMON
@given(PROGRAM)
def test_fmt_output_obeys_the_visible_rules(src):
    out, _ = format_source(src, "gen.tf")
    for line in out.split("\n"):
        assert "\t" not in line
        assert line == line.rstrip(" "), repr(line)
        indent = len(line) - len(line.lstrip(" "))
        if line.strip() and not line.lstrip().startswith(("*", "/*")):
            assert indent % 2 == 0, repr(line)


# ------------------------------------------- layout independence, validate robustness
@COMMON
@given(PROGRAM, st.integers(0, 10 ** 6))
def test_fmt_answer_does_not_depend_on_the_layout_it_starts_from(src, seed):
    """Format, destroy the white space two ways (tests/scramble.py), format again: the
    same text must come back.  The generated-program form of the recovery test that
    tests/test_reference_evidence.py runs on the reference's own files."""
    import random
    from scramble import damage_lines, damage_tokens
    want, diags = format_source(src, "gen.tf")
    assert not diags
    for fn in (damage_lines, damage_tokens):
        bad = fn(want, random.Random(seed))
        assert signature(bad) == signature(want)
        got, diags = format_source(bad, "gen.tf")
        assert not diags and got == want, (fn.__name__, bad)


RELEASES = st.sampled_from(["0.13", "0.14", "0.15", "1.0", "1.2", "1.3", "1.5", "1.8", "1.9", "1.10"])
TOP_LEVEL = st.sampled_from(['variable "v" ', 'output "o" ', "locals ", 'resource "a_b" "c" ', 'data "a_b" "c" ',
                             'module "m" ', "terraform ", 'provider "p" ', "moved ", "import ", 'check "k" ',
                             "removed ", 'ephemeral "a_b" "e" ', 'variable "w" ', 'resource "a_b" "d" '])
INNER_NAME = st.sampled_from(["type", "default", "value", "count", "for_each", "source", "depends_on", "condition",
                              "error_message", "sensitive", "nullable", "from", "to", "id", "provider", "version",
                              "required_version", "experiments", "description", "x", "when", "on_failure", "iterator",
                              "labels", "ignore_changes", "ephemeral", "validation", "lifecycle"])
INNER_BLOCK = st.sampled_from(["validation", "lifecycle", "precondition", "postcondition", "provisioner \"p\"",
                               "connection", "dynamic \"d\"", "content", "assert", "backend \"b\"",
                               "required_providers", "data \"a_b\" \"n\"", "cloud", "nested"])


def module_bodies(depth=2):
    attr = st.tuples(INNER_NAME, exprs()).map(lambda t: f"  {t[0]} = {t[1]}\n")
    items = [attr, attr]
    if depth:
        items.append(st.tuples(INNER_BLOCK, module_bodies(depth - 1)).map(lambda t: f"  {t[0]} {{\n{t[1]}  }}\n"))
    return st.lists(st.one_of(*items), max_size=5).map("".join)


_DIRS: dict = {}


def _one_dir(factory, name):
    """One directory per property, rewritten for every example (thousands of examples
    must not leave thousands of directories behind)."""
    if name not in _DIRS:
        _DIRS[name] = factory.mktemp(name)
    target = _DIRS[name] / "main.tf"
    if target.exists():
        target.unlink()          # (truncating in place is very slow on some container file systems)
    return _DIRS[name]


MODULE = st.lists(st.tuples(TOP_LEVEL, module_bodies()).map(lambda t: f"{t[0]}{{\n{t[1]}}}\n"), max_size=6).map("".join)


@COMMON
@given(MODULE, RELEASES)
def test_validate_reports_instead_of_raising(tmp_path_factory, src, release):
    """Structurally plausible nonsense under every restated release: diagnostics, never a
    traceback, and every diagnostic carries a summary, a file and a line."""
    from nvidia_terraform_modules_b200.check import clear_caches
    from nvidia_terraform_modules_b200.config.validate import validate_module
    from nvidia_terraform_modules_b200.lint.rules import run_rules
    d = _one_dir(tmp_path_factory, "fz")
    (d / "main.tf").write_text(src)
    clear_caches()
    res = validate_module(str(d), str(d), release)
    assert res.module.restated.text == release
    for diag in res.diags:
        assert diag.summary and diag.file == "main.tf" and diag.line >= 0
    run_rules(res)            # the lint rules walk the same tree: they must not raise either
    from nvidia_terraform_modules_b200.docs.tfdocs import inject, parse_readme, render
    tables = render(res.module)
    assert inject(inject("# T\n", tables), tables) == inject("# T\n", tables)
    assert set(parse_readme(tables)) <= {"Requirements", "Providers", "Modules", "Resources", "Inputs", "Outputs"}


@COMMON
@given(NOISE, RELEASES)
def test_validate_survives_noise(tmp_path_factory, src, release):
    from nvidia_terraform_modules_b200.check import clear_caches
    from nvidia_terraform_modules_b200.config.validate import validate_module
    d = _one_dir(tmp_path_factory, "fn")
    (d / "main.tf").write_text(src, encoding="utf-8", newline="")
    clear_caches()
    validate_module(str(d), str(d), release)


# ---------------------------------------------------- richer, Terraform-shaped expressions
REFS = st.sampled_from([
    "var.v", "var.w", "var.nope", "local.x", "local.y", "a_b.c", "a_b.c.id", "a_b.c[0].id", "a_b.d[\"k\"].id",
    "a_b.c[*].id", "a_b.c.*.id", "data.a_b.c.id", "data.a_b.c[0]", "module.m", "module.m.out", "module.m[0].out",
    "module.m[\"k\"]", "ephemeral.a_b.e.v", "count.index", "each.key", "each.value.x", "self.id", "path.module",
    "path.nope", "terraform.workspace", "terraform.applying", "var", "local", "module", "data.a_b", "a_b", "each",
    "x.value", "x.key", "null", "true", "12", "\"s\"", "\"a${var.v}b\"", "\"${local.x}\"", "<<EOT\nh ${var.v}\nEOT\n",
])
FUNCS = st.sampled_from(["upper", "length", "merge", "one", "try", "can", "lookup", "format", "list", "map", "nosuch",
                         "startswith", "provider::p::f", "core::upper", "templatestring", "alltrue", "sensitive"])


def rich_exprs():
    def grow(inner):
        return st.one_of(
            st.tuples(FUNCS, st.lists(inner, max_size=3), st.booleans()).map(
                lambda t: f"{t[0]}({', '.join(t[1])}{'...' if t[2] and t[1] else ''})"),
            st.tuples(inner, inner, inner).map(lambda t: f"{t[0]} ? {t[1]} : {t[2]}"),
            st.tuples(inner, st.sampled_from(["+", "-", "*", "/", "%", "==", "!=", "<", ">=", "&&", "||"]), inner).map(
                lambda t: f"{t[0]} {t[1]} {t[2]}"),
            inner.map(lambda e: f"({e})"), inner.map(lambda e: f"!{e}"), inner.map(lambda e: f"-{e}"),
            st.lists(inner, max_size=3).map(lambda xs: "[" + ", ".join(xs) + "]"),
            st.lists(st.tuples(st.sampled_from(["a", "b", "\"q\"", "(var.v)", "a.b", "true"]), inner), max_size=3).map(
                lambda kvs: "{ " + ", ".join(f"{k} = {v}" for k, v in kvs) + " }" if kvs else "{}"),
            st.tuples(inner, inner).map(lambda t: f"[for x in {t[0]} : {t[1]}]"),
            st.tuples(inner, inner, inner).map(lambda t: f"{{ for k, x in {t[0]} : {t[1]} => {t[2]} }}"),
            st.tuples(inner, inner).map(lambda t: f"{t[0]}[{t[1]}]"),
            inner.map(lambda e: f"{e}.attr"), inner.map(lambda e: f"{e}[*].attr"),
        )
    return st.recursive(REFS, grow, max_leaves=6)


def rich_bodies(depth=2):
    attr = st.tuples(INNER_NAME, rich_exprs()).map(lambda t: f"  {t[0]} = {t[1]}\n")
    items = [attr, attr, attr]
    if depth:
        items.append(st.tuples(INNER_BLOCK, rich_bodies(depth - 1)).map(lambda t: f"  {t[0]} {{\n{t[1]}  }}\n"))
    return st.lists(st.one_of(*items), max_size=5).map("".join)


RICH_MODULE = st.lists(st.tuples(TOP_LEVEL, rich_bodies()).map(lambda t: f"{t[0]}{{\n{t[1]}}}\n"), max_size=6).map("".join)


@COMMON
@given(RICH_MODULE, RELEASES)
def test_validate_lint_docs_and_fmt_survive_terraform_shaped_modules(tmp_path_factory, src, release):
    from nvidia_terraform_modules_b200.check import clear_caches
    from nvidia_terraform_modules_b200.config.validate import validate_module
    from nvidia_terraform_modules_b200.docs.tfdocs import render
    from nvidia_terraform_modules_b200.lint.rules import run_rules
    d = _one_dir(tmp_path_factory, "rich")
    (d / "main.tf").write_text(src)
    clear_caches()
    res = validate_module(str(d), str(d), release)
    for diag in res.diags:
        assert diag.summary and diag.line >= 0
    run_rules(res)
    render(res.module)
    out, diags = format_source(src, "main.tf")
    if not has_errors(diags):
        assert format_source(out, "main.tf")[0] == out or "${" in src      # (nested "${"${x}"}" unwraps twice)
