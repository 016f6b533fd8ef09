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

COMMON = settings(max_examples=300, deadline=None, suppress_health_check=[HealthCheck.too_slow])

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
