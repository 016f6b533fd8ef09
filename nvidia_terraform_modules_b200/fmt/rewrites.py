"""Terraform-level source rewrites that ``terraform fmt`` applies on top of the
generic HCL layout pass.

Upstream these live in Terraform core's ``fmt`` command (not part of
/root/reference; the reference merely requires contributors to run it,
CONTRIBUTING.md:12).  Restated here -- PARITY UNPINNED, see DESIGN.md:

* an attribute whose whole value is a single interpolation, ``"${expr}"``, is
  unwrapped to ``expr`` (0.11-style syntax; the reference has none, but the
  check must not miss it in a contribution);
* ``variable`` blocks: legacy quoted/bare type keywords become type
  expressions (``"string"`` -> ``string``, ``"list"`` -> ``list(string)``,
  ``"map"`` -> ``map(string)``, bare ``list`` -> ``list(any)``, ``map`` -> ``map(any)``);
* block labels are written as quoted strings (``resource aws_x y {`` ->
  ``resource "aws_x" "y" {``).

All three work on token sub-ranges recorded by the parser and return a new
token list; white space is then fixed by ``canonical.format_tokens``.
"""
from __future__ import annotations

from ..hcl import ast as A
from ..hcl.tokens import T, Token


def _tok(typ: T, text: str, spaces: int = 0) -> Token:
    return Token(typ, text, spaces_before=spaces, raw_before=" " * spaces)


def _quote_label(value: str) -> list[Token]:
    esc = (value.replace("\\", "\\\\").replace('"', '\\"').replace("\n", "\\n")
           .replace("\r", "\\r").replace("\t", "\\t").replace("${", "$${").replace("%{", "%%{"))
    toks = [_tok(T.OQUOTE, '"', 1)]
    if esc:
        toks.append(_tok(T.QUOTED_LIT, esc))
    toks.append(_tok(T.CQUOTE, '"'))
    return toks


def _trim_newlines(tokens: list[Token]) -> list[Token]:
    start, end = 0, len(tokens)
    while start < end and tokens[start].type is T.NEWLINE:
        start += 1
    while end > start and tokens[end - 1].type is T.NEWLINE:
        end -= 1
    return tokens[start:end]


def unwrap_value_expr(tokens: list[Token]) -> list[Token]:
    """``"${ expr }"`` -> ``expr`` when that is provably the whole value."""
    if len(tokens) < 5:
        return tokens
    if (tokens[0].type is not T.OQUOTE or tokens[1].type is not T.TEMPLATE_INTERP
            or tokens[-2].type is not T.TEMPLATE_SEQ_END or tokens[-1].type is not T.CQUOTE):
        return tokens
    inside = tokens[2:-2]
    quotes = 0
    for tok in inside:
        if tok.type is T.OQUOTE:
            quotes += 1
            continue
        if tok.type is T.CQUOTE:
            quotes -= 1
            continue
        if quotes > 0:
            continue          # nested string inside the expression is fine
        if tok.type in (T.TEMPLATE_INTERP, T.TEMPLATE_SEQ_END):
            return tokens     # "${a}${b}" is not a single interpolation
        if tok.type is T.QUOTED_LIT:
            return tokens     # literal text beside the interpolation
    trimmed = _trim_newlines(inside)
    if not trimmed:
        return tokens
    multi_line = any(t.type is T.NEWLINE for t in trimmed)
    parenthesised = trimmed[0].type is T.OPAREN and trimmed[-1].type is T.CPAREN
    if multi_line and not parenthesised:
        return [_tok(T.OPAREN, "(")] + trimmed + [_tok(T.CPAREN, ")")]
    return trimmed


def modernise_type_expr(tokens: list[Token]) -> list[Token]:
    if len(tokens) == 1 and tokens[0].type is T.IDENT:
        kw = tokens[0].text
        if kw in ("list", "map"):
            return [_tok(T.IDENT, kw), _tok(T.OPAREN, "("), _tok(T.IDENT, "any"), _tok(T.CPAREN, ")")]
        return tokens
    if (len(tokens) == 3 and tokens[0].type is T.OQUOTE and tokens[1].type is T.QUOTED_LIT
            and tokens[2].type is T.CQUOTE):
        kw = tokens[1].text
        if kw == "string":
            return [_tok(T.IDENT, "string")]
        if kw in ("list", "map"):
            return [_tok(T.IDENT, kw), _tok(T.OPAREN, "("), _tok(T.IDENT, "string"), _tok(T.CPAREN, ")")]
    return tokens


def collect_rewrites(body: A.Body, tokens: list[Token]) -> dict[tuple[int, int], list[Token]]:
    """Map of token ranges -> replacement tokens for a parsed file."""
    edits: dict[tuple[int, int], list[Token]] = {}

    def visit(b: A.Body, in_blocks: tuple[str, ...]) -> None:
        for item in b.items:
            if isinstance(item, A.Attribute):
                rng = (item.expr_tok_start, item.expr_tok_end)
                old = tokens[rng[0]:rng[1]]
                if in_blocks == ("variable",) and item.name == "type":
                    new = modernise_type_expr(old)
                else:
                    new = unwrap_value_expr(old)
                if new is not old and new != old:
                    new = list(new)
                    # the value keeps the gap that separated it from '='
                    first = Token(new[0].type, new[0].text, new[0].line, new[0].col, new[0].offset,
                                  old[0].spaces_before, old[0].raw_before)
                    edits[rng] = [first] + new[1:]
            else:
                for lab in item.labels:
                    if not lab.quoted:
                        edits[(lab.tok_start, lab.tok_end)] = _quote_label(lab.value)
                visit(item.body, in_blocks + (item.type,))

    visit(body, ())
    return edits


def apply_rewrites(tokens: list[Token], edits: dict[tuple[int, int], list[Token]]) -> list[Token]:
    if not edits:
        return tokens
    out: list[Token] = []
    i = 0
    for (start, end) in sorted(edits):
        out.extend(tokens[i:start])
        out.extend(edits[(start, end)])
        i = end
    out.extend(tokens[i:])
    return out
