"""Canonical white-space layout for HCL2 token streams.

What ``terraform fmt`` enforces (the reference's stated gate, CONTRIBUTING.md:12;
BASELINE.json config 1) is implemented upstream in hashicorp/hcl v2,
``hclwrite/format.go`` -- a dependency of the Terraform binary that is NOT under
/root/reference and is not installed in this image.  This module restates that
published algorithm.  PARITY UNPINNED: there is no ``terraform`` binary here to
run beside it and the reference ships no golden files; see DESIGN.md.

The algorithm only ever changes ``spaces_before`` of tokens.  It runs in three
passes over "format lines" (runs of tokens ending in a newline token or in a
``#``/``//`` comment, which swallows its newline):

1. indentation  -- two spaces per unclosed bracket group, counted per line;
2. spacing      -- one rule table deciding 0 or 1 space between neighbours;
3. alignment    -- each line is split into lead / assign / comment cells;
                   ``=`` signs of consecutive single-line attributes are aligned,
                   and so are trailing ``#`` comments.
"""
from __future__ import annotations

from dataclasses import dataclass, field

from ..hcl.tokens import NIL_TOKEN, T, Token


@dataclass
class FormatLine:
    lead: list[Token] = field(default_factory=list)
    assign: list[Token] = field(default_factory=list)
    comment: list[Token] = field(default_factory=list)


def columns(tokens: list[Token]) -> int:
    return sum(t.spaces_before + len(t.text) for t in tokens)


# --------------------------------------------------------------------------
# line splitting
# --------------------------------------------------------------------------
def lines_for_format(tokens: list[Token]) -> list[FormatLine]:
    lines: list[FormatLine] = []
    current: list[Token] = []
    for tok in tokens:
        if tok.type is T.EOF:
            break
        current.append(tok)
        if tok.is_line_end:
            lines.append(FormatLine(lead=current))
            current = []
    lines.append(FormatLine(lead=current))

    in_heredoc = False
    for line in lines:
        if not line.lead:
            continue
        if in_heredoc:
            # everything on a heredoc continuation line stays in "lead"
            if any(t.type is T.CHEREDOC for t in line.lead):
                in_heredoc = False
            continue
        opened = False
        for t in line.lead:
            if t.type is T.OHEREDOC:
                opened = True
            elif t.type is T.CHEREDOC:
                opened = False
        in_heredoc = opened

        if len(line.lead) > 1 and line.lead[-1].type is T.COMMENT:
            line.comment = line.lead[-1:]
            line.lead = line.lead[:-1]

        for i, tok in enumerate(line.lead):
            if i > 0 and tok.type is T.EQUAL:
                # Only a right-hand side that is a whole expression on this
                # line is moved to the "assign" cell; an unclosed bracket means
                # a multi-line expression starts here.
                if sum(t.bracket_change for t in line.lead[i:]) == 0:
                    line.assign = line.lead[i:]
                    line.lead = line.lead[:i]
                break
    return lines


# --------------------------------------------------------------------------
# pass 1: indentation
# --------------------------------------------------------------------------
def format_indent(lines: list[FormatLine]) -> None:
    indents: list[int] = []   # brackets opened by each line that deepened the nesting
    in_heredoc = False
    for line in lines:
        if not line.lead:
            continue
        if in_heredoc:
            if any(t.type is T.CHEREDOC for t in line.lead):
                in_heredoc = False
            continue            # heredoc text is never re-indented
        if line.lead[0].type is T.NEWLINE:
            line.lead[0].spaces_before = 0   # blank line: no trailing blanks
            continue

        net = 0
        for tok in line.lead + line.assign:
            net += tok.bracket_change
            if tok.type is T.OHEREDOC:
                in_heredoc = True
            elif tok.type is T.CHEREDOC:
                in_heredoc = False

        if net > 0:
            line.lead[0].spaces_before = 2 * len(indents)
            indents.append(net)
        elif net < 0:
            closed = -net
            while closed > 0 and indents:
                top = indents[-1]
                if closed > top:
                    closed -= top
                    indents.pop()
                elif closed < top:
                    indents[-1] = top - closed
                    closed = 0
                else:
                    indents.pop()
                    closed = 0
            line.lead[0].spaces_before = 2 * len(indents)
        else:
            line.lead[0].spaces_before = 2 * len(indents)


# --------------------------------------------------------------------------
# pass 2: spacing between neighbouring tokens
# --------------------------------------------------------------------------
_STRINGISH_SUBJECT = {T.QUOTED_LIT, T.STRING_LIT, T.OQUOTE, T.OHEREDOC}
_STRINGISH_AFTER = {T.QUOTED_LIT, T.STRING_LIT, T.CQUOTE, T.CHEREDOC}
_NEGATION_CONTEXT = {
    T.NIL,
    T.OPAREN, T.OBRACE, T.OBRACK, T.EQUAL, T.COLON, T.COMMA, T.QUESTION,
    T.PLUS, T.STAR, T.SLASH, T.PERCENT, T.MINUS,
    T.EQUAL_OP, T.NOT_EQUAL, T.GREATER_THAN, T.GREATER_THAN_EQ, T.LESS_THAN, T.LESS_THAN_EQ,
    T.AND, T.OR, T.BANG,
}


def space_after(subject: Token, before: Token, after: Token) -> bool:
    """Should one space separate ``subject`` from ``after``?  ``before`` is the
    token preceding ``subject`` in the same cell (NIL at the start)."""
    s, a = subject.type, after.type

    if a is T.NEWLINE or a is T.NIL:
        return False                      # never pad before a newline
    if s is T.IDENT and a is T.OPAREN:
        return False                      # function call: name(
    if (s is T.IDENT and a is T.DOUBLE_COLON) or (s is T.DOUBLE_COLON and a is T.IDENT):
        return False                      # provider::ns::fn
    if s is T.DOT or a is T.DOT:
        return False                      # attribute access
    if a is T.COMMA or a is T.ELLIPSIS:
        return False
    if s is T.COMMA:
        return True
    if s in _STRINGISH_SUBJECT or a in _STRINGISH_AFTER:
        return False                      # nothing is inserted inside templates
    if s is T.IDENT and subject.text == "in" and before.type is T.IDENT:
        return True                       # [for x in [a, b] : x]
    if a is T.OBRACK and (s is T.IDENT or s is T.NUMBER_LIT or subject.bracket_change < 0):
        return False                      # index: foo[0], f()[0], a[0][1]
    if s is T.MINUS:
        # subtraction keeps its space, negation does not; guess from what
        # precedes the minus sign.
        return before.type not in _NEGATION_CONTEXT
    if s is T.BANG:
        return False
    if s is T.OBRACE or a is T.CBRACE:
        # braces are padded on the inside -- { a = 1 } -- except when empty
        return not (s is T.OBRACE and a is T.CBRACE)
    if s in (T.TEMPLATE_INTERP, T.TEMPLATE_CONTROL) and a is T.OBRACE:
        return True                       # "${ { a = 1 } }" stays readable
    if s is T.CBRACE and a is T.TEMPLATE_SEQ_END:
        return True
    if s is T.TEMPLATE_SEQ_END and a in (T.TEMPLATE_INTERP, T.TEMPLATE_CONTROL):
        return False
    if subject.bracket_change > 0:
        return False                      # no space after an opener
    if after.bracket_change < 0:
        return False                      # no space before a closer
    return True


def _space_cell(cell: list[Token]) -> None:
    for i, tok in enumerate(cell):
        if i + 1 >= len(cell):
            break
        before = cell[i - 1] if i > 0 else NIL_TOKEN
        after = cell[i + 1]
        after.spaces_before = 1 if space_after(tok, before, after) else 0


def format_spaces(lines: list[FormatLine]) -> None:
    for line in lines:
        _space_cell(line.lead)
        if line.assign:
            line.assign[0].spaces_before = 1
            _space_cell(line.assign)
        if line.comment:
            line.comment[0].spaces_before = 1


# --------------------------------------------------------------------------
# pass 3: vertical alignment of "=" and of trailing comments
# --------------------------------------------------------------------------
def format_cells(lines: list[FormatLine]) -> None:
    def align(cell_of, width_of) -> None:
        chain: list[FormatLine] = []

        def close() -> None:
            if chain:
                widest = max(width_of(l) for l in chain)
                for l in chain:
                    cell_of(l)[0].spaces_before = widest - width_of(l) + 1
                chain.clear()

        for line in lines:
            if cell_of(line):
                chain.append(line)
            else:
                close()
        close()

    align(lambda l: l.assign, lambda l: columns(l.lead))
    align(lambda l: l.comment, lambda l: columns(l.lead) + columns(l.assign))


# --------------------------------------------------------------------------
def format_tokens(tokens: list[Token]) -> list[Token]:
    """Lay the token stream out canonically, in place; returns it."""
    lines = lines_for_format(tokens)
    format_indent(lines)
    format_spaces(lines)
    format_cells(lines)
    return tokens
