"""Token model for HCL2 native syntax.

The reference (/root/reference) is written in HCL2 native syntax (32 ``.tf``,
6 ``.tfvars``, 6 ``.hcl`` files).  The tool that gives those
files meaning -- Terraform core and the hashicorp/hcl v2 library it embeds -- is
NOT part of the reference tree.  The token vocabulary below restates the one the
HCL2 syntax specification defines, because the canonical-format algorithm
(``fmt/canonical.py``) is specified in terms of these token kinds.

Every token remembers the horizontal white space that preceded it
(``spaces_before``) so that writing the tokens back out reproduces the source
byte for byte when nothing was changed.
"""
from __future__ import annotations

from dataclasses import dataclass
from enum import Enum


class T(str, Enum):
    """Token kinds (names follow the HCL2 syntax specification)."""

    OBRACE = "{"
    CBRACE = "}"
    OBRACK = "["
    CBRACK = "]"
    OPAREN = "("
    CPAREN = ")"
    OQUOTE = 'o"'
    CQUOTE = 'c"'
    OHEREDOC = "<<"
    CHEREDOC = ">>"

    STAR = "*"
    SLASH = "/"
    PLUS = "+"
    MINUS = "-"
    PERCENT = "%"

    EQUAL = "="
    EQUAL_OP = "=="
    NOT_EQUAL = "!="
    LESS_THAN = "<"
    LESS_THAN_EQ = "<="
    GREATER_THAN = ">"
    GREATER_THAN_EQ = ">="

    AND = "&&"
    OR = "||"
    BANG = "!"

    DOT = "."
    COMMA = ","
    ELLIPSIS = "..."
    FAT_ARROW = "=>"
    QUESTION = "?"
    COLON = ":"
    DOUBLE_COLON = "::"

    TEMPLATE_INTERP = "${"
    TEMPLATE_CONTROL = "%{"
    TEMPLATE_SEQ_END = "}t"

    QUOTED_LIT = "qlit"   # literal run inside "..."
    STRING_LIT = "slit"   # literal run inside a heredoc
    NUMBER_LIT = "num"
    IDENT = "ident"

    COMMENT = "comment"
    NEWLINE = "nl"
    EOF = "eof"
    INVALID = "invalid"
    NIL = "nil"           # sentinel "no token" used by the spacing rules


_OPENERS = {T.OBRACE, T.OBRACK, T.OPAREN, T.TEMPLATE_CONTROL, T.TEMPLATE_INTERP}
_CLOSERS = {T.CBRACE, T.CBRACK, T.CPAREN, T.TEMPLATE_SEQ_END}


@dataclass
class Token:
    type: T
    text: str
    line: int = 0          # 1-based line of the first character
    col: int = 0           # 1-based column (in characters) of the first character
    offset: int = 0        # character offset of the first character
    spaces_before: int = 0  # horizontal white space columns preceding the token
    raw_before: str = ""   # the exact white space text (tabs are kept here)

    @property
    def bracket_change(self) -> int:
        if self.type in _OPENERS:
            return 1
        if self.type in _CLOSERS:
            return -1
        return 0

    @property
    def is_line_end(self) -> bool:
        """True for a newline token and for a ``#``/``//`` comment, whose text
        swallows its own terminating newline."""
        if self.type is T.NEWLINE:
            return True
        return self.type is T.COMMENT and self.text.endswith("\n")

    def __repr__(self) -> str:  # pragma: no cover - debugging aid
        return f"Token({self.type.name}, {self.text!r}, {self.line}:{self.col})"


NIL_TOKEN = Token(T.NIL, "")


def render(tokens: list[Token], *, keep_raw: bool = False) -> str:
    """Write tokens back to text.  With ``keep_raw`` the original white space
    (tabs included) is reproduced; otherwise ``spaces_before`` spaces are used."""
    out: list[str] = []
    for tok in tokens:
        if keep_raw:
            out.append(tok.raw_before)
        elif tok.spaces_before:
            out.append(" " * tok.spaces_before)
        out.append(tok.text)
    return "".join(out)
