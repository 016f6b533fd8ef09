"""Scanner for HCL2 native syntax.

Restates the lexical grammar of the HCL2 native syntax specification
(hashicorp/hcl v2, ``hclsyntax/spec.md``: "Lexical Elements", "Comments and
Whitespace", "Identifiers", "Numeric Literals", "Templates").  That library is a
third-party dependency of Terraform and is not present under /root/reference;
the reference only *uses* the language (e.g. templates at eks/main.tf:11,
heredocs at aks/variables.tf:99-103, ``/* */`` comments at eks/main.tf:4-6,
``//`` comments at aks/main.tf:51).

Three scanning modes nest inside each other:

* normal   -- expressions and body structure;
* quoted   -- between ``"`` and ``"``; literal runs plus ``${``/``%{`` sequences;
* heredoc  -- between ``<<EOT``/``<<-EOT`` and the closing marker line.

A ``${``/``%{`` sequence re-enters normal mode until the brace that balances it.
"""
from __future__ import annotations

import re
import unicodedata

from .tokens import T, Token

_TWO_CHAR = {
    "==": T.EQUAL_OP,
    "!=": T.NOT_EQUAL,
    "<=": T.LESS_THAN_EQ,
    ">=": T.GREATER_THAN_EQ,
    "&&": T.AND,
    "||": T.OR,
    "=>": T.FAT_ARROW,
    "::": T.DOUBLE_COLON,
}
_ONE_CHAR = {
    "{": T.OBRACE,
    "}": T.CBRACE,
    "[": T.OBRACK,
    "]": T.CBRACK,
    "(": T.OPAREN,
    ")": T.CPAREN,
    "*": T.STAR,
    "/": T.SLASH,
    "+": T.PLUS,
    "-": T.MINUS,
    "%": T.PERCENT,
    "=": T.EQUAL,
    "<": T.LESS_THAN,
    ">": T.GREATER_THAN,
    "!": T.BANG,
    ".": T.DOT,
    ",": T.COMMA,
    "?": T.QUESTION,
    ":": T.COLON,
}


_NORMAL_RE = re.compile(r"""
      (?P<ws>[ \t]+)
    | (?P<ident>[A-Za-z_][A-Za-z0-9_\-]*)
    | (?P<nl>\r?\n)
    | (?P<comment>(?:\#|//)[^\n]*\n? | /\*.*?\*/)
    | (?P<num>[0-9]+(?:\.[0-9]+)?(?:[eE][+-]?[0-9]+)?)
    | (?P<dots>\.\.\.)
    | (?P<op2>==|!=|<=|>=|&&|\|\||=>|::)
    | (?P<op1>[\[\]()*+\-%=>!.,?:] | /(?!\*))
""", re.VERBOSE | re.DOTALL)


def is_id_start(ch: str) -> bool:
    if ch == "_" or ("a" <= ch <= "z") or ("A" <= ch <= "Z"):
        return True
    if ch < "\x80":
        return False
    return unicodedata.category(ch) in ("Lu", "Ll", "Lt", "Lm", "Lo", "Nl")


def is_id_continue(ch: str) -> bool:
    if is_id_start(ch) or ("0" <= ch <= "9") or ch == "-":
        return True
    if ch < "\x80":
        return False
    return unicodedata.category(ch) in ("Mn", "Mc", "Nd", "Pc")


def is_valid_identifier(name: str) -> bool:
    return bool(name) and is_id_start(name[0]) and all(is_id_continue(c) for c in name[1:])


class Scanner:
    def __init__(self, src: str):
        if src.startswith("\ufeff"):
            src = src[1:]
        self.src = src
        self.n = len(src)
        self.pos = 0
        self.line = 1
        self.col = 1
        self.tokens: list[Token] = []
        self._pending_ws = ""

    # -- helpers -----------------------------------------------------------
    def _peek(self, k: int = 0) -> str:
        i = self.pos + k
        return self.src[i] if i < self.n else ""

    def _advance(self, text: str) -> None:
        if "\n" in text:
            self.line += text.count("\n")
            self.col = len(text) - text.rfind("\n")
        else:
            self.col += len(text)
        self.pos += len(text)

    def _emit(self, typ: T, text: str) -> Token:
        ws = self._pending_ws
        tok = Token(typ, text, self.line, self.col, self.pos, len(ws), ws)
        self._pending_ws = ""
        self.tokens.append(tok)
        self._advance(text)
        return tok

    def _skip_ws(self) -> None:
        start = self.pos
        while self._peek() in (" ", "\t") and self.pos < self.n:
            self.pos += 1
            self.col += 1
        self._pending_ws += self.src[start:self.pos]

    # -- entry point -------------------------------------------------------
    def scan(self) -> list[Token]:
        self._normal(in_template=False)
        self._emit(T.EOF, "")
        return self.tokens

    # -- normal mode -------------------------------------------------------
    def _normal(self, in_template: bool) -> None:
        """One master regular expression recognises the frequent tokens; the
        branches below it handle what needs state (strings, heredocs, the brace
        that ends a template sequence) or is rare (non-ASCII identifiers)."""
        depth = 0
        src, n = self.src, self.n
        match = _NORMAL_RE.match
        while True:
            m = match(src, self.pos)
            if m is not None:
                kind = m.lastgroup
                text = m.group(kind)
                if kind == "ws":
                    self._pending_ws += text
                    self.pos = m.end()
                    self.col += len(text)
                    continue
                if kind == "ident":
                    end = m.end()
                    if end < n and src[end] >= "\x80" and is_id_continue(src[end]):
                        while end < n and is_id_continue(src[end]):
                            end += 1
                        text = src[self.pos:end]
                    self._emit(T.IDENT, text)
                elif kind == "nl":
                    self._emit(T.NEWLINE, text)
                elif kind == "comment":
                    self._emit(T.COMMENT, text)
                elif kind == "num":
                    self._emit(T.NUMBER_LIT, text)
                elif kind == "op2":
                    self._emit(_TWO_CHAR[text], text)
                elif kind == "dots":
                    self._emit(T.ELLIPSIS, text)
                else:   # op1: single-character operators other than braces
                    self._emit(_ONE_CHAR[text], text)
                continue
            if self.pos >= n:
                return
            ch = src[self.pos]
            nx = src[self.pos + 1] if self.pos + 1 < n else ""
            if ch == '"':
                self._emit(T.OQUOTE, '"')
                self._quoted()
            elif ch == "<" and nx == "<" and self._heredoc_intro():
                pass
            elif ch == "{":
                depth += 1
                self._emit(T.OBRACE, "{")
            elif ch == "~" and nx == "}" and in_template and depth == 0:
                self._emit(T.TEMPLATE_SEQ_END, "~}")
                return
            elif ch == "}":
                if in_template and depth == 0:
                    self._emit(T.TEMPLATE_SEQ_END, "}")
                    return
                depth -= 1
                self._emit(T.CBRACE, "}")
            elif ch == "/" and nx == "*":
                self._emit(T.INVALID, src[self.pos:])       # unterminated block comment
            elif ch == "<":
                self._emit(T.LESS_THAN, "<")
            elif is_id_start(ch):
                j = self.pos + 1
                while j < n and is_id_continue(src[j]):
                    j += 1
                self._emit(T.IDENT, src[self.pos:j])
            elif ch == "$" and nx == "{":
                # only meaningful inside templates; invalid in normal mode
                self._emit(T.INVALID, "${")
            else:
                self._emit(T.INVALID, ch)

    # -- quoted template mode ----------------------------------------------
    def _lit(self, typ: T, end: int) -> None:
        """Emit the literal run from the current position up to ``end``."""
        if end > self.pos:
            self._emit(typ, self.src[self.pos:end])

    def _quoted(self) -> None:
        i = self.pos
        src, n = self.src, self.n
        while True:
            if i >= n:
                self._lit(T.QUOTED_LIT, i)
                return
            ch = src[i]
            nx = src[i + 1] if i + 1 < n else ""
            if ch == "\\" and nx:
                i += 2
            elif ch == '"':
                self._lit(T.QUOTED_LIT, i)
                self._emit(T.CQUOTE, '"')
                return
            elif ch == "\n":
                # a quoted string may not span lines; stop here so the parser
                # can report the problem at a useful position.
                self._lit(T.QUOTED_LIT, i)
                return
            elif ch in "$%" and nx == ch and src.startswith("{", i + 2):
                i += 3                 # $${ and %%{ are escapes for ${ and %{
            elif ch in "$%" and nx == "{":
                self._lit(T.QUOTED_LIT, i)
                self._template_seq(ch)
                i = self.pos
            else:
                i += 1

    def _template_seq(self, intro: str) -> None:
        text = intro + "{"
        if self._peek(2) == "~":
            text += "~"
        self._emit(T.TEMPLATE_INTERP if intro == "$" else T.TEMPLATE_CONTROL, text)
        self._normal(in_template=True)

    # -- heredoc mode ------------------------------------------------------
    def _heredoc_intro(self) -> bool:
        j = self.pos + 2
        if j < self.n and self.src[j] == "-":
            j += 1
        k = j
        if not (k < self.n and is_id_start(self.src[k])):
            return False
        while k < self.n and is_id_continue(self.src[k]):
            k += 1
        marker = self.src[j:k]
        if self.src.startswith("\r\n", k):
            k += 2
        elif self.src.startswith("\n", k):
            k += 1
        else:
            return False
        self._emit(T.OHEREDOC, self.src[self.pos:k])
        self._heredoc_body(marker)
        return True

    def _heredoc_body(self, marker: str) -> None:
        src, n = self.src, self.n
        while self.pos < n:
            eol = src.find("\n", self.pos)
            line = src[self.pos:(n if eol < 0 else eol)].rstrip("\r")
            if line.strip(" \t") == marker:
                self._emit(T.CHEREDOC, line)
                return
            # literal / template content of one body line, newline included
            stop = n if eol < 0 else eol + 1
            i = self.pos
            while i < stop:
                ch = src[i]
                nx = src[i + 1] if i + 1 < n else ""
                if ch in "$%" and nx == ch and src.startswith("{", i + 2):
                    i += 3
                elif ch in "$%" and nx == "{":
                    self._lit(T.STRING_LIT, i)
                    self._template_seq(ch)
                    i = self.pos
                    # the sequence may have spanned lines; recompute the stop
                    eol = src.find("\n", i)
                    stop = n if eol < 0 else eol + 1
                else:
                    i += 1
            self._lit(T.STRING_LIT, i)


def scan(src: str) -> list[Token]:
    """Tokenise HCL2 native-syntax source text."""
    return Scanner(src).scan()
