"""Recursive-descent parser for HCL2 native syntax.

Restates the grammar of the HCL2 native syntax specification ("Structural
Elements", "Expressions", "Templates").  The grammar is small; what needs care
is newline handling: newlines end attributes and separate object-constructor
items, but are insignificant inside ``( )``, ``[ ]``, ``${ }`` and ``for``
expressions.  ``_nl_ignore`` is the stack that tracks which regime applies.

Error wording follows Terraform's, because the reference's contribution guide
(CONTRIBUTING.md:12) makes "``terraform validate`` does not break" the gate and
a maintainer will compare messages.
"""
from __future__ import annotations

import os
from typing import Optional

from . import ast as A
from .diagnostics import Diagnostic, ParseError, ERROR
from .scanner import scan
from .tokens import T, Token

_BINARY_LEVELS: list[dict[T, str]] = [
    {T.OR: "||"},
    {T.AND: "&&"},
    {T.EQUAL_OP: "==", T.NOT_EQUAL: "!="},
    {T.GREATER_THAN: ">", T.GREATER_THAN_EQ: ">=", T.LESS_THAN: "<", T.LESS_THAN_EQ: "<="},
    {T.PLUS: "+", T.MINUS: "-"},
    {T.STAR: "*", T.SLASH: "/", T.PERCENT: "%"},
]

_PRECEDENCE: dict[T, int] = {typ: level for level, ops in enumerate(_BINARY_LEVELS) for typ in ops}

_ESCAPES = {"n": "\n", "r": "\r", "t": "\t", '"': '"', "\\": "\\"}


def unescape_quoted(text: str) -> str:
    """Literal value of a quoted-template literal run."""
    out: list[str] = []
    i, n = 0, len(text)
    while i < n:
        ch = text[i]
        if ch == "\\" and i + 1 < n:
            nx = text[i + 1]
            if nx in _ESCAPES:
                out.append(_ESCAPES[nx])
                i += 2
                continue
            if nx == "u" and i + 6 <= n:
                try:
                    out.append(chr(int(text[i + 2:i + 6], 16)))
                    i += 6
                    continue
                except ValueError:
                    pass
            if nx == "U" and i + 10 <= n:
                try:
                    out.append(chr(int(text[i + 2:i + 10], 16)))
                    i += 10
                    continue
                except ValueError:
                    pass
            out.append(ch)
            i += 1
        elif ch in "$%" and text.startswith(ch + "{", i + 1):
            out.append(ch + "{")
            i += 3
        else:
            out.append(ch)
            i += 1
    return "".join(out)


def unescape_heredoc(text: str) -> str:
    return text.replace("$${", "${").replace("%%{", "%{")


class Parser:
    def __init__(self, src: str, filename: str = ""):
        self.filename = filename
        self.all = scan(src)
        # parser view: comments dropped; a '#'/'//' comment stands for the
        # newline it swallowed.
        self.idx: list[int] = []
        self.toks: list[Token] = []
        for i, t in enumerate(self.all):
            if t.type is T.COMMENT:
                if t.text.endswith("\n"):
                    self.idx.append(i)
                    self.toks.append(Token(T.NEWLINE, "\n", t.line, t.col + len(t.text) - 1, t.offset))
                continue
            self.idx.append(i)
            self.toks.append(t)
        self.p = 0
        self._nl_ignore = [False]
        self.diags: list[Diagnostic] = []

    # -- token access ------------------------------------------------------
    def _skip_nl(self) -> None:
        if self._nl_ignore[-1]:
            while self.toks[self.p].type is T.NEWLINE:
                self.p += 1

    def peek(self) -> Token:
        self._skip_nl()
        return self.toks[self.p]

    def peek2(self) -> Token:
        self._skip_nl()
        q = self.p + 1
        if self._nl_ignore[-1]:
            while q < len(self.toks) and self.toks[q].type is T.NEWLINE:
                q += 1
        return self.toks[min(q, len(self.toks) - 1)]

    def next(self) -> Token:
        tok = self.peek()
        if tok.type is not T.EOF:
            self.p += 1
        return tok

    def _full_index(self, p: Optional[int] = None) -> int:
        return self.idx[min(self.p if p is None else p, len(self.idx) - 1)]

    def error(self, summary: str, detail: str = "", tok: Optional[Token] = None) -> ParseError:
        tok = tok or self.peek()
        return ParseError(Diagnostic(ERROR, summary, detail, self.filename, tok.line, tok.col, "syntax"))

    def expect(self, typ: T, summary: str, detail: str = "") -> Token:
        tok = self.peek()
        if tok.type is not typ:
            raise self.error(summary, detail, tok)
        return self.next()

    def _mark(self) -> int:
        """Parser position of the next significant token (cheap: positions are
        only turned into line/column when a node is actually finished)."""
        self._skip_nl()
        return self.p

    def _finish(self, node, mark: int):
        tok = self.toks[mark]
        node.tok_start = self.idx[mark]
        node.pos = A.Pos(tok.line, tok.col)
        # end = full index just past the last consumed parser token
        node.tok_end = self.idx[self.p - 1] + 1 if self.p > 0 else 0
        return node

    # -- structure ---------------------------------------------------------
    def parse_file(self) -> A.Body:
        for t in self.all:
            if t.type is T.INVALID:
                self.diags.append(Diagnostic(ERROR, "Invalid character",
                                             f"This character is not used within the language: {t.text[:10]!r}",
                                             self.filename, t.line, t.col, "syntax"))
        body = self.parse_body(T.EOF)
        return body

    def parse_body(self, end: T) -> A.Body:
        mark = self._mark()
        body = A.Body()
        nl_depth = len(self._nl_ignore)
        while True:
            while self.toks[self.p].type is T.NEWLINE:
                self.p += 1
            tok = self.toks[self.p]
            if tok.type is end or tok.type is T.EOF:
                break
            start_p = self.p
            try:
                if tok.type is not T.IDENT:
                    raise self.error("Argument or block definition required",
                                     "An argument or block definition is required here.", tok)
                body.items.append(self.parse_body_item())
            except ParseError as e:
                self.diags.append(e.diag)
                del self._nl_ignore[nl_depth:]
                self._recover(start_p, end)
        return self._finish(body, mark)

    def _recover(self, start_p: int, end: T) -> None:
        """Skip to the end of the line on which bracket nesting returns to the
        level it had where the broken item began."""
        self.p = max(self.p, start_p)
        depth = 0
        while True:
            t = self.toks[self.p]
            if t.type is T.EOF:
                return
            if t.type in (T.OBRACE, T.OBRACK, T.OPAREN, T.TEMPLATE_INTERP, T.TEMPLATE_CONTROL):
                depth += 1
            elif t.type in (T.CBRACE, T.CBRACK, T.CPAREN, T.TEMPLATE_SEQ_END):
                if depth == 0:
                    if t.type is end:
                        return      # closing brace of the enclosing body
                    self.p += 1     # stray closer: step over it
                    continue
                depth -= 1
            elif t.type is T.NEWLINE and depth == 0 and self.p > start_p:
                return
            self.p += 1

    def parse_body_item(self):
        mark = self._mark()
        name_p = self.p
        name = self.next()
        nxt = self.toks[self.p]
        if nxt.type is T.EQUAL:
            return self._finish_attribute(name, name_p, mark, one_line=False)
        if nxt.type in (T.OQUOTE, T.IDENT, T.OBRACE):
            return self._finish_block(name, name_p, mark)
        if nxt.type is T.NEWLINE or nxt.type is T.EOF:
            raise self.error("Invalid block definition",
                             "A block definition must have block content delimited by \"{\" and \"}\", "
                             "starting on the same line as the block header.", nxt)
        raise self.error("Argument or block definition required",
                         "An argument or block definition is required here. To set an argument, "
                         "use the equals sign \"=\" to introduce the argument value.", nxt)

    def _finish_attribute(self, name: Token, name_p: int, mark, one_line: bool) -> A.Attribute:
        eq_p = self.p
        self.next()  # '='
        expr_start = self._full_index()
        expr = self.parse_expr()
        expr_end = self.idx[self.p - 1] + 1
        attr = A.Attribute(name.text, expr, name_tok=self.idx[name_p], eq_tok=self.idx[eq_p],
                           expr_tok_start=expr_start, expr_tok_end=expr_end)
        self._finish(attr, mark)
        if not one_line:
            end = self.toks[self.p]
            if end.type not in (T.NEWLINE, T.EOF):
                raise self.error("Missing newline after argument",
                                 "An argument definition must end with a newline.", end)
        return attr

    def _finish_block(self, typ: Token, type_p: int, mark) -> A.Block:
        labels: list[A.Label] = []
        while True:
            tok = self.toks[self.p]
            if tok.type is T.IDENT:
                labels.append(A.Label(tok.text, False, self.idx[self.p], self.idx[self.p] + 1))
                self.p += 1
            elif tok.type is T.OQUOTE:
                start = self.idx[self.p]
                self.p += 1
                value = ""
                lit = self.toks[self.p]
                if lit.type is T.QUOTED_LIT:
                    value = unescape_quoted(lit.text)
                    self.p += 1
                close = self.toks[self.p]
                if close.type is not T.CQUOTE:
                    raise self.error("Invalid string literal",
                                     "Template sequences are not allowed in this string. "
                                     "Block labels must be literal strings.", close)
                self.p += 1
                labels.append(A.Label(value, True, start, self.idx[self.p - 1] + 1))
            elif tok.type is T.OBRACE:
                break
            elif tok.type is T.EQUAL:
                raise self.error("Invalid block definition",
                                 "The equals sign \"=\" indicates an argument definition, and must "
                                 "not be used when defining a block.", tok)
            elif tok.type is T.NEWLINE:
                raise self.error("Invalid block definition",
                                 "A block definition must have block content delimited by \"{\" and "
                                 "\"}\", starting on the same line as the block header.", tok)
            else:
                raise self.error("Invalid block definition",
                                 "Either a quoted string block label or an opening brace (\"{\") is "
                                 "expected here.", tok)
        obrace_p = self.p
        self.p += 1  # '{'
        one_line = False
        nxt = self.toks[self.p]
        if nxt.type is T.NEWLINE:
            self._nl_ignore.append(False)
            body = self.parse_body(T.CBRACE)
            self._nl_ignore.pop()
            close = self.toks[self.p]
            if close.type is not T.CBRACE:
                raise self.error("Unclosed configuration block",
                                 "There is no closing brace for this block before the end of the file. "
                                 "This may be caused by incorrect brace nesting elsewhere in this file.",
                                 self.toks[obrace_p])
        else:
            one_line = True
            bmark = self._mark()
            body = A.Body()
            if nxt.type is T.IDENT:
                amark = self._mark()
                name_p = self.p
                name = self.next()
                if self.toks[self.p].type is not T.EQUAL:
                    if self.toks[self.p].type in (T.OBRACE, T.OQUOTE, T.IDENT):
                        raise self.error("Invalid single-argument block definition",
                                         "A single-line block definition cannot contain another block "
                                         "definition. Use the multi-line block syntax instead.",
                                         self.toks[self.p])
                    raise self.error("Invalid single-argument block definition",
                                     "A single-line block definition must contain exactly one "
                                     "argument definition.", self.toks[self.p])
                body.items.append(self._finish_attribute(name, name_p, amark, one_line=True))
            self._finish(body, bmark)
            close = self.toks[self.p]
            if close.type is not T.CBRACE:
                if close.type is T.COMMA:
                    raise self.error("Invalid single-argument block definition",
                                     "Single-line block syntax can include only one argument definition. "
                                     "To define multiple arguments, use the multi-line block syntax with "
                                     "one argument definition per line.", close)
                raise self.error("Invalid single-argument block definition",
                                 "A single-line block definition must end with a closing brace "
                                 "immediately after its single argument definition.", close)
        cbrace_p = self.p
        self.p += 1  # '}'
        block = A.Block(typ.text, labels, body, one_line, type_tok=self.idx[type_p],
                        obrace_tok=self.idx[obrace_p], cbrace_tok=self.idx[cbrace_p])
        self._finish(block, mark)
        end = self.toks[self.p]
        if end.type not in (T.NEWLINE, T.EOF):
            raise self.error("Missing newline after block definition",
                             "A block definition must end with a newline.", end)
        return block

    # -- expressions -------------------------------------------------------
    def parse_expr(self) -> A.Expr:
        mark = self._mark()
        cond = self._binary(0)
        if self.toks[self.p].type is not T.QUESTION and not self._nl_ignore[-1]:
            return cond                  # common case: nothing follows on this line
        if self.peek().type is T.QUESTION:
            self.next()
            t = self.parse_expr()
            colon = self.peek()
            if colon.type is not T.COLON:
                raise self.error("Missing false expression in conditional",
                                 "The conditional operator (...?...:...) requires a false expression, "
                                 "delimited by a colon.", colon)
            self.next()
            f = self.parse_expr()
            return self._finish(A.Conditional(cond, t, f), mark)
        return cond

    def _binary(self, min_prec: int) -> A.Expr:
        """Precedence climbing over the six binary operator levels."""
        mark = self._mark()
        lhs = self._unary()
        while True:
            tok = self.peek()
            prec = _PRECEDENCE.get(tok.type)
            if prec is None or prec < min_prec:
                return lhs
            self.next()
            rhs = self._binary(prec + 1)
            lhs = self._finish(A.Binary(tok.text, lhs, rhs), mark)

    def _unary(self) -> A.Expr:
        tok = self.peek()
        if tok.type in (T.MINUS, T.BANG):
            mark = self._mark()
            self.next()
            return self._finish(A.Unary(tok.text, self._unary()), mark)
        return self._postfix()

    def _postfix(self) -> A.Expr:
        mark = self._mark()
        expr = self._term()
        while True:
            tok = self.peek()
            if tok.type is T.DOT:
                self.next()
                nt = self.peek()
                if nt.type is T.IDENT:
                    self.next()
                    expr = self._add_step(expr, A.Step("attr", nt.text, line=nt.line, col=nt.col), mark)
                elif nt.type is T.NUMBER_LIT:
                    self.next()
                    for part in nt.text.split("."):
                        if not part.isdigit():
                            raise self.error("Invalid legacy index syntax",
                                             "When using the legacy index syntax, chaining two indexes "
                                             "together is not permitted. Use the [...] index syntax instead.", nt)
                        expr = self._add_step(expr, A.Step("index", int(part), legacy=True,
                                                           line=nt.line, col=nt.col), mark)
                elif nt.type is T.STAR:
                    self.next()
                    steps: list[A.Step] = []
                    while self.peek().type is T.DOT and self.peek2().type is T.IDENT:
                        self.next()
                        a = self.next()
                        steps.append(A.Step("attr", a.text, line=a.line, col=a.col))
                    expr = self._finish(A.Splat(expr, steps, full=False), mark)
                else:
                    raise self.error("Invalid attribute name",
                                     "An attribute name is required after a dot.", nt)
            elif tok.type is T.OBRACK:
                self.next()
                self._nl_ignore.append(True)
                if self.peek().type is T.STAR and self.peek2().type is T.CBRACK:
                    self.next()
                    self.next()
                    self._nl_ignore.pop()
                    steps = []
                    while True:
                        t2 = self.peek()
                        if t2.type is T.DOT and self.peek2().type is T.IDENT:
                            self.next()
                            a = self.next()
                            steps.append(A.Step("attr", a.text, line=a.line, col=a.col))
                        elif t2.type is T.OBRACK and self._literal_index_ahead():
                            self.next()
                            key = self._literal_key()
                            self.expect(T.CBRACK, "Missing close bracket on index")
                            steps.append(A.Step("index", key, line=t2.line, col=t2.col))
                        else:
                            break
                    expr = self._finish(A.Splat(expr, steps, full=True), mark)
                    continue
                key_expr = self.parse_expr()
                close = self.peek()
                if close.type is not T.CBRACK:
                    raise self.error("Missing close bracket on index",
                                     "The index operator must end with a closing bracket (\"]\").", close)
                self.next()
                self._nl_ignore.pop()
                lit = self._static_key(key_expr)
                if lit is not None:
                    expr = self._add_step(expr, A.Step("index", lit, line=tok.line, col=tok.col), mark)
                else:
                    expr = self._finish(A.Index(expr, key_expr), mark)
            else:
                return expr

    def _literal_index_ahead(self) -> bool:
        q = self.p + 1
        t = self.toks[q]
        if t.type is T.NUMBER_LIT:
            return self.toks[q + 1].type is T.CBRACK
        if t.type is T.OQUOTE:
            return (self.toks[q + 1].type is T.QUOTED_LIT and self.toks[q + 2].type is T.CQUOTE
                    and self.toks[q + 3].type is T.CBRACK)
        return False

    def _literal_key(self):
        t = self.next()
        if t.type is T.NUMBER_LIT:
            return _number_value(t.text)
        lit = self.next()
        self.next()
        return unescape_quoted(lit.text)

    @staticmethod
    def _static_key(e: A.Expr):
        if isinstance(e, A.Literal) and isinstance(e.value, (int, float)) and not isinstance(e.value, bool):
            return e.value
        if isinstance(e, A.Template) and not e.heredoc:
            return e.literal_value()
        return None

    def _add_step(self, expr: A.Expr, step: A.Step, mark) -> A.Expr:
        if isinstance(expr, A.ScopeTraversal):
            expr.steps.append(step)
            return self._finish(expr, mark)
        if isinstance(expr, A.RelativeTraversal):
            expr.steps.append(step)
            return self._finish(expr, mark)
        return self._finish(A.RelativeTraversal(expr, [step]), mark)

    def _term(self) -> A.Expr:
        mark = self._mark()
        tok = self.peek()
        typ = tok.type
        if typ is T.NUMBER_LIT:
            self.next()
            return self._finish(A.Literal(_number_value(tok.text), tok.text), mark)
        if typ is T.IDENT:
            self.next()
            nxt = self.peek()
            if nxt.type is T.OPAREN or nxt.type is T.DOUBLE_COLON:
                return self._call(tok, mark)
            if tok.text == "true":
                return self._finish(A.Literal(True, "true"), mark)
            if tok.text == "false":
                return self._finish(A.Literal(False, "false"), mark)
            if tok.text == "null":
                return self._finish(A.Literal(None, "null"), mark)
            return self._finish(A.ScopeTraversal(tok.text, []), mark)
        if typ is T.OQUOTE:
            return self._quoted_template(mark)
        if typ is T.OHEREDOC:
            return self._heredoc_template(mark)
        if typ is T.OPAREN:
            self.next()
            self._nl_ignore.append(True)
            inner = self.parse_expr()
            close = self.peek()
            if close.type is not T.CPAREN:
                raise self.error("Unbalanced parentheses",
                                 "Expected a closing parenthesis to terminate the expression.", close)
            self.next()
            self._nl_ignore.pop()
            return self._finish(A.Parens(inner), mark)
        if typ is T.OBRACK:
            return self._tuple(mark)
        if typ is T.OBRACE:
            return self._object(mark)
        if typ is T.EOF or typ is T.NEWLINE:
            raise self.error("Invalid expression", "Expected the start of an expression, but found "
                             "the end of the line." if typ is T.NEWLINE else
                             "Expected the start of an expression, but found the end of the file.", tok)
        raise self.error("Invalid expression",
                         "Expected the start of an expression, but found an invalid expression token.", tok)

    def _call(self, name_tok: Token, mark) -> A.Expr:
        name = name_tok.text
        while self.peek().type is T.DOUBLE_COLON:
            self.next()
            seg = self.expect(T.IDENT, "Missing function name",
                              "Function scope resolution symbol :: must be followed by a function name.")
            name += "::" + seg.text
        self.expect(T.OPAREN, "Missing open parenthesis")
        self._nl_ignore.append(True)
        args: list[A.Expr] = []
        expand = False
        while True:
            if self.peek().type is T.CPAREN:
                break
            args.append(self.parse_expr())
            sep = self.peek()
            if sep.type is T.COMMA:
                self.next()
                continue
            if sep.type is T.ELLIPSIS:
                self.next()
                expand = True
                if self.peek().type is not T.CPAREN:
                    raise self.error("Missing closing parenthesis",
                                     "An expanded function argument (with ...) must be immediately "
                                     "followed by closing parentheses.")
                break
            if sep.type is T.CPAREN:
                break
            raise self.error("Missing argument separator",
                             "A comma is required to separate each function argument from the next.", sep)
        self.next()  # ')'
        self._nl_ignore.pop()
        return self._finish(A.FunctionCall(name, args, expand), mark)

    def _tuple(self, mark) -> A.Expr:
        open_tok = self.next()
        self._nl_ignore.append(True)
        if self.peek().type is T.IDENT and self.peek().text == "for" and self._looks_like_for():
            expr = self._for_expr(mark, T.CBRACK)
            self._nl_ignore.pop()
            return expr
        items: list[A.Expr] = []
        while True:
            if self.peek().type is T.CBRACK:
                break
            items.append(self.parse_expr())
            sep = self.peek()
            if sep.type is T.COMMA:
                self.next()
                continue
            if sep.type is T.CBRACK:
                break
            if sep.type is T.EOF:
                raise self.error("Unterminated tuple constructor expression",
                                 "There is no corresponding closing bracket before the end of the file.",
                                 open_tok)
            raise self.error("Missing item separator",
                             "Expected a comma to mark the beginning of the next item.", sep)
        self.next()
        self._nl_ignore.pop()
        return self._finish(A.TupleCons(items), mark)

    def _looks_like_for(self) -> bool:
        # "for" IDENT ("," IDENT)? "in"
        q = self.p + 1

        def skip(q):
            while self.toks[q].type is T.NEWLINE:
                q += 1
            return q
        q = skip(q)
        if self.toks[q].type is not T.IDENT:
            return False
        q = skip(q + 1)
        if self.toks[q].type is T.COMMA:
            q = skip(q + 1)
            if self.toks[q].type is not T.IDENT:
                return False
            q = skip(q + 1)
        return self.toks[q].type is T.IDENT and self.toks[q].text == "in"

    def _object(self, mark) -> A.Expr:
        open_tok = self.next()
        self._nl_ignore.append(True)
        if self.peek().type is T.IDENT and self.peek().text == "for" and self._looks_like_for():
            expr = self._for_expr(mark, T.CBRACE)
            self._nl_ignore.pop()
            return expr
        self._nl_ignore.pop()
        self._nl_ignore.append(False)
        items: list[tuple[A.ObjectKey, A.Expr]] = []
        colon_keys: list[bool] = []
        while True:
            while self.toks[self.p].type is T.NEWLINE:
                self.p += 1
            tok = self.toks[self.p]
            if tok.type is T.CBRACE:
                break
            if tok.type is T.EOF:
                raise self.error("Unterminated object constructor expression",
                                 "There is no corresponding closing brace before the end of the file. "
                                 "This may be caused by incorrect brace nesting elsewhere in this file.",
                                 open_tok)
            kmark = self._mark()
            forced = tok.type is T.OPAREN
            key = self.parse_expr()
            bare = None
            if isinstance(key, A.ScopeTraversal):
                # A dotted key (``aws.west = aws.usw2`` in a module's ``providers`` map) is
                # accepted here: HCL objects to it only when the key is EVALUATED, so that
                # callers reading the map statically may allow it.  config/validate.py
                # raises "Ambiguous attribute key" wherever the object is a value.
                if not key.steps:
                    bare = key.root
            elif isinstance(key, A.Literal) and key.text in ("true", "false", "null"):
                bare = key.text
            okey = self._finish(A.ObjectKey(key, bare, forced), kmark)
            sep = self.toks[self.p]
            if sep.type not in (T.EQUAL, T.COLON):
                if sep.type in (T.NEWLINE, T.COMMA):
                    raise self.error("Missing attribute value",
                                     "Expected an attribute value, introduced by an equals sign (\"=\").", sep)
                raise self.error("Missing key/value separator",
                                 "Expected an equals sign (\"=\") to mark the beginning of the attribute value.",
                                 sep)
            self.p += 1
            colon_keys.append(sep.type is T.COLON)
            value = self.parse_expr()
            items.append((okey, value))
            end = self.toks[self.p]
            if end.type is T.COMMA:
                self.p += 1
                continue
            if end.type is T.NEWLINE:
                continue
            if end.type is T.CBRACE:
                break
            if end.type is T.EOF:
                raise self.error("Unterminated object constructor expression",
                                 "There is no corresponding closing brace before the end of the file. "
                                 "This may be caused by incorrect brace nesting elsewhere in this file.",
                                 open_tok)
            raise self.error("Missing attribute separator",
                             "Expected a newline or comma to mark the beginning of the next attribute.", end)
        self.p += 1  # '}'
        self._nl_ignore.pop()
        return self._finish(A.ObjectCons(items, colon_keys), mark)

    def _for_expr(self, mark, close: T) -> A.Expr:
        self.next()  # 'for'
        first = self.expect(T.IDENT, "Invalid 'for' expression",
                            "For expression requires variable name after 'for'.")
        key_var: Optional[str] = None
        val_var = first.text
        if self.peek().type is T.COMMA:
            self.next()
            second = self.expect(T.IDENT, "Invalid 'for' expression",
                                 "For expression requires value variable name after comma.")
            key_var, val_var = first.text, second.text
        kw = self.peek()
        if not (kw.type is T.IDENT and kw.text == "in"):
            raise self.error("Invalid 'for' expression",
                             "For expression requires the 'in' keyword after its name declarations.", kw)
        self.next()
        coll = self.parse_expr()
        self.expect(T.COLON, "Invalid 'for' expression",
                    "For expression requires a colon after the collection expression.")
        key_expr: Optional[A.Expr] = None
        val_expr = self.parse_expr()
        group = False
        if close is T.CBRACE:
            arrow = self.peek()
            if arrow.type is not T.FAT_ARROW:
                raise self.error("Invalid 'for' expression",
                                 "For expression in an object constructor requires => symbol after the key expression.",
                                 arrow)
            self.next()
            key_expr, val_expr = val_expr, self.parse_expr()
            if self.peek().type is T.ELLIPSIS:
                self.next()
                group = True
        elif self.peek().type is T.FAT_ARROW:
            raise self.error("Invalid 'for' expression",
                             "Key expression is not valid when building a tuple.")
        cond: Optional[A.Expr] = None
        kw = self.peek()
        if kw.type is T.IDENT and kw.text == "if":
            self.next()
            cond = self.parse_expr()
        end = self.peek()
        if end.type is not close:
            raise self.error("Invalid 'for' expression",
                             "Extra characters after the end of the 'for' expression.", end)
        self.next()
        return self._finish(A.ForExpr(key_var, val_var, coll, key_expr, val_expr, cond, group), mark)

    # -- templates ---------------------------------------------------------
    def _quoted_template(self, mark) -> A.Expr:
        open_tok = self.next()  # OQUOTE
        parts = self._template_parts(T.CQUOTE, T.QUOTED_LIT, unescape_quoted, open_tok)
        return self._finish(A.Template(_merge_literals(parts)), mark)

    def _heredoc_template(self, mark) -> A.Expr:
        open_tok = self.next()  # OHEREDOC
        strip = open_tok.text.startswith("<<-")
        parts = self._template_parts(T.CHEREDOC, T.STRING_LIT, unescape_heredoc, open_tok)
        parts = _merge_literals(parts)
        if strip:
            parts = _strip_indent(parts)
        return self._finish(A.Template(parts, heredoc=True, indent_strip=strip), mark)

    def _template_parts(self, close: T, lit: T, unescape, open_tok: Token,
                        stop_keywords: tuple[str, ...] = ()) -> list:
        """Parse template content up to ``close`` (or, for directive bodies, up
        to a ``%{`` whose keyword is in ``stop_keywords`` -- that directive is
        left unconsumed and its keyword returned as the final list element in a
        1-tuple)."""
        parts: list = []
        while True:
            tok = self.toks[self.p]
            if tok.type is close:
                if stop_keywords:
                    raise self.error("Unterminated template directive",
                                     f"Expected %{{ {stop_keywords[0]} }} before the end of the template.", tok)
                self.p += 1
                return parts
            if tok.type is lit:
                parts.append(unescape(tok.text))
                self.p += 1
            elif tok.type is T.TEMPLATE_INTERP:
                self.p += 1
                self._nl_ignore.append(True)
                e = self.parse_expr()
                end = self.peek()
                if end.type is not T.TEMPLATE_SEQ_END:
                    raise self.error("Extra characters after interpolation expression",
                                     "Expected a closing brace to end the interpolation expression, "
                                     "but found extra characters.", end)
                self.next()
                self._nl_ignore.pop()
                parts.append(e)
            elif tok.type is T.TEMPLATE_CONTROL:
                save = self.p
                self.p += 1
                self._nl_ignore.append(True)
                kw = self.peek()
                if kw.type is not T.IDENT:
                    raise self.error("Invalid template directive",
                                     "A template directive keyword (\"if\", \"for\", etc) is expected "
                                     "at the beginning of a %{ sequence.", kw)
                if kw.text in stop_keywords:
                    self._nl_ignore.pop()
                    self.p = save
                    parts.append((kw.text,))
                    return parts
                parts.append(self._directive(kw, close, lit, unescape, open_tok))
            elif tok.type in (T.EOF, T.NEWLINE):
                raise self.error("Unterminated template string" if close is T.CQUOTE
                                 else "Unterminated heredoc",
                                 "No closing marker was found for the string.", open_tok)
            else:
                raise self.error("Invalid template", "Unexpected token inside template.", tok)

    def _end_seq(self) -> None:
        end = self.peek()
        if end.type is not T.TEMPLATE_SEQ_END:
            raise self.error("Extra characters in template directive",
                             "Expected a closing brace to end the sequence.", end)
        self.next()
        self._nl_ignore.pop()

    def _consume_directive(self, keyword: str) -> None:
        """Consume ``%{ keyword }`` at the current position."""
        self.p += 1
        self._nl_ignore.append(True)
        kw = self.next()
        assert kw.text == keyword
        self._end_seq()

    def _directive(self, kw: Token, close, lit, unescape, open_tok):
        mark = self._mark()
        if kw.text == "if":
            self.next()
            cond = self.parse_expr()
            self._end_seq()
            then = self._template_parts(close, lit, unescape, open_tok, ("else", "endif"))
            stop = then.pop()[0]
            otherwise: list = []
            if stop == "else":
                self._consume_directive("else")
                otherwise = self._template_parts(close, lit, unescape, open_tok, ("endif",))
                otherwise.pop()
            self._consume_directive("endif")
            return self._finish(A.TemplateIf(cond, _merge_literals(then), _merge_literals(otherwise)), mark)
        if kw.text == "for":
            self.next()
            first = self.expect(T.IDENT, "Invalid 'for' directive",
                                "For directive requires variable name after 'for'.")
            key_var, val_var = None, first.text
            if self.peek().type is T.COMMA:
                self.next()
                second = self.expect(T.IDENT, "Invalid 'for' directive",
                                     "For directive requires value variable name after comma.")
                key_var, val_var = first.text, second.text
            inkw = self.peek()
            if not (inkw.type is T.IDENT and inkw.text == "in"):
                raise self.error("Invalid 'for' directive",
                                 "For directive requires 'in' keyword after names.", inkw)
            self.next()
            coll = self.parse_expr()
            self._end_seq()
            body = self._template_parts(close, lit, unescape, open_tok, ("endfor",))
            body.pop()
            self._consume_directive("endfor")
            return self._finish(A.TemplateFor(key_var, val_var, coll, _merge_literals(body)), mark)
        raise self.error("Invalid template control keyword",
                         f"\"{kw.text}\" is not a valid template control keyword.", kw)


def _number_value(text: str):
    try:
        return int(text)
    except ValueError:
        return float(text)


def _merge_literals(parts: list) -> list:
    out: list = []
    for p in parts:
        if isinstance(p, str) and out and isinstance(out[-1], str):
            out[-1] += p
        else:
            out.append(p)
    return out


def _strip_indent(parts: list) -> list:
    """``<<-`` heredocs drop the indentation common to all lines (HCL2 spec,
    "Template Expressions")."""
    text_lines: list[str] = []
    at_line_start = True
    for p in parts:
        if isinstance(p, str):
            for seg in p.splitlines(keepends=True):
                if at_line_start:
                    text_lines.append(seg)
                at_line_start = seg.endswith("\n")
        else:
            if at_line_start:
                text_lines.append("")   # line begins with an interpolation: indent 0
            at_line_start = False
    indents = [len(l) - len(l.lstrip(" \t")) for l in text_lines if l.strip("\r\n") != "" or l == ""]
    cut = min(indents) if indents else 0
    if cut == 0:
        return parts
    out: list = []
    at_line_start = True
    for p in parts:
        if isinstance(p, str):
            segs = []
            for seg in p.splitlines(keepends=True):
                if at_line_start:
                    lead = len(seg) - len(seg.lstrip(" \t"))
                    seg = seg[min(cut, lead):]
                at_line_start = seg.endswith("\n")
                segs.append(seg)
            out.append("".join(segs))
        else:
            out.append(p)
            at_line_start = False
    return out


def parse(src: str, filename: str = "") -> tuple[A.Body, list[Diagnostic], list[Token]]:
    """Parse one file.  Returns (body, diagnostics, full token list)."""
    parser = Parser(src, filename)
    try:
        body = parser.parse_file()
    except RecursionError:
        # hundreds of brackets inside one another: refuse the file rather than die in it
        tok = parser.toks[min(parser.p, len(parser.toks) - 1)]
        diag = Diagnostic(ERROR, "Expression nested too deeply",
                          "This file nests brackets more deeply than this checker will follow.",
                          filename, tok.line, tok.col, "syntax")
        return A.Body(), parser.diags + [diag], parser.all
    return body, parser.diags, parser.all


_FILE_CACHE: dict[tuple, tuple[str, A.Body, list[Diagnostic], list[Token]]] = {}


def parse_path(path: str, display: str) -> tuple[str, A.Body, list[Diagnostic], list[Token]]:
    """Read and parse a file, remembering the result until the file changes
    (size or modification time).  One ``check`` run looks at each file several
    times -- fmt, the module that owns it, modules that call that module -- and
    they all share one parse.  Callers must treat the result as read-only;
    ``fmt`` restores the ``spaces_before`` values it touches."""
    try:
        st = os.stat(path)
    except OSError as e:
        return _unreadable(display, e)
    key = (os.path.abspath(path), display, st.st_mtime_ns, st.st_size)
    hit = _FILE_CACHE.get(key)
    if hit is None:
        try:
            with open(path, "rb") as fh:
                raw = fh.read()
        except OSError as e:
            return _unreadable(display, e)
        try:
            src = raw.decode("utf-8")
        except UnicodeDecodeError as e:
            # HCL's scanner reports this for the one file and carries on with the rest
            line = raw.count(b"\n", 0, e.start) + 1
            col = e.start - (raw.rfind(b"\n", 0, e.start) + 1) + 1
            diag = Diagnostic(ERROR, "Invalid character encoding",
                              "All input files must be UTF-8 encoded. Ensure that UTF-8 encoding is selected "
                              "in your editor.", display, line, col, "syntax")
            return "", A.Body(), [diag], []
        body, diags, tokens = parse(src, display)
        if len(_FILE_CACHE) > 4096:
            _FILE_CACHE.clear()
        hit = _FILE_CACHE[key] = (src, body, diags, tokens)
    return hit


def _unreadable(display: str, err: OSError) -> tuple[str, A.Body, list[Diagnostic], list[Token]]:
    """A file that cannot be opened (dangling symlink, permissions) is that file's
    problem only: one error diagnostic, no tokens, and the run goes on."""
    why = err.strerror or str(err)
    diag = Diagnostic(ERROR, "Failed to read file", f"The file \"{display}\" could not be read: {why}.",
                      display, 0, 0, "syntax")
    return "", A.Body(), [diag], []
