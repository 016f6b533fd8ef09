"""Syntax tree for HCL2 native syntax (bodies, blocks, attributes, expressions).

Node kinds follow the HCL2 native syntax specification's "Structural Elements"
and "Expressions" sections.  Each node carries the index range of the tokens it
was parsed from (``tok_start`` inclusive, ``tok_end`` exclusive, indices into the
*full* token list including comments and newlines) so later passes -- the
Terraform-level rewrites in ``fmt/rewrites.py`` and the diagnostics in
``config/validate.py`` -- can point back into the source.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Iterator, Optional, Union


@dataclass
class Pos:
    line: int
    col: int


@dataclass
class Node:
    tok_start: int = field(default=0, kw_only=True)
    tok_end: int = field(default=0, kw_only=True)
    pos: Pos = field(default_factory=lambda: Pos(0, 0), kw_only=True)

    def children(self) -> Iterator["Node"]:
        return iter(())

    def walk(self) -> Iterator["Node"]:
        yield self
        for ch in self.children():
            yield from ch.walk()


# --------------------------------------------------------------------------
# expressions
# --------------------------------------------------------------------------
class Expr(Node):
    pass


@dataclass
class Literal(Expr):
    """null, true, false or a number."""
    value: Union[None, bool, int, float]
    text: str = ""


@dataclass
class TemplateIf(Node):
    cond: Expr
    then: list
    otherwise: list

    def children(self):
        yield self.cond
        for p in self.then + self.otherwise:
            if isinstance(p, Node):
                yield p


@dataclass
class TemplateFor(Node):
    key_var: Optional[str]
    val_var: str
    coll: Expr
    body: list

    def children(self):
        yield self.coll
        for p in self.body:
            if isinstance(p, Node):
                yield p


@dataclass
class Template(Expr):
    """A quoted or heredoc template.  ``parts`` holds ``str`` (already
    unescaped literal text), ``Expr`` (interpolations) and TemplateIf/For."""
    parts: list
    heredoc: bool = False
    indent_strip: bool = False   # <<- flavour

    def children(self):
        for p in self.parts:
            if isinstance(p, Node):
                yield p

    def literal_value(self) -> Optional[str]:
        if all(isinstance(p, str) for p in self.parts):
            return "".join(self.parts)
        return None

    def is_single_interpolation(self) -> bool:
        """True for ``"${expr}"`` -- the form ``terraform fmt`` unwraps."""
        return (not self.heredoc and len(self.parts) == 1
                and isinstance(self.parts[0], Expr))


@dataclass
class TupleCons(Expr):
    items: list[Expr]

    def children(self):
        return iter(self.items)


@dataclass
class ObjectKey(Expr):
    """Key position of an object constructor.  A bare identifier there is a
    literal key, not a variable reference (HCL2 spec, "Collection Values")."""
    wrapped: Expr
    bare_name: Optional[str] = None     # set when the key is a bare identifier
    forced_expr: bool = False           # key was written in parentheses

    def children(self):
        if self.bare_name is None:
            yield self.wrapped

    def literal_key(self) -> Optional[str]:
        if self.bare_name is not None:
            return self.bare_name
        w = self.wrapped
        if isinstance(w, Template):
            return w.literal_value()
        if isinstance(w, Literal):
            return w.text
        return None


@dataclass
class ObjectCons(Expr):
    items: list[tuple[ObjectKey, Expr]]
    colon_keys: list[bool] = field(default_factory=list)  # item used ':' not '='

    def children(self):
        for k, v in self.items:
            yield k
            yield v

    def get(self, name: str) -> Optional[Expr]:
        for k, v in self.items:
            if k.literal_key() == name:
                return v
        return None


@dataclass
class FunctionCall(Expr):
    name: str
    args: list[Expr]
    expand_final: bool = False

    def children(self):
        return iter(self.args)


@dataclass
class ForExpr(Expr):
    key_var: Optional[str]
    val_var: str
    coll: Expr
    key_expr: Optional[Expr]    # set for the object form
    val_expr: Expr
    cond: Optional[Expr]
    group: bool = False

    def children(self):
        yield self.coll
        if self.key_expr is not None:
            yield self.key_expr
        yield self.val_expr
        if self.cond is not None:
            yield self.cond


@dataclass
class Conditional(Expr):
    cond: Expr
    true: Expr
    false: Expr

    def children(self):
        yield self.cond
        yield self.true
        yield self.false


@dataclass
class Binary(Expr):
    op: str
    lhs: Expr
    rhs: Expr

    def children(self):
        yield self.lhs
        yield self.rhs


@dataclass
class Unary(Expr):
    op: str
    val: Expr

    def children(self):
        yield self.val


@dataclass
class Parens(Expr):
    inner: Expr

    def children(self):
        yield self.inner


@dataclass
class Step:
    """One step of a static traversal: ``.name`` or ``[literal]``."""
    kind: str                      # "attr" | "index"
    value: Union[str, int, float]
    legacy: bool = False           # ``.0`` index syntax inherited from HCL1
    line: int = 0
    col: int = 0


@dataclass
class ScopeTraversal(Expr):
    """A variable name followed by zero or more static steps, e.g.
    ``module.vpc[0].vpc_id`` (eks/main.tf:59 in the reference)."""
    root: str
    steps: list[Step]

    def dotted(self, n: Optional[int] = None) -> str:
        parts = [self.root]
        for s in self.steps[:n]:
            parts.append(f".{s.value}" if s.kind == "attr" else f"[{s.value!r}]")
        return "".join(parts)


@dataclass
class RelativeTraversal(Expr):
    source: Expr
    steps: list[Step]

    def children(self):
        yield self.source


@dataclass
class Index(Expr):
    coll: Expr
    key: Expr

    def children(self):
        yield self.coll
        yield self.key


@dataclass
class Splat(Expr):
    """``source.*.a.b`` (attribute-only, legacy) or ``source[*].a[0]`` (full)."""
    source: Expr
    steps: list[Step]
    full: bool = False

    def children(self):
        yield self.source


# --------------------------------------------------------------------------
# structure
# --------------------------------------------------------------------------
@dataclass
class Attribute(Node):
    name: str
    expr: Expr
    name_tok: int = 0            # index of the name token
    eq_tok: int = 0              # index of the '=' token
    expr_tok_start: int = 0      # token range of the expression itself
    expr_tok_end: int = 0

    def children(self):
        yield self.expr


@dataclass
class Label:
    value: str
    quoted: bool
    tok_start: int
    tok_end: int


@dataclass
class Block(Node):
    type: str
    labels: list[Label]
    body: "Body"
    one_line: bool = False
    type_tok: int = 0
    obrace_tok: int = 0
    cbrace_tok: int = 0

    def children(self):
        yield self.body

    @property
    def label_values(self) -> list[str]:
        return [l.value for l in self.labels]


@dataclass
class Body(Node):
    items: list[Union[Attribute, Block]] = field(default_factory=list)

    def children(self):
        return iter(self.items)

    @property
    def attributes(self) -> dict[str, Attribute]:
        out: dict[str, Attribute] = {}
        for it in self.items:
            if isinstance(it, Attribute):
                out.setdefault(it.name, it)
        return out

    @property
    def blocks(self) -> list[Block]:
        return [it for it in self.items if isinstance(it, Block)]

    def blocks_of(self, typ: str) -> list[Block]:
        return [b for b in self.blocks if b.type == typ]
