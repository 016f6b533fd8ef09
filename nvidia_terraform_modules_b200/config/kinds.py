"""Static "kind" of an expression, where it can be known without evaluation.

``terraform validate`` type-checks expressions against unknown values; most of
that needs provider schemas, but a few mistakes are visible from literals alone:
the two arms of a conditional that can never share a type, arithmetic on a
collection, logic on a number.  Rules restated from Terraform's "Conditional
Expressions" and "Arithmetic and Logical Operators" documentation.  Everything
here is deliberately conservative: ``None`` means "cannot tell", and no check
fires unless both sides are known.
"""
from __future__ import annotations

from typing import Optional

from ..hcl import ast as A

PRIMITIVES = {"string", "number", "bool"}

# result kind of built-in functions whose result type never depends on the input
_FUNCTION_KINDS = {
    "length": "number", "abs": "number", "ceil": "number", "floor": "number", "max": "number", "min": "number",
    "parseint": "number", "pow": "number", "signum": "number", "sum": "number", "index": "number", "log": "number",
    "tonumber": "number", "timecmp": "number",
    "upper": "string", "lower": "string", "title": "string", "trimspace": "string", "trim": "string",
    "trimprefix": "string", "trimsuffix": "string", "chomp": "string", "format": "string", "join": "string",
    "replace": "string", "substr": "string", "indent": "string", "strrev": "string", "jsonencode": "string",
    "yamlencode": "string", "base64encode": "string", "base64decode": "string", "urlencode": "string",
    "file": "string", "templatefile": "string", "md5": "string", "sha1": "string", "sha256": "string",
    "sha512": "string", "uuid": "string", "timestamp": "string", "formatdate": "string", "basename": "string",
    "dirname": "string", "abspath": "string", "pathexpand": "string", "cidrhost": "string",
    "cidrnetmask": "string", "cidrsubnet": "string", "tostring": "string", "bcrypt": "string",
    "contains": "bool", "startswith": "bool", "endswith": "bool", "strcontains": "bool", "alltrue": "bool",
    "anytrue": "bool", "can": "bool", "fileexists": "bool", "tobool": "bool", "issensitive": "bool",
    "concat": "list", "compact": "list", "distinct": "list", "flatten": "list", "reverse": "list",
    "sort": "list", "split": "list", "keys": "list", "values": "list", "range": "list", "slice": "list",
    "chunklist": "list", "formatlist": "list", "cidrsubnets": "list", "regexall": "list", "tolist": "list",
    "coalescelist": "list", "setproduct": "list",
    "toset": "set", "setunion": "set", "setintersection": "set", "setsubtract": "set", "fileset": "set",
    "merge": "map", "zipmap": "map", "tomap": "map", "transpose": "map",
}
_SEQUENCES = {"tuple", "list", "set"}
_MAPPINGS = {"object", "map"}


def kind_of(e: A.Node) -> Optional[str]:
    if isinstance(e, A.Parens):
        return kind_of(e.inner)
    if isinstance(e, A.Literal):
        if e.value is None:
            return "null"
        return "bool" if isinstance(e.value, bool) else "number"
    if isinstance(e, A.Template):
        if e.is_single_interpolation():
            return kind_of(e.parts[0])       # "${x}" passes x through unchanged
        return "string"
    if isinstance(e, A.TupleCons):
        return "tuple"
    if isinstance(e, A.ObjectCons):
        return "object"
    if isinstance(e, A.ForExpr):
        return "object" if e.key_expr is not None else "tuple"
    if isinstance(e, A.FunctionCall):
        return _FUNCTION_KINDS.get(e.name)
    if isinstance(e, A.Unary):
        return "bool" if e.op == "!" else "number"
    if isinstance(e, A.Binary):
        return "number" if e.op in ("+", "-", "*", "/", "%") else "bool"
    if isinstance(e, A.Conditional):
        a, b = kind_of(e.true), kind_of(e.false)
        if a == b:
            return a
        return None
    return None


def family(kind: Optional[str]) -> Optional[str]:
    if kind in PRIMITIVES:
        return "primitive"
    if kind in _SEQUENCES:
        return "sequence"
    if kind in _MAPPINGS:
        return "mapping"
    return None            # unknown or null: compatible with everything


def conditional_mismatch(e: A.Conditional) -> Optional[tuple[str, str]]:
    """(true kind, false kind) when the arms can never be unified."""
    a, b = kind_of(e.true), kind_of(e.false)
    fa, fb = family(a), family(b)
    if fa and fb and fa != fb:
        return a, b
    if {a, b} == {"number", "bool"}:
        return a, b            # the only primitive pair with no common type
    return None


def operand_problem(e: A.Node) -> Optional[tuple[str, A.Node, str]]:
    """(side, operand, required type) for an operator applied to a literal of
    the wrong family."""
    if isinstance(e, A.Binary):
        if e.op in ("+", "-", "*", "/", "%", "<", ">", "<=", ">="):
            need = "number"
        elif e.op in ("&&", "||"):
            need = "bool"
        else:
            return None        # == and != accept anything
        for side, operand in (("left", e.lhs), ("right", e.rhs)):
            k = kind_of(operand)
            if _clashes(k, need, operand):
                return side, operand, need
    elif isinstance(e, A.Unary):
        need = "bool" if e.op == "!" else "number"
        k = kind_of(e.val)
        if _clashes(k, need, e.val):
            return "", e.val, need
    return None


def _clashes(kind: Optional[str], need: str, operand: A.Node) -> bool:
    if kind is None or kind == "null":
        return False
    if family(kind) in ("sequence", "mapping"):
        return True
    if kind == need:
        return False
    if kind == "string":
        # a literal string converts when its text allows it; anything else is unknown at this point
        inner = operand.inner if isinstance(operand, A.Parens) else operand
        text = inner.literal_value() if isinstance(inner, A.Template) else None
        if text is None:
            return False
        if need == "number":
            try:
                float(text)
                return False
            except ValueError:
                return True
        return text not in ("true", "false")
    return True                # number where bool is needed, or the reverse
