"""Type-constraint expressions (``variable "x" { type = ... }``) and the check
that a literal ``default`` is compatible with them.

The reference declares 101 variables, e.g.
``type = list(any)`` at aks/variables.tf:98 and ``object({...})`` at
eks/variables.tf (existing_vpc_details).  Grammar and conversion rules restated
from Terraform's "Type Constraints" language documentation: primitives
``string``/``number``/``bool``, ``any``, ``list(T)``, ``set(T)``, ``map(T)``,
``object({k = T, ...})`` with ``optional(T[, default])`` attributes, and
``tuple([T, ...])``.  Primitive values convert between string and number/bool
when the text allows it; ``null`` is valid for every type.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional

from ..hcl import ast as A


class TypeError_(Exception):
    def __init__(self, message: str, node: A.Node):
        super().__init__(message)
        self.node = node


@dataclass
class Type:
    kind: str                                   # string number bool any list set map object tuple
    elem: Optional["Type"] = None               # list/set/map
    attrs: dict[str, "Type"] = field(default_factory=dict)    # object
    optional: set[str] = field(default_factory=set)           # object
    elems: list["Type"] = field(default_factory=list)         # tuple

    def __str__(self) -> str:
        if self.kind in ("list", "set", "map"):
            return f"{self.kind}({self.elem})"
        if self.kind == "object":
            inner = ", ".join(f"{k} = {'optional(' + str(v) + ')' if k in self.optional else v}"
                              for k, v in self.attrs.items())
            return "object({" + inner + "})"
        if self.kind == "tuple":
            return "tuple([" + ", ".join(map(str, self.elems)) + "])"
        return self.kind


class _Unknown:
    """A value that is not a literal (depends on a reference or a call)."""
    def __repr__(self):
        return "<unknown>"


UNKNOWN = _Unknown()


def parse_type(expr: A.Expr) -> Type:
    if isinstance(expr, A.ScopeTraversal) and not expr.steps:
        name = expr.root
        if name in ("string", "number", "bool", "any"):
            return Type(name)
        if name in ("list", "set", "map"):
            # bare collection keyword: 0.11 shorthand, still accepted
            return Type(name, elem=Type("any"))
        if name in ("object", "tuple"):
            raise TypeError_(f"The {name} type constructor requires one argument specifying the "
                             f"{'attribute types and values as a map' if name == 'object' else 'element types as a list'}.",
                             expr)
        raise TypeError_(f"The keyword \"{name}\" is not a valid type specification.", expr)
    if isinstance(expr, A.Template):
        raise TypeError_("A type specification is either a primitive type keyword (bool, number, string) "
                         "or a complex type constructor call, like list(string). Quoted type constraints "
                         "are not valid.", expr)
    if isinstance(expr, A.FunctionCall):
        name = expr.name
        if name in ("list", "set", "map"):
            if len(expr.args) != 1:
                raise TypeError_(f"The {name} type constructor requires one argument specifying the element type.", expr)
            return Type(name, elem=parse_type(expr.args[0]))
        if name == "object":
            if len(expr.args) != 1 or not isinstance(expr.args[0], A.ObjectCons):
                raise TypeError_("The object type constructor requires one argument specifying the "
                                 "attribute types and values as a map.", expr)
            t = Type("object")
            for key, val in expr.args[0].items:
                k = key.literal_key()
                if k is None:
                    raise TypeError_("Object constructor map keys must be attribute names.", key)
                if isinstance(val, A.FunctionCall) and val.name == "optional":
                    if not 1 <= len(val.args) <= 2:
                        raise TypeError_("Optional attribute modifier requires the attribute type as its "
                                         "first argument, and optionally a default value as its second.", val)
                    t.attrs[k] = parse_type(val.args[0])
                    t.optional.add(k)
                else:
                    t.attrs[k] = parse_type(val)
            return t
        if name == "tuple":
            if len(expr.args) != 1 or not isinstance(expr.args[0], A.TupleCons):
                raise TypeError_("The tuple type constructor requires one argument specifying the "
                                 "element types as a list.", expr)
            return Type("tuple", elems=[parse_type(a) for a in expr.args[0].items])
        if name == "optional":
            raise TypeError_("Keyword \"optional\" is valid only as a modifier for object type attributes.", expr)
        raise TypeError_(f"Keyword \"{name}\" is not a valid type constructor.", expr)
    raise TypeError_("A type specification is either a primitive type keyword (bool, number, string) "
                     "or a complex type constructor call, like list(string).", expr)


def literal_value(expr: A.Expr):
    """Python value of an expression made only of literals; UNKNOWN otherwise.
    Objects become dicts, tuples become lists."""
    if isinstance(expr, A.Literal):
        return expr.value
    if isinstance(expr, A.Template):
        v = expr.literal_value()
        return UNKNOWN if v is None else v
    if isinstance(expr, A.Parens):
        return literal_value(expr.inner)
    if isinstance(expr, A.Unary) and expr.op == "-":
        v = literal_value(expr.val)
        if isinstance(v, (int, float)) and not isinstance(v, bool):
            return -v
        return UNKNOWN
    if isinstance(expr, A.TupleCons):
        return [literal_value(i) for i in expr.items]
    if isinstance(expr, A.ObjectCons):
        out = {}
        for k, v in expr.items:
            name = k.literal_key()
            if name is None:
                return UNKNOWN
            out[name] = literal_value(v)
        return out
    return UNKNOWN


def _is_num(v) -> bool:
    return isinstance(v, (int, float)) and not isinstance(v, bool)


def conformance_error(value, typ: Type, path: str = "") -> Optional[str]:
    """Why ``value`` cannot convert to ``typ`` -- None when it can (or when the
    value is not a literal and so cannot be judged)."""
    where = f"{path}: " if path else ""
    if value is None or value is UNKNOWN or typ.kind == "any":
        return None
    k = typ.kind
    if k == "string":
        if isinstance(value, (str, bool)) or _is_num(value):
            return None
        return f"{where}string required"
    if k == "number":
        if _is_num(value):
            return None
        if isinstance(value, str):
            try:
                float(value)
                return None
            except ValueError:
                return f"{where}a number is required"
        return f"{where}number required"
    if k == "bool":
        if isinstance(value, bool) or (isinstance(value, str) and value in ("true", "false")):
            return None
        return f"{where}bool required"
    if k in ("list", "set"):
        if not isinstance(value, list):
            return f"{where}{k} of {typ.elem} required"
        for i, v in enumerate(value):
            err = conformance_error(v, typ.elem, f"{path}[{i}]" if path else f"element {i}")
            if err:
                return err
        return None
    if k == "map":
        if not isinstance(value, dict):
            return f"{where}map of {typ.elem} required"
        for key, v in value.items():
            err = conformance_error(v, typ.elem, f'{path}["{key}"]' if path else f'element "{key}"')
            if err:
                return err
        return None
    if k == "object":
        if not isinstance(value, dict):
            return f"{where}object required"
        for name, t in typ.attrs.items():
            if name not in value:
                if name in typ.optional:
                    continue
                return f"{where}attribute \"{name}\" is required"
            err = conformance_error(value[name], t, f"{path}.{name}" if path else f"attribute \"{name}\"")
            if err:
                return err
        return None
    if k == "tuple":
        if not isinstance(value, list):
            return f"{where}tuple required"
        if len(value) != len(typ.elems):
            return f"{where}tuple required with {len(typ.elems)} elements"
        for i, (v, t) in enumerate(zip(value, typ.elems)):
            err = conformance_error(v, t, f"{path}[{i}]" if path else f"element {i}")
            if err:
                return err
        return None
    return None
