"""Terraform's built-in function table: name -> (fixed parameter count, variadic?).

``terraform validate`` rejects calls to unknown functions and calls with the
wrong number of arguments without needing any provider schema, so this check is
within reach offline.  The reference calls ``merge`` (eks/main.tf:85), ``concat``
(gke/main.tf:83), ``length``/``one`` (gke/main.tf:40), ``jsonencode``
(eks/examples/cnpack/aws-amp.tf:48), ``jsondecode``
(aks/examples/cnpack/azure-prometheus.tf:54) and ``base64decode``
(eks/provider.tf:7).  Signatures restated from Terraform's "Built-in Functions"
documentation.
"""
from __future__ import annotations

from typing import Optional

# name: (number of fixed parameters, accepts further variadic arguments)
FUNCTIONS: dict[str, tuple[int, bool]] = {
    # numeric
    "abs": (1, False), "ceil": (1, False), "floor": (1, False), "log": (2, False),
    "max": (0, True), "min": (0, True), "parseint": (2, False), "pow": (2, False),
    "signum": (1, False), "sum": (1, False),
    # string
    "chomp": (1, False), "endswith": (2, False), "format": (1, True), "formatlist": (1, True),
    "indent": (2, False), "join": (1, True), "lower": (1, False), "regex": (2, False),
    "regexall": (2, False), "replace": (3, False), "split": (2, False), "startswith": (2, False),
    "strcontains": (2, False), "strrev": (1, False), "substr": (3, False),
    "templatestring": (2, False), "title": (1, False), "trim": (2, False),
    "trimprefix": (2, False), "trimspace": (1, False), "trimsuffix": (2, False), "upper": (1, False),
    # collection
    "alltrue": (1, False), "anytrue": (1, False), "chunklist": (2, False), "coalesce": (0, True),
    "coalescelist": (0, True), "compact": (1, False), "concat": (0, True), "contains": (2, False),
    "distinct": (1, False), "element": (2, False), "flatten": (1, False), "index": (2, False),
    "keys": (1, False), "length": (1, False), "lookup": (2, True), "matchkeys": (3, False),
    "merge": (0, True), "one": (1, False), "range": (0, True), "reverse": (1, False),
    "setintersection": (1, True), "setproduct": (0, True), "setsubtract": (2, False),
    "setunion": (1, True), "slice": (3, False), "sort": (1, False), "transpose": (1, False),
    "values": (1, False), "zipmap": (2, False),
    # encoding
    "base64decode": (1, False), "base64encode": (1, False), "base64gzip": (1, False),
    "csvdecode": (1, False), "jsondecode": (1, False), "jsonencode": (1, False),
    "textdecodebase64": (2, False), "textencodebase64": (2, False), "urlencode": (1, False),
    "yamldecode": (1, False), "yamlencode": (1, False),
    # filesystem
    "abspath": (1, False), "basename": (1, False), "dirname": (1, False), "file": (1, False),
    "filebase64": (1, False), "fileexists": (1, False), "fileset": (2, False),
    "pathexpand": (1, False), "templatefile": (2, False),
    # date and time
    "formatdate": (2, False), "plantimestamp": (0, False), "timeadd": (2, False),
    "timecmp": (2, False), "timestamp": (0, False),
    # hash and crypto
    "base64sha256": (1, False), "base64sha512": (1, False), "bcrypt": (1, True),
    "filebase64sha256": (1, False), "filebase64sha512": (1, False), "filemd5": (1, False),
    "filesha1": (1, False), "filesha256": (1, False), "filesha512": (1, False), "md5": (1, False),
    "rsadecrypt": (2, False), "sha1": (1, False), "sha256": (1, False), "sha512": (1, False),
    "uuid": (0, False), "uuidv5": (2, False),
    # ip network
    "cidrhost": (2, False), "cidrnetmask": (1, False), "cidrsubnet": (3, False),
    "cidrsubnets": (1, True),
    # type conversion
    "can": (1, False), "ephemeralasnull": (1, False), "issensitive": (1, False),
    "nonsensitive": (1, False), "sensitive": (1, False), "tobool": (1, False),
    "tolist": (1, False), "tomap": (1, False), "tonumber": (1, False), "toset": (1, False),
    "tostring": (1, False), "try": (0, True), "type": (1, False),
}

# removed in Terraform 0.12+, with the advice Terraform prints
REMOVED: dict[str, str] = {
    "list": "The \"list\" function was deprecated in Terraform v0.12 and is no longer available. "
            "To construct a list from a collection of values, use tolist([ ... ]) syntax.",
    "map": "The \"map\" function was deprecated in Terraform v0.12 and is no longer available. "
           "To construct a map from a collection of key/value pairs, use tomap({ ... }) syntax.",
}


def lookup(name: str) -> Optional[tuple[int, bool]]:
    if name.startswith("core::"):
        name = name[len("core::"):]
    return FUNCTIONS.get(name)


def arity_error(name: str, nargs: int, expanded: bool) -> Optional[tuple[str, str]]:
    """(summary, detail) when ``nargs`` arguments cannot satisfy ``name``."""
    sig = lookup(name)
    if sig is None:
        return None
    fixed, variadic = sig
    if expanded:
        return None        # f(xs...) -- count only known at run time
    if nargs < fixed:
        return ("Not enough function arguments",
                f"Function \"{name}\" expects {fixed} argument(s). Missing value for parameter {nargs + 1}.")
    if nargs > fixed and not variadic:
        return ("Too many function arguments",
                f"Function \"{name}\" expects only {fixed} argument(s).")
    return None
