"""Version numbers and version-constraint strings as Terraform reads them.

Used for ``required_version`` (eks/versions.tf:16, gke/versions.tf:15,
aks/versions.tf:16 in the reference), ``required_providers`` entries
(eks/versions.tf:6-15), module ``version`` arguments (eks/main.tf:9,53) and the
``version``/``constraints`` pairs of dependency lock files
(eks/.terraform.lock.hcl).  Syntax restated from Terraform's "Version
Constraints" language documentation: comma-separated terms, each an optional
operator (``=``, ``!=``, ``>``, ``>=``, ``<``, ``<=``, ``~>``) and a version.
"""
from __future__ import annotations

import re
from dataclasses import dataclass
from functools import total_ordering
from typing import Optional

_VERSION_RE = re.compile(
    r"^v?(\d+)(?:\.(\d+))?(?:\.(\d+))?(?:-([0-9A-Za-z.-]+))?(?:\+([0-9A-Za-z.-]+))?$")
_TERM_RE = re.compile(r"^\s*(~>|>=|<=|!=|=|>|<)?\s*(\S+)\s*$")


class VersionError(ValueError):
    pass


@total_ordering
@dataclass(frozen=True)
class Version:
    major: int
    minor: int = 0
    patch: int = 0
    pre: str = ""
    segments: int = 3        # how many numeric segments were written (matters for ~>)

    @classmethod
    def parse(cls, text: str) -> "Version":
        m = _VERSION_RE.match(text.strip())
        if not m:
            raise VersionError(f"Malformed version: {text!r}")
        segs = 1 + (m.group(2) is not None) + (m.group(3) is not None)
        return cls(int(m.group(1)), int(m.group(2) or 0), int(m.group(3) or 0), m.group(4) or "", segs)

    def _key(self):
        # a pre-release sorts before the release it precedes
        pre_key: tuple = (1,) if not self.pre else (0, tuple(
            (0, int(p)) if p.isdigit() else (1, p) for p in self.pre.split(".")))
        return (self.major, self.minor, self.patch, pre_key)

    def __eq__(self, other):
        return isinstance(other, Version) and self._key() == other._key()

    def __lt__(self, other):
        return self._key() < other._key()

    def __hash__(self):
        return hash(self._key())

    def __str__(self):
        s = f"{self.major}.{self.minor}.{self.patch}"
        return s + (f"-{self.pre}" if self.pre else "")


@dataclass(frozen=True)
class Term:
    op: str
    version: Version

    def allows(self, v: Version) -> bool:
        w = self.version
        if v.pre and not w.pre and self.op != "=":
            # pre-releases only match exact constraints
            return False
        if self.op == "=":
            return v == w
        if self.op == "!=":
            return v != w
        if self.op == ">":
            return v > w
        if self.op == ">=":
            return v >= w
        if self.op == "<":
            return v < w
        if self.op == "<=":
            return v <= w
        if self.op == "~>":
            if v < w:
                return False
            if w.segments >= 3:
                return v.major == w.major and v.minor == w.minor
            if w.segments == 2:
                return v.major == w.major  # ~> 1.2 pins the major
            return True                    # ~> 1 has no segment left to pin
        raise AssertionError(self.op)

    def __str__(self):
        text = f"{self.version.major}"
        if self.version.segments >= 2:
            text += f".{self.version.minor}"
        if self.version.segments >= 3:
            text += f".{self.version.patch}"
        if self.version.pre:
            text += f"-{self.version.pre}"
        return text if self.op == "=" else f"{self.op} {text}"


@dataclass(frozen=True)
class Constraints:
    terms: tuple[Term, ...]

    @classmethod
    def parse(cls, text: str) -> "Constraints":
        if not text.strip():
            raise VersionError("Empty version constraint")
        terms = []
        for part in text.split(","):
            m = _TERM_RE.match(part)
            if not m:
                raise VersionError(f"Malformed constraint: {part.strip()!r}")
            terms.append(Term(m.group(1) or "=", Version.parse(m.group(2))))
        return cls(tuple(terms))

    def allows(self, v: Version) -> bool:
        return all(t.allows(v) for t in self.terms)

    def canonical(self) -> str:
        """The spelling ``terraform init`` records in a lock file's
        ``constraints`` argument (terms normalised and sorted)."""
        return ", ".join(sorted({str(t) for t in self.terms}))

    def is_exact(self) -> bool:
        return len(self.terms) == 1 and self.terms[0].op == "="

    def upper_bounded(self) -> bool:
        return any(t.op in ("=", "~>", "<", "<=") for t in self.terms)


def try_parse_constraints(text: str) -> tuple[Optional[Constraints], str]:
    try:
        return Constraints.parse(text), ""
    except VersionError as e:
        return None, str(e)
