"""Which Terraform release ``validate`` restates, and what each release knows.

``terraform validate`` is not one behaviour: the language grew between releases
(``moved`` blocks in 1.1, ``optional()`` defaults in 1.3, ``check`` in 1.5, variable
validation that may look at other objects in 1.9, ...).  The reference pins three
different floors -- ``>= 1.2.4`` (eks/versions.tf:16), ``>= 0.14`` (gke/versions.tf:15)
and ``>= 1.3.4`` (aks/versions.tf:16) -- so a stand-in has to say which release it is
standing in for, and behave like that one release throughout.

One run restates exactly ONE release, chosen in this order:

1. ``--terraform-version X.Y`` on the command line;
2. the lowest release the module's own ``required_version`` admits -- the oldest
   Terraform its authors promise it works with, so the strictest reading;
3. ``NEWEST`` when the module states nothing.

The choice is applied to the whole configuration, child modules included, just as one
Terraform binary reads every module of a configuration.  Releases older than ``OLDEST``
are read as ``OLDEST`` and releases newer than ``NEWEST`` as ``NEWEST`` (nothing newer
is restated here); both clamps are said out loud in the result.

Release numbers below are restated from Terraform's published changelogs ("NEW
FEATURES" / "ENHANCEMENTS" of each v1.x.0) and upgrade guides.  PARITY UNPINNED like
the rest of ``validate``: no binary of any of these releases
is available offline to confirm them.
"""
from __future__ import annotations

import re
from dataclasses import dataclass
from typing import Iterable, Optional

from .versions import Constraints, Version

Release = tuple[int, int]

# every minor series from the first HCL2 release on; the tail beyond NEWEST exists
# only so that a floor such as ">= 1.12" can be recognised (and then clamped)
SERIES: tuple[Release, ...] = ((0, 12), (0, 13), (0, 14), (0, 15)) + tuple((1, n) for n in range(0, 16))
OLDEST: Release = (0, 13)
NEWEST: Release = (1, 10)

# feature -> first release that has it
SINCE: dict[str, Release] = {
    # top-level blocks
    "block:moved": (1, 1),
    "block:import": (1, 5),
    "block:check": (1, 5),
    "block:removed": (1, 7),
    "block:ephemeral": (1, 10),
    # arguments and nested blocks
    "variable:sensitive": (0, 14),
    "variable:nullable": (1, 1),
    "variable:ephemeral": (1, 10),
    "output:ephemeral": (1, 10),
    "output:precondition": (1, 2),
    "lifecycle:precondition": (1, 2),
    "lifecycle:postcondition": (1, 2),
    "lifecycle:replace_triggered_by": (1, 2),
    "terraform:cloud": (1, 1),
    "required_providers:configuration_aliases": (0, 15),
    "import:for_each": (1, 7),
    "removed:provisioner": (1, 9),
    # language behaviour
    "optional-attributes": (1, 3),            # object({ a = optional(string, "x") }) without an experiment
    "validation-refers-to-other-objects": (1, 9),
    "provider-defined-functions": (1, 8),     # provider::ns::fn(...)
    "terraform.applying": (1, 10),
}

# built-in functions that arrived after 0.13 (everything else in functions.FUNCTIONS
# was already there)
FUNCTION_SINCE: dict[str, Release] = {
    "alltrue": (0, 14), "anytrue": (0, 14), "textdecodebase64": (0, 14), "textencodebase64": (0, 14),
    "one": (0, 15), "sensitive": (0, 15), "nonsensitive": (0, 15),
    "startswith": (1, 3), "endswith": (1, 3), "timecmp": (1, 3),
    "strcontains": (1, 5), "plantimestamp": (1, 5),
    "issensitive": (1, 8),
    "templatestring": (1, 9),
    "ephemeralasnull": (1, 10),
}
# list() and map() still worked, with a deprecation warning, until 0.15 removed them
LEGACY_FUNCTIONS_REMOVED: Release = (0, 15)


class TerraformVersionError(ValueError):
    pass


def text(v: Release) -> str:
    return f"{v[0]}.{v[1]}"


def parse_release(s: str) -> Release:
    """``"1.9"``, ``"v1.9"`` or ``"1.9.3"`` -> ``(1, 9)``."""
    m = re.match(r"^v?(\d+)\.(\d+)(?:\.\d+)?(?:-[0-9A-Za-z.-]+)?$", s.strip())
    if not m:
        raise TerraformVersionError(f"not a Terraform version: {s!r} (expected something like 1.5 or 1.5.7)")
    return int(m.group(1)), int(m.group(2))


def lowest_admitted(constraints: Iterable[Constraints]) -> Optional[Release]:
    """The oldest release series holding a version that every constraint allows,
    or None when no lower bound can be read off (``< 2.0``, or constraints that
    contradict each other)."""
    cons = list(constraints)
    if not any(t.op in (">", ">=", "=", "~>") for c in cons for t in c.terms):
        return None                        # nothing bounds it from below
    for major, minor in SERIES:
        for patch in range(0, 40):
            if all(c.allows(Version(major, minor, patch)) for c in cons):
                return major, minor
    return None


@dataclass(frozen=True)
class Restated:
    """The release one run restates, and why."""
    release: Release
    source: str                    # human-readable: where the choice came from
    note: str = ""                 # set when the choice had to be clamped
    requested: Optional[Release] = None     # what --terraform-version named, before any clamp

    def has(self, feature: str) -> bool:
        return self.release >= SINCE[feature]

    def has_function(self, name: str) -> bool:
        return self.release >= FUNCTION_SINCE.get(name, OLDEST)

    @property
    def text(self) -> str:
        return text(self.release)

    def describe(self) -> str:
        s = f"Terraform {self.text} ({self.source})"
        return s + (f"; {self.note}" if self.note else "")


def _clamp(release: Release, source: str) -> Restated:
    if release < OLDEST:
        return Restated(OLDEST, source, f"releases before {text(OLDEST)} are not restated; read as {text(OLDEST)}")
    if release > NEWEST:
        return Restated(NEWEST, source, f"nothing newer than {text(NEWEST)} is restated; read as {text(NEWEST)}")
    return Restated(release, source)


def from_flag(value: str) -> Restated:
    asked = parse_release(value)
    r = _clamp(asked, "--terraform-version")
    return Restated(r.release, r.source, r.note, asked)


def series_admitted(release: Release, constraints: Iterable[Constraints]) -> bool:
    """Does any patch release of ``release`` satisfy every constraint?"""
    cons = list(constraints)
    return any(all(c.allows(Version(release[0], release[1], patch)) for c in cons) for patch in range(0, 40))


def from_required_version(entries: list[tuple[str, str]]) -> Restated:
    """``entries`` = ``(constraint text, "file:line")`` of every ``required_version``
    in the module.  Falls back to NEWEST when they give no floor."""
    cons = []
    for constraint_text, _ in entries:
        try:
            cons.append(Constraints.parse(constraint_text))
        except ValueError:
            continue
    low = lowest_admitted(cons)
    if low is None:
        why = "no required_version in the module" if not entries else \
            f"required_version \"{entries[0][0]}\" at {entries[0][1]} sets no floor"
        return Restated(NEWEST, f"default: {why}")
    first_text, first_where = entries[0]
    return _clamp(low, f"lowest release admitted by required_version \"{first_text}\" at {first_where}")
