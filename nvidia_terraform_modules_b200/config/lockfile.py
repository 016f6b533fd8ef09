"""Dependency lock files (``.terraform.lock.hcl``) and their agreement with the
module's provider requirements.

The reference commits six lock files (eks/.terraform.lock.hcl and peers; SURVEY.md
section 1, layer L1).  ``terraform init`` refuses to proceed when a locked
version no longer satisfies the configuration's constraints, which is one of
the few ``init``-time failures that can be predicted offline.  Format restated
from Terraform's "Dependency Lock File" documentation: one ``provider "<fqn>"``
block per provider with ``version``, optional ``constraints`` and ``hashes``.
"""
from __future__ import annotations

import os
import re
from dataclasses import dataclass, field
from typing import Optional

from ..hcl import ast as A
from ..hcl.diagnostics import Diagnostic, ERROR, NOTICE, WARNING
from ..hcl.parser import parse_path
from . import typeexpr
from .model import Module
from .versions import Constraints, Version, VersionError, try_parse_constraints

LOCK_NAME = ".terraform.lock.hcl"
_HASH_RE = re.compile(r"^(h1:[A-Za-z0-9+/]{43}=|zh:[0-9a-f]{64})$")


@dataclass
class LockedProvider:
    fqn: str
    version: Optional[Version]
    constraints_text: Optional[str]
    hashes: list[str]
    line: int


@dataclass
class LockFile:
    path: str
    providers: dict[str, LockedProvider] = field(default_factory=dict)
    diags: list[Diagnostic] = field(default_factory=list)


def load_lockfile(directory: str, display_root: Optional[str] = None) -> Optional[LockFile]:
    path = os.path.join(directory, LOCK_NAME)
    if not os.path.isfile(path):
        return None
    disp = os.path.relpath(path, display_root) if display_root else path
    _, body, diags, _ = parse_path(path, disp)
    lock = LockFile(disp, diags=list(diags))

    def bad(summary: str, detail: str, node: A.Node) -> None:
        lock.diags.append(Diagnostic(ERROR, summary, detail, disp, node.pos.line, node.pos.col, "lock-file"))

    for item in body.items:
        if isinstance(item, A.Attribute):
            bad("Unsupported argument", f"An argument named \"{item.name}\" is not expected here.", item)
            continue
        if item.type != "provider" or len(item.labels) != 1:
            bad("Unsupported block type", f"Blocks of type \"{item.type}\" are not expected here.", item)
            continue
        fqn = item.labels[0].value.lower()      # provider addresses compare case-insensitively
        if len(fqn.split("/")) != 3:
            bad("Invalid provider source address",
                f"The provider source address \"{fqn}\" must be in the fully-qualified form hostname/namespace/type.",
                item)
            continue
        if fqn in lock.providers:
            bad("Duplicate provider lock", f"This lock file already contains a lock for provider {fqn}.", item)
            continue
        attrs = item.body.attributes
        version: Optional[Version] = None
        if "version" not in attrs:
            bad("Missing required argument", "The argument \"version\" is required, but no definition was found.", item)
        else:
            text = typeexpr.literal_value(attrs["version"].expr)
            try:
                version = Version.parse(text) if isinstance(text, str) else None
            except VersionError:
                version = None
            if version is None:
                bad("Invalid provider version number",
                    "The selected version number for this provider is not a valid version string.", attrs["version"])
        ctext = None
        if "constraints" in attrs:
            ctext = typeexpr.literal_value(attrs["constraints"].expr)
            if not isinstance(ctext, str) or try_parse_constraints(ctext)[0] is None:
                bad("Invalid provider version constraints",
                    "The recorded version constraints for this provider are not valid.", attrs["constraints"])
                ctext = None
        hashes: list[str] = []
        if "hashes" in attrs:
            hv = typeexpr.literal_value(attrs["hashes"].expr)
            if isinstance(hv, list) and all(isinstance(h, str) for h in hv):
                hashes = hv
                for h in hv:
                    if not _HASH_RE.match(h):
                        bad("Invalid provider hash", f"The string \"{h}\" is not a valid h1: or zh: package hash.",
                            attrs["hashes"])
            else:
                bad("Invalid provider hash set", "The \"hashes\" argument must be a set of strings.", attrs["hashes"])
        lock.providers[fqn] = LockedProvider(fqn, version, ctext, hashes, item.pos.line)
    return lock


def check_lock_consistency(m: Module, lock: Optional[LockFile],
                           children: tuple[Module, ...] = ()) -> list[Diagnostic]:
    """Compare a module's provider needs with its lock file.  ``children`` are
    the local-path child modules (transitively); their needs count too because
    one lock file covers the whole module tree."""
    diags: list[Diagnostic] = []
    if lock is None:
        return diags
    diags.extend(lock.diags)

    # what the module needs: explicit requirements + providers implied by
    # resource types and provider blocks (hashicorp/<name> unless declared)
    needed: dict[str, tuple[Optional[Constraints], str, int, str]] = {}
    for mod in (m,) + tuple(children):
        for name, req in mod.required_providers.items():
            if req.fqn in needed and needed[req.fqn][0] is not None and req.constraints is not None:
                both = Constraints(needed[req.fqn][0].terms + req.constraints.terms)
                needed[req.fqn] = (both,) + needed[req.fqn][1:]
            elif req.fqn not in needed or needed[req.fqn][0] is None:
                needed[req.fqn] = (req.constraints, req.file, req.line, name)
    for mod in (m,) + tuple(children):
        used = mod.used_provider_names()
        implied = set(used) | {p.name for p in mod.providers}
        for name in sorted(implied):
            if name in mod.required_providers or name == "terraform":
                continue
            fqn = f"registry.terraform.io/hashicorp/{name}"
            first = (used.get(name) or [None])[0]
            where = (first.file, first.line) if first else next(
                ((p.file, p.line) for p in mod.providers if p.name == name), ("", 0))
            needed.setdefault(fqn, (None, where[0], where[1], name))

    for fqn, (cons, file, line, name) in sorted(needed.items()):
        locked = lock.providers.get(fqn)
        if locked is None:
            diags.append(Diagnostic(
                ERROR, "Inconsistent dependency lock file",
                f"provider {fqn}: required by this configuration but no version is selected. "
                "To update the locked dependency selections to match a changed configuration, run: "
                "terraform init -upgrade", file, line, 0, "lock-missing"))
            continue
        if cons is not None and locked.version is not None and not cons.allows(locked.version):
            diags.append(Diagnostic(
                ERROR, "Inconsistent dependency lock file",
                f"provider {fqn}: locked version selection {locked.version} doesn't match the updated "
                f"version constraints \"{cons.canonical()}\". To update the locked dependency selections to "
                "match a changed configuration, run: terraform init -upgrade", file, line, 0, "lock-version"))
        elif cons is not None and locked.constraints_text is not None:
            recorded = try_parse_constraints(locked.constraints_text)[0]
            # registry modules add their own terms, so the lock may record MORE
            # than this module asks for; only a term the lock lacks is drift.
            if recorded is not None and not ({str(t) for t in cons.terms}
                                             <= {str(t) for t in recorded.terms}):
                diags.append(Diagnostic(
                    NOTICE, "Lock file records different constraints",
                    f"provider {fqn}: the lock file records constraints \"{locked.constraints_text}\" but the "
                    f"configuration now asks for \"{cons.canonical()}\"; the next \"terraform init\" will rewrite "
                    "the lock file.", lock.path, locked.line, 0, "lock-constraints"))
    for fqn, locked in sorted(lock.providers.items()):
        if fqn not in needed:
            diags.append(Diagnostic(
                NOTICE, "Locked provider is not required by the configuration",
                f"provider {fqn} {locked.version}: nothing in this module (or, for a root module with child "
                "modules, nothing visible offline) needs it; it may be a leftover of a removed resource or be "
                "required by a registry module.", lock.path, locked.line, 0, "lock-unused"))
        if not locked.hashes:
            diags.append(Diagnostic(WARNING, "Locked provider has no package hashes",
                                    f"provider {fqn}: without hashes the lock does not pin package contents.",
                                    lock.path, locked.line, 0, "lock-hashes"))
    return diags
