"""Command line: ``python -m nvidia_terraform_modules_b200 <command> ...``.

Sub-commands mirror the tools the reference's contribution guide names
(CONTRIBUTING.md:12-14) so a maintainer can swap them into a hook unchanged:

    fmt       [-check] [-diff] [-list=false] [-write=false] [-recursive] [TARGET...]
    validate  [-json] [DIR]
    lint      [--only RULES] [--skip RULES] [--json] [DIR...]
    lock      [DIR...]            lock file vs. required_providers
    docs      [--check|--write] [DIR]   terraform-docs style tables / README drift / update README in place
    check     [--json] [--strict] [ROOT]   everything, every module under ROOT
    report    [-o FILE] [ROOT]    markdown issue list of `check`

Exit codes: fmt as ``terraform fmt`` (0, 2 syntax error, 3 -check found diffs);
validate 1 when invalid; lint/lock/docs/check 1 when anything of severity
warning or above (check: error or unformatted file; with --strict also warnings)
was found.  A comment ``# tfcheck-ignore: rule-a, rule-b`` (or ``all``) silences lint
rules on its own line and the next one.
"""
from __future__ import annotations

import json
import os
import sys

from .check import STAND_IN, check_tree
from .config.lockfile import check_lock_consistency, load_lockfile
from .config.validate import validate_module
from .docs.tfdocs import InjectRefused, check_readme, find_readme, inject, render
from .fmt import command as fmt_command
from .hcl.diagnostics import ERROR, WARNING
from .lint.rules import RULES, run_rules
from .report import render_markdown


def _print_diags(diags, stream=sys.stderr) -> None:
    for d in diags:
        stream.write(d.format() + "\n\n")


def cmd_validate(argv: list[str]) -> int:
    as_json = any(a in ("-json", "--json") for a in argv)
    dirs = [a for a in argv if not a.startswith("-")] or ["."]
    res = validate_module(dirs[0], ".")
    errors = sum(1 for d in res.diags if d.severity == ERROR)
    warnings = sum(1 for d in res.diags if d.severity == WARNING)
    if as_json:
        print(json.dumps({
            "format_version": "1.0", "valid": res.valid, "error_count": errors, "warning_count": warnings,
            "stand_in": STAND_IN,
            "diagnostics": [{"severity": d.severity, "summary": d.summary, "detail": d.detail,
                             "range": {"filename": d.file, "start": {"line": d.line, "column": d.col}}}
                            for d in res.diags]}, indent=2))
    else:
        _print_diags(res.diags)
        if res.valid:
            print("Success! The configuration is valid" + (", but there were some validation warnings as shown "
                                                          "above." if warnings else "."))
        print(f"\n(note: {STAND_IN}; provider schemas were not consulted)")
    return 0 if res.valid else 1


def _split(arg: str) -> set[str]:
    return {x for x in arg.split(",") if x}


def cmd_lint(argv: list[str]) -> int:
    only = skip = None
    as_json = False
    dirs: list[str] = []
    it = iter(argv)
    for a in it:
        if a == "--only":
            only = _split(next(it, ""))
        elif a == "--skip":
            skip = _split(next(it, ""))
        elif a in ("--json", "-json"):
            as_json = True
        elif a == "--list-rules":
            for r in RULES:
                print(f"{r.id:32s} {r.severity:8s} {r.summary}")
            return 0
        else:
            dirs.append(a)
    found = []
    for d in dirs or ["."]:
        found.extend(run_rules(validate_module(d, "."), only, skip))
    if as_json:
        print(json.dumps([d.to_dict() for d in found], indent=2))
    else:
        _print_diags(found, sys.stdout)
        print(f"{len(found)} finding(s)")
    return 1 if any(d.severity in (ERROR, WARNING) for d in found) else 0


def cmd_lock(argv: list[str]) -> int:
    found = []
    for d in [a for a in argv if not a.startswith("-")] or ["."]:
        res = validate_module(d, ".")
        for mc in res.module.module_calls.values():
            res.validator.child(mc)
        kids = tuple(c for c in res.validator.children.values() if c is not None)
        lock = load_lockfile(d, ".")
        if lock is None:
            print(f"{d}: no .terraform.lock.hcl")
            continue
        found.extend(check_lock_consistency(res.module, lock, kids))
    _print_diags(found, sys.stdout)
    print(f"{len(found)} finding(s)")
    return 1 if any(d.severity in (ERROR, WARNING) for d in found) else 0


def cmd_docs(argv: list[str]) -> int:
    check = "--check" in argv
    dirs = [a for a in argv if not a.startswith("-")] or ["."]
    res = validate_module(dirs[0], ".")
    if "--write" in argv:
        readme = find_readme(dirs[0]) or os.path.join(dirs[0], "README.md")
        old = open(readme, encoding="utf-8").read() if os.path.isfile(readme) else ""
        try:
            new = inject(old, render(res.module, load_lockfile(dirs[0], ".")))
        except InjectRefused as e:
            sys.stderr.write(f"{readme}: not updated: {e}\n")
            return 1
        if new != old:
            with open(readme, "w", encoding="utf-8") as fh:
                fh.write(new)
            print(f"{readme} updated")
        return 0
    if not check:
        sys.stdout.write(render(res.module, load_lockfile(dirs[0], ".")))
        return 0
    readme = find_readme(dirs[0])
    if readme is None:
        print(f"{dirs[0]}: no README found")
        return 1
    found = check_readme(res.module, readme)
    _print_diags(found, sys.stdout)
    print(f"{len(found)} finding(s)")
    return 1 if found else 0


def cmd_check(argv: list[str]) -> int:
    as_json = any(a in ("-json", "--json") for a in argv)
    strict = "--strict" in argv
    roots = [a for a in argv if not a.startswith("-")] or ["."]
    rep = check_tree(os.path.abspath(roots[0]))
    if as_json:
        print(json.dumps(rep.to_dict(), indent=2))
    else:
        for p in rep.fmt.changed:
            print(f"fmt: {p} is not in canonical format")
        for m in rep.modules:
            _print_diags([d for d in m.all() if d.severity in (ERROR, WARNING)], sys.stdout)
        c = rep.to_dict()["counts"]
        print(f"{len(rep.modules)} modules, {len(rep.fmt.checked)} files, {rep.bytes_scanned} bytes in "
              f"{rep.seconds * 1000:.0f} ms: {c['fmt_differing']} unformatted, {c['error']} errors, "
              f"{c['warning']} warnings, {c['notice']} notices")
        print(f"(note: {STAND_IN})")
    if strict and rep.count(WARNING):
        return 1
    return 0 if rep.ok else 1


def cmd_report(argv: list[str]) -> int:
    out = None
    roots: list[str] = []
    it = iter(argv)
    for a in it:
        if a == "-o":
            out = next(it, None)
        else:
            roots.append(a)
    rep = check_tree(os.path.abspath(roots[0] if roots else "."))
    text = render_markdown(rep)
    if out:
        with open(out, "w", encoding="utf-8") as fh:
            fh.write(text)
    else:
        sys.stdout.write(text)
    return 0


COMMANDS = {"fmt": fmt_command.main, "validate": cmd_validate, "lint": cmd_lint, "lock": cmd_lock,
            "docs": cmd_docs, "check": cmd_check, "report": cmd_report}


def main(argv: list[str] | None = None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] in ("-h", "--help", "help"):
        print(__doc__)
        return 0
    cmd = COMMANDS.get(argv[0])
    if cmd is None:
        sys.stderr.write(f"unknown command {argv[0]!r}; try --help\n")
        return 1
    return cmd(argv[1:])


if __name__ == "__main__":
    sys.exit(main())
