"""Command line: ``python -m nvidia_terraform_modules_b200 <command> ...``.

Sub-commands mirror the tools the reference's contribution guide names
(CONTRIBUTING.md:12-14) so a maintainer can swap them into a hook unchanged:

    fmt       [-check] [-diff] [-list=false] [-write=false] [-recursive] [TARGET...]
    validate  [-json] [--terraform-version X.Y] [DIR]
    lint      [--only RULES] [--skip RULES] [--json] [--terraform-version X.Y] [DIR...]
    lock      [DIR...]            lock file vs. required_providers
    docs      [--check|--write] [DIR]   terraform-docs style tables / README drift / update README in place
    check     [--json] [--strict] [--terraform-version X.Y] [ROOT]   everything, every module under ROOT
    report    [-o FILE] [--terraform-version X.Y] [ROOT]    markdown issue list of `check`

``--terraform-version X.Y`` names the ONE Terraform release whose rules ``validate``
restates for the run (0.13 ... 1.10).  Without it each module is read under the oldest
release its own ``required_version`` admits, or 1.10 when it states none; the choice
is printed with every result.

File names in messages are relative to the directory given on the command line (a
module's own files appear as ``main.tf``, like Terraform run inside that directory);
``-json``/``--json`` output carries the absolute path as well.

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
from typing import Optional

from .check import STAND_IN, check_tree
from .config import tfversion
from .config.lockfile import check_lock_consistency, load_lockfile
from .config.validate import validate_module
from .docs.tfdocs import InjectRefused, check_readme, find_readme, inject, render
from .fmt import command as fmt_command
from .hcl.diagnostics import Diagnostic, ERROR, WARNING
from .lint.rules import RULES, run_rules
from .report import render_markdown


def _print_diags(diags, stream=sys.stderr) -> None:
    for d in diags:
        stream.write(d.format() + "\n\n")


class UsageError(Exception):
    pass


def _options(argv: list[str], flags: tuple[str, ...] = (), valued: tuple[str, ...] = ()) -> tuple[dict, list[str]]:
    """Split ``argv`` into options and positional arguments.  Options may be written
    with one dash or two, and valued ones as ``--name value`` or ``--name=value``."""
    opts: dict = {}
    rest: list[str] = []
    it = iter(argv)
    for a in it:
        if not a.startswith("-") or a == "-":
            rest.append(a)
            continue
        name, eq, val = a.lstrip("-").partition("=")
        if name == "no-color":
            continue
        if name in flags and not eq:
            opts[name] = True
        elif name in valued:
            if not eq:
                val = next(it, None)
                if val is None:
                    raise UsageError(f"option --{name} needs a value")
            opts[name] = val
        else:
            raise UsageError(f"unknown option {a!r}")
    if "terraform-version" in opts:
        try:
            tfversion.from_flag(opts["terraform-version"])
        except tfversion.TerraformVersionError as e:
            raise UsageError(str(e))
    return opts, rest


def _display_base(dirs: list[str]) -> str:
    """What file names are shown relative to: the directory argument itself, or the
    deepest directory holding all of them when several were given (so that two modules'
    ``main.tf`` stay distinguishable) -- never a path that climbs out with ``..``."""
    if len(dirs) == 1:
        return dirs[0]
    return os.path.commonpath([os.path.abspath(d) for d in dirs])


def _to_dict(d: Diagnostic, base: str) -> dict:
    """A diagnostic for machines: ``file`` as shown to people (relative to the
    directory argument) plus ``abs_file``."""
    out = d.to_dict()
    out["abs_file"] = os.path.normpath(os.path.join(os.path.abspath(base), d.file)) if d.file else ""
    return out


def cmd_validate(argv: list[str]) -> int:
    opts, dirs = _options(argv, flags=("json",), valued=("terraform-version",))
    directory = dirs[0] if dirs else "."
    res = validate_module(directory, directory, opts.get("terraform-version"))
    errors = sum(1 for d in res.diags if d.severity == ERROR)
    warnings = sum(1 for d in res.diags if d.severity == WARNING)
    restated = res.module.restated
    if opts.get("json"):
        print(json.dumps({
            "format_version": "1.0", "valid": res.valid, "error_count": errors, "warning_count": warnings,
            "stand_in": STAND_IN, "terraform_version": restated.text,
            "terraform_version_source": restated.describe(),
            "diagnostics": [{"severity": d.severity, "summary": d.summary, "detail": d.detail,
                             "range": {"filename": d.file,
                                       "abs_filename": _to_dict(d, directory)["abs_file"],
                                       "start": {"line": d.line, "column": d.col}}}
                            for d in res.diags]}, indent=2))
    else:
        _print_diags(res.diags)
        if res.valid:
            print("Success! The configuration is valid" + (", but there were some validation warnings as shown "
                                                          "above." if warnings else "."))
        print(f"\n(note: {STAND_IN}; this run restated {restated.describe()}; "
              "provider schemas were not consulted)")
    return 0 if res.valid else 1


def _split(arg: str) -> set[str]:
    return {x for x in arg.split(",") if x}


def cmd_lint(argv: list[str]) -> int:
    opts, dirs = _options(argv, flags=("json", "list-rules"), valued=("only", "skip", "terraform-version"))
    if opts.get("list-rules"):
        for r in RULES:
            print(f"{r.id:32s} {r.severity:8s} {r.summary}")
        return 0
    only = _split(opts["only"]) if "only" in opts else None
    skip = _split(opts["skip"]) if "skip" in opts else None
    found = []
    dirs = dirs or ["."]
    base = _display_base(dirs)
    for d in dirs:
        found.extend(run_rules(validate_module(d, base, opts.get("terraform-version")), only, skip))
    as_dicts = [_to_dict(x, base) for x in found]
    if opts.get("json"):
        print(json.dumps(as_dicts, indent=2))
    else:
        _print_diags(found, sys.stdout)
        print(f"{len(found)} finding(s)")
    return 1 if any(d.severity in (ERROR, WARNING) for d in found) else 0


def cmd_lock(argv: list[str]) -> int:
    _, dirs = _options(argv)
    found = []
    dirs = dirs or ["."]
    base = _display_base(dirs)
    for d in dirs:
        res = validate_module(d, base)
        for mc in res.module.module_calls.values():
            res.validator.child(mc)
        kids = tuple(c for c in res.validator.children.values() if c is not None)
        lock = load_lockfile(d, base)
        if lock is None:
            print(f"{d}: no .terraform.lock.hcl")
            continue
        found.extend(check_lock_consistency(res.module, lock, kids))
    _print_diags(found, sys.stdout)
    print(f"{len(found)} finding(s)")
    return 1 if any(d.severity in (ERROR, WARNING) for d in found) else 0


def cmd_docs(argv: list[str]) -> int:
    opts, dirs = _options(argv, flags=("check", "write"))
    directory = dirs[0] if dirs else "."
    res = validate_module(directory, directory)
    if opts.get("write"):
        readme = find_readme(directory) or os.path.join(directory, "README.md")
        old = open(readme, encoding="utf-8").read() if os.path.isfile(readme) else ""
        try:
            new = inject(old, render(res.module, load_lockfile(directory, directory)))
        except InjectRefused as e:
            sys.stderr.write(f"{readme}: not updated: {e}\n")
            return 1
        if new != old:
            with open(readme, "w", encoding="utf-8") as fh:
                fh.write(new)
            print(f"{readme} updated")
        return 0
    if not opts.get("check"):
        sys.stdout.write(render(res.module, load_lockfile(directory, directory)))
        return 0
    readme = find_readme(directory)
    if readme is None:
        print(f"{directory}: no README found")
        return 1
    found = check_readme(res.module, readme, os.path.relpath(readme, directory))
    _print_diags(found, sys.stdout)
    print(f"{len(found)} finding(s)")
    return 1 if found else 0


def cmd_check(argv: list[str]) -> int:
    opts, roots = _options(argv, flags=("json", "strict"), valued=("terraform-version",))
    root = os.path.abspath(roots[0] if roots else ".")
    rep = check_tree(root, terraform_version=opts.get("terraform-version"))
    if opts.get("json"):
        out = rep.to_dict()
        for m in out["modules"]:
            for d in m["diagnostics"]:
                d["abs_file"] = os.path.normpath(os.path.join(root, d["file"])) if d["file"] else ""
        print(json.dumps(out, indent=2))
    else:
        for p in rep.fmt.changed:
            print(f"fmt: {p} is not in canonical format")
        for m in rep.modules:
            _print_diags([d for d in m.all() if d.severity in (ERROR, WARNING)], sys.stdout)
        c = rep.to_dict()["counts"]
        print(f"{len(rep.modules)} modules, {len(rep.fmt.checked)} files, {rep.bytes_scanned} bytes in "
              f"{rep.seconds * 1000:.0f} ms: {c['fmt_differing']} unformatted, {c['error']} errors, "
              f"{c['warning']} warnings, {c['notice']} notices")
        for m in rep.modules:
            print(f"  {m.path}: validated as {m.terraform_version_source}")
        print(f"(note: {STAND_IN})")
    if opts.get("strict") and rep.count(WARNING):
        return 1
    return 0 if rep.ok else 1


def cmd_report(argv: list[str]) -> int:
    opts, roots = _options(argv, valued=("o", "terraform-version"))
    rep = check_tree(os.path.abspath(roots[0] if roots else "."), terraform_version=opts.get("terraform-version"))
    text = render_markdown(rep)
    out: Optional[str] = opts.get("o")
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
    try:
        return cmd(argv[1:])
    except UsageError as e:
        sys.stderr.write(f"{argv[0]}: {e}\n")
        return 1


if __name__ == "__main__":
    sys.exit(main())
