"""``fmt`` command: same flags, file selection, output and exit codes as
``terraform fmt``.

This is BASELINE.json config 1 -- "``terraform fmt -check -recursive`` over eks/,
aks/, gke/" -- runnable without the Terraform binary (absent from this image).
It is a STAND-IN, not ``terraform fmt``: see DESIGN.md "Parity".

Behaviour restated from Terraform's CLI documentation:

* targets: files or directories; default ``.``; ``-`` reads stdin and writes stdout;
* a directory target formats ``*.tf``, ``*.tfvars`` and ``*.tftest.hcl`` in it
  (lock files such as eks/.terraform.lock.hcl are not touched);
  ``-recursive`` also descends into sub-directories, skipping hidden ones;
* ``-list`` (default true) prints the path of each file whose canonical form
  differs; ``-write`` (default true) rewrites it; ``-diff`` prints a unified diff;
* ``-check`` disables writing and exits 3 when any file differs;
* a file with syntax errors is never reformatted; diagnostics go to stderr and
  the exit status is 2.
"""
from __future__ import annotations

import difflib
import os
import shutil
import subprocess
import sys
import tempfile
from dataclasses import dataclass, field
from typing import Iterable, Optional, TextIO

from ..hcl.diagnostics import Diagnostic, has_errors
from ..hcl.parser import parse, parse_path
from ..hcl.tokens import render
from .canonical import format_tokens
from .rewrites import apply_rewrites, collect_rewrites

FMT_SUFFIXES = (".tf", ".tfvars", ".tftest.hcl")


def format_source(src: str, filename: str = "", parsed=None) -> tuple[str, list[Diagnostic]]:
    """Canonical form of one file's text.  On syntax errors the text is
    returned unchanged together with the diagnostics.  ``parsed`` may hand in
    an existing ``(body, diagnostics, tokens)`` for ``src``; its tokens are
    left exactly as they were."""
    body, diags, tokens = parsed if parsed is not None else parse(src, filename)
    if has_errors(diags):
        return src, diags
    saved = [t.spaces_before for t in tokens]
    try:
        laid_out = apply_rewrites(tokens, collect_rewrites(body, tokens))
        format_tokens(laid_out)
        out = render(laid_out)
    finally:
        for t, n in zip(tokens, saved):
            t.spaces_before = n
    if src.startswith("\ufeff"):
        out = "\ufeff" + out
    return out, diags


def is_formattable(name: str) -> bool:
    return name.endswith(FMT_SUFFIXES)


def discover(target: str, recursive: bool) -> list[str]:
    """Files ``fmt`` would look at for one directory target, in stable order."""
    found: list[str] = []
    try:
        entries = sorted(os.listdir(target))
    except OSError:
        return found
    for name in entries:
        path = os.path.join(target, name)
        if os.path.isdir(path):
            if recursive and not name.startswith("."):
                found.extend(discover(path, recursive))
        elif is_formattable(name) and not name.startswith("."):
            found.append(path)
    return found


@dataclass
class FmtResult:
    changed: list[str] = field(default_factory=list)      # files whose canonical form differs
    checked: list[str] = field(default_factory=list)
    diags: list[Diagnostic] = field(default_factory=list)
    diffs: dict[str, str] = field(default_factory=dict)
    bytes_checked: int = 0

    @property
    def exit_code(self) -> int:
        if has_errors(self.diags):
            return 2
        return 0


_DIFF_PROGRAM = shutil.which("diff")


def _external_diff(path: str, old: str, new: str) -> Optional[str]:
    with tempfile.TemporaryDirectory(prefix="tffmt_") as d:
        files = []
        for name, text in (("old", old), ("new", new)):
            files.append(os.path.join(d, name))
            with open(files[-1], "w", encoding="utf-8", newline="") as fh:
                fh.write(text)
        try:
            p = subprocess.run([_DIFF_PROGRAM, f"--label=old/{path}", f"--label=new/{path}", "-u", *files],
                               capture_output=True, timeout=60)
        except (OSError, subprocess.SubprocessError):
            return None
    if p.returncode not in (0, 1):          # 1 = "files differ"
        return None
    return p.stdout.decode("utf-8", errors="replace")


def _hunk_range(start: int, stop: int) -> str:
    """``start,count`` the way ``diff -u`` writes it: the count is left out when it is
    1, and an empty range names the line before it."""
    count = stop - start
    if count == 1:
        return str(start + 1)
    if count == 0:
        return f"{start},0"
    return f"{start + 1},{count}"


def unified_diff(path: str, old: str, new: str, external: bool = True) -> str:
    """What ``terraform fmt -diff`` prints.  Terraform shells out to
    ``diff --label=old/PATH --label=new/PATH -u``; so does this when a ``diff`` program
    is installed (byte-identical output by construction).  Without one -- Terraform
    itself fails there -- the same GNU unified format is produced in-process; the two
    agree whenever the changed lines are not repeated elsewhere in the file, which is
    the case ``diff`` and ``difflib`` may pair up differently."""
    if external and _DIFF_PROGRAM:
        text = _external_diff(path, old, new)
        if text is not None:
            return text
    a = old.splitlines(keepends=True)
    b = new.splitlines(keepends=True)
    out = [f"--- old/{path}\n", f"+++ new/{path}\n"]

    def emit(tag: str, line: str) -> None:
        out.append(tag + line if line.endswith("\n") else tag + line + "\n\\ No newline at end of file\n")

    # autojunk off: blank lines are frequent in .tf files and must still match
    matcher = difflib.SequenceMatcher(None, a, b, autojunk=False)
    for group in matcher.get_grouped_opcodes(3):
        a0, a1, b0, b1 = group[0][1], group[-1][2], group[0][3], group[-1][4]
        out.append(f"@@ -{_hunk_range(a0, a1)} +{_hunk_range(b0, b1)} @@\n")
        for tag, i1, i2, j1, j2 in group:
            if tag == "equal":
                for ln in a[i1:i2]:
                    emit(" ", ln)
                continue
            if tag in ("replace", "delete"):
                for ln in a[i1:i2]:
                    emit("-", ln)
            if tag in ("replace", "insert"):
                for ln in b[j1:j2]:
                    emit("+", ln)
    return "".join(out)


def _display(path: str) -> str:
    return path[2:] if path.startswith("./") else path


def run_fmt(targets: Iterable[str], *, recursive: bool = False, write: bool = True,
            want_diff: bool = False, check: bool = False, external_diff: bool = True) -> FmtResult:
    res = FmtResult()
    if check:
        write = False
    files: list[str] = []
    for target in targets:
        if os.path.isdir(target):
            files.extend(discover(target, recursive))
        elif os.path.isfile(target):
            if not is_formattable(os.path.basename(target)):
                res.diags.append(Diagnostic("error", "Only .tf, .tfvars, and .tftest.hcl files can be "
                                            f"processed with terraform fmt", file=target, rule="fmt-target"))
                continue
            files.append(target)
        else:
            res.diags.append(Diagnostic("error", f"No file or directory at {target}", file=target,
                                        rule="fmt-target"))
    seen: set[str] = set()
    for path in files:
        real = os.path.normpath(path)
        if real in seen:                 # overlapping targets ("." and "sub") name a file once
            continue
        seen.add(real)
        src, body, pdiags, tokens = parse_path(path, _display(path))
        res.checked.append(_display(path))
        res.bytes_checked += len(src.encode("utf-8"))
        out, diags = format_source(src, _display(path), (body, pdiags, tokens))
        res.diags.extend(diags)
        if has_errors(diags) or out == src:
            continue
        res.changed.append(_display(path))
        if want_diff:
            res.diffs[_display(path)] = unified_diff(_display(path), src, out, external_diff)
        if write:
            with open(path, "w", encoding="utf-8", newline="") as fh:
                fh.write(out)
    return res


def main(argv: list[str], stdout: TextIO = sys.stdout, stderr: TextIO = sys.stderr,
         stdin: TextIO = sys.stdin) -> int:
    opts = {"list": True, "write": True, "diff": False, "check": False, "recursive": False}
    targets: list[str] = []
    for arg in argv:
        if arg.startswith("-") and arg != "-":
            name, _, val = arg.lstrip("-").partition("=")
            if name == "no-color":
                continue
            if name not in opts:
                stderr.write(f"Error parsing command-line flags: flag provided but not defined: -{name}\n")
                return 1
            opts[name] = True if val == "" else val.lower() not in ("false", "0", "f")
        else:
            targets.append(arg)
    if not targets:
        targets = ["."]

    if targets == ["-"]:
        src = stdin.read()
        out, diags = format_source(src, "<stdin>")
        for d in diags:
            stderr.write(d.format() + "\n\n")
        if has_errors(diags):
            return 2
        if opts["check"]:
            return 3 if out != src else 0
        if opts["diff"]:
            stdout.write(unified_diff("<stdin>", src, out))
        else:
            stdout.write(out)
        return 0

    res = run_fmt(targets, recursive=opts["recursive"], write=opts["write"],
                  want_diff=opts["diff"], check=opts["check"])
    for d in res.diags:
        stderr.write(d.format() + "\n\n")
    if has_errors(res.diags) and opts["check"]:
        # Terraform buffers -check output and gives up before printing it when any
        # file failed to parse; without -check the names stream out as files are done
        return 2
    for path in res.changed:
        if opts["list"]:
            stdout.write(path + "\n")
        if opts["diff"]:
            stdout.write(res.diffs[path])
    if has_errors(res.diags):
        return 2
    if opts["check"] and res.changed:
        return 3
    return 0
