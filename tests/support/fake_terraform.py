#!/usr/bin/env python3
"""NOT Terraform.  A shim with Terraform's command-line shape that answers with this
repository's own stand-ins.

Its one purpose is to let ``tests/test_harness_plumbing.py`` run the dormant parity
harness (tests/test_against_terraform.py) end to end where no real binary exists, so
that the harness's own plumbing -- process calls, JSON shapes, file comparison, corpus
building -- is known to work before the day a real ``terraform`` is dropped in.  A
green run through this shim compares the code with itself and says NOTHING about
parity.

Speaks: ``version -json``, ``fmt [flags] [targets|-]``, ``init ...`` (no-op),
``validate -json``.  The release it pretends to be is ``$FAKE_TERRAFORM_VERSION``
(default 1.10.0).
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from nvidia_terraform_modules_b200.config.validate import validate_module  # noqa: E402
from nvidia_terraform_modules_b200.fmt import command as fmt_command  # noqa: E402
from nvidia_terraform_modules_b200.hcl.diagnostics import ERROR, WARNING  # noqa: E402

VERSION = os.environ.get("FAKE_TERRAFORM_VERSION", "1.10.0")


def main(argv: list[str]) -> int:
    if not argv:
        return 127
    cmd, args = argv[0], argv[1:]
    if cmd == "version":
        print(json.dumps({"terraform_version": VERSION, "platform": "fake", "provider_selections": {},
                          "terraform_outdated": False}))
        return 0
    if cmd == "fmt":
        return fmt_command.main(args)
    if cmd == "init":
        print("Terraform has been successfully initialized! (fake)")
        return 0
    if cmd == "validate":
        res = validate_module(".", ".", ".".join(VERSION.split(".")[:2]))
        print(json.dumps({
            "format_version": "1.0", "valid": res.valid,
            "error_count": sum(d.severity == ERROR for d in res.diags),
            "warning_count": sum(d.severity == WARNING for d in res.diags),
            "diagnostics": [{"severity": d.severity, "summary": d.summary, "detail": d.detail,
                             "range": {"filename": d.file, "start": {"line": d.line, "column": d.col, "byte": 0},
                                       "end": {"line": d.line, "column": d.col, "byte": 0}}}
                            for d in res.diags]}))
        return 0 if res.valid else 1
    sys.stderr.write(f"fake terraform: unknown command {cmd}\n")
    return 127


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
