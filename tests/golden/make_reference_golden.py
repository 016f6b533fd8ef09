#!/usr/bin/env python
"""Regenerate tests/golden/reference_findings.json from the reference tree.

Run in the development container, where the reference is mounted read-only:

    python tests/golden/make_reference_golden.py [/root/reference]

The JSON holds RESULTS only (paths, line numbers, rule ids, digests of the
formatter's output) -- never reference source text -- so it can travel to
machines where the reference does not exist.
"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from nvidia_terraform_modules_b200.check import check_tree  # noqa: E402
from nvidia_terraform_modules_b200.config.lockfile import load_lockfile  # noqa: E402
from nvidia_terraform_modules_b200.config.model import load_module  # noqa: E402
from nvidia_terraform_modules_b200.docs.tfdocs import find_readme, render  # noqa: E402
from nvidia_terraform_modules_b200.fmt.command import discover, format_source  # noqa: E402


def build(root: str) -> dict:
    rep = check_tree(root)
    out = {
        "fmt_checked": rep.fmt.checked,
        "fmt_changed": rep.fmt.changed,
        "fmt_output_sha256": {},
        "modules": {},
        "docs_rows": {},
    }
    cwd = os.getcwd()
    os.chdir(root)
    try:
        for top in ("aks", "eks", "gke"):
            for path in discover(top, True):
                with open(path, newline="") as fh:
                    text, _ = format_source(fh.read(), path)
                out["fmt_output_sha256"][path] = hashlib.sha256(text.encode()).hexdigest()
        for m in rep.modules:
            out["modules"][m.path] = {
                "counts": m.counts,
                "findings": sorted([d.rule, d.file, d.line] for d in m.all()),
            }
            mod = load_module(m.path, ".")
            readme = find_readme(m.path)
            rows = [ln for ln in render(mod, load_lockfile(m.path, ".")).splitlines()
                    if ln.startswith("| ") and not ln.startswith("| Name")]
            have = set(open(readme).read().splitlines())
            out["docs_rows"][m.path] = {"generated": len(rows),
                                        "identical_in_readme": sum(1 for r in rows if r in have)}
    finally:
        os.chdir(cwd)
    return out


if __name__ == "__main__":
    root = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    data = build(os.path.abspath(root))
    with open(os.path.join(HERE, "reference_findings.json"), "w") as fh:
        json.dump(data, fh, indent=1, sort_keys=True)
        fh.write("\n")
    print("modules:", len(data["modules"]), "fmt changed:", data["fmt_changed"])
