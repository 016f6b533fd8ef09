#!/usr/bin/env python3
"""NOT terraform-docs.  Same purpose as fake_terraform.py: lets the plumbing of the
dormant harness's terraform-docs leg be exercised where the real tool is absent.
Answers ``markdown table DIR`` with this repository's own ``docs`` output, preceded by
a header line the way terraform-docs may print one."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from nvidia_terraform_modules_b200.config.lockfile import load_lockfile  # noqa: E402
from nvidia_terraform_modules_b200.config.model import load_module  # noqa: E402
from nvidia_terraform_modules_b200.docs.tfdocs import render  # noqa: E402

if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("-")]
    if args[:2] != ["markdown", "table"] or len(args) != 3:
        sys.exit("fake terraform-docs: only `markdown table DIR`")
    d = args[2]
    sys.stdout.write("<!-- header -->\n" + render(load_module(d, d), load_lockfile(d, d)))
