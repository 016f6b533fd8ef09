"""Markdown write-up of a TreeReport: the "fix-up patch list" SURVEY.md section 8
("Next 2") asks for -- every finding with file:line, grouped so a maintainer of
the reference can turn it into issues or one clean-up PR.
"""
from __future__ import annotations

from collections import Counter

from .check import STAND_IN, TreeReport
from .hcl.diagnostics import Diagnostic, ERROR, NOTICE, WARNING
from .lint.rules import RULES

_ORDER = {ERROR: 0, WARNING: 1, NOTICE: 2}


def _row(d: Diagnostic) -> str:
    detail = d.detail.replace("\n", " ").replace("|", "\\|")
    if len(detail) > 220:
        detail = detail[:217] + "..."
    return f"| `{d.file}:{d.line}` | {d.severity} | `{d.rule}` | {d.summary.replace('|', '\\|')} | {detail} |"


def render_markdown(rep: TreeReport, title: str = "Static-check findings") -> str:
    out: list[str] = [f"# {title}", ""]
    out.append(f"Checker status: **{STAND_IN}**.")
    out.append("")
    c = rep.to_dict()["counts"]
    out.append(f"Scanned {len(rep.fmt.checked)} formattable files ({rep.bytes_scanned} bytes) in "
               f"{len(rep.modules)} modules in {rep.seconds * 1000:.0f} ms on one CPU core: "
               f"{c['fmt_differing']} files not in canonical format, {c['error']} errors, "
               f"{c['warning']} warnings, {c['notice']} notices.")
    out.append("")

    out.append("## BASELINE.json configs")
    out.append("")
    out.append("| # | Config | Status | Result |")
    out.append("|---|---|---|---|")
    for i, cfg in enumerate(rep.baseline_configs(), 1):
        if "files_differing" in cfg:
            result = (f"exit {cfg['exit_code']}; {len(cfg['files_differing'])} of {cfg['files_checked']} files differ: "
                      + ", ".join(f"`{p}`" for p in cfg["files_differing"]))
        elif cfg.get("errors") is not None:
            result = f"{cfg['errors']} errors"
        else:
            result = cfg["note"]
        out.append(f"| {i} | `{cfg['config']}` | {cfg['status']} | {result} |")
    out.append("")

    out.append("## Formatting (`fmt -check -recursive`)")
    out.append("")
    if not rep.fmt.changed:
        out.append("All files are in canonical format.")
    for path in rep.fmt.changed:
        out.append(f"### `{path}`")
        out.append("")
        out.append("```diff")
        out.append(rep.fmt.diffs[path].rstrip("\n"))
        out.append("```")
        out.append("")

    out.append("## Findings by module")
    out.append("")
    for m in rep.modules:
        diags = sorted(m.all(), key=lambda d: (_ORDER.get(d.severity, 3), d.file, d.line, d.rule))
        by_sev = Counter(d.severity for d in diags)
        out.append(f"### `{m.path}/` -- {by_sev.get(ERROR, 0)} errors, {by_sev.get(WARNING, 0)} warnings, "
                   f"{by_sev.get(NOTICE, 0)} notices")
        out.append("")
        out.append("Declares " + ", ".join(f"{v} {k.replace('_', ' ')}" for k, v in m.counts.items()) + ".")
        if m.terraform_version_source:
            out.append(f"Validated as {m.terraform_version_source}.")
        out.append("")
        major = [d for d in diags if d.severity in (ERROR, WARNING)]
        if major:
            out.append("| Where | Severity | Rule | Finding | Detail |")
            out.append("|---|---|---|---|---|")
            out.extend(_row(d) for d in major)
            out.append("")
        minor = [d for d in diags if d.severity == NOTICE]
        if minor:
            counts = Counter(d.rule for d in minor)
            out.append("Notices: " + ", ".join(f"`{r}` x{n}" for r, n in sorted(counts.items())) + ".")
            out.append("")

    out.append("## Rules")
    out.append("")
    out.append("| Rule | Default severity | What it catches | Closest tflint rule |")
    out.append("|---|---|---|---|")
    for r in RULES:
        out.append(f"| `{r.id}` | {r.severity} | {r.summary} | {r.tflint or '-'} |")
    out.append("")
    return "\n".join(out)
