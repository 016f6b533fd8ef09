"""Diagnostics shared by the parser, the validator and the lint rules.

Shape follows Terraform's own diagnostics (severity / summary / detail / source
range) so messages can be worded the way ``terraform validate`` words them.
"""
from __future__ import annotations

from dataclasses import dataclass, asdict

ERROR = "error"
WARNING = "warning"
NOTICE = "notice"


@dataclass
class Diagnostic:
    severity: str
    summary: str
    detail: str = ""
    file: str = ""
    line: int = 0
    col: int = 0
    rule: str = ""           # stable machine-readable id, e.g. "undeclared-variable"

    def to_dict(self) -> dict:
        return asdict(self)

    def format(self) -> str:
        where = f"{self.file}:{self.line}" if self.file else f"line {self.line}"
        tag = f" [{self.rule}]" if self.rule else ""
        head = f"{self.severity.capitalize()}: {self.summary}{tag}\n  on {where}"
        return head + (f"\n  {self.detail}" if self.detail else "")


class ParseError(Exception):
    def __init__(self, diag: Diagnostic):
        super().__init__(diag.summary)
        self.diag = diag


def has_errors(diags: list[Diagnostic]) -> bool:
    return any(d.severity == ERROR for d in diags)
