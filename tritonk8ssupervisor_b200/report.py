"""Findings of the static checkers (errors fail a config, warnings do not)."""
from __future__ import annotations

import dataclasses
from typing import List, Optional


@dataclasses.dataclass
class Finding:
    severity: str  # "error" | "warning"
    code: str
    where: str
    message: str

    def __str__(self) -> str:
        return "%s[%s] %s: %s" % (self.severity, self.code, self.where, self.message)


@dataclasses.dataclass
class Report:
    findings: List[Finding] = dataclasses.field(default_factory=list)

    def error(self, code: str, where: str, message: str) -> None:
        self.findings.append(Finding("error", code, where, message))

    def warn(self, code: str, where: str, message: str) -> None:
        self.findings.append(Finding("warning", code, where, message))

    @property
    def errors(self) -> List[Finding]:
        return [f for f in self.findings if f.severity == "error"]

    @property
    def warnings(self) -> List[Finding]:
        return [f for f in self.findings if f.severity == "warning"]

    @property
    def ok(self) -> bool:
        return not self.errors

    def codes(self, severity: Optional[str] = None) -> List[str]:
        return [f.code for f in self.findings if severity is None or f.severity == severity]
