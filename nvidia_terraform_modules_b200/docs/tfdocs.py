"""README reference tables: generate them and check the committed ones.

The reference asks contributors to regenerate each README's tables with
``terraform-docs markdown .`` (CONTRIBUTING.md:14); that tool is absent here and
the committed tables have drifted (SURVEY.md section 8 "Next 3":
eks/examples/cnpack/Readme.md:229 vs variables.tf:58, and
gke/examples/cnpack/README.md:101 documents an input that does not exist).

``render`` writes the six sections terraform-docs' markdown-table format emits
(Requirements, Providers, Modules, Resources, Inputs, Outputs) from the module
model; ``check_readme`` parses the same sections out of an existing README and
reports every row that disagrees with the ``.tf`` files.  Table layout restated
from terraform-docs' documented "markdown table" output.
"""
from __future__ import annotations

import json
import os
import re
from dataclasses import dataclass, field
from typing import Optional

from ..config import typeexpr
from ..config.lockfile import LockFile
from ..config.model import Module, Variable
from ..hcl import ast as A
from ..hcl.diagnostics import Diagnostic, WARNING
from ..hcl.tokens import T

SECTIONS = ("Requirements", "Providers", "Modules", "Resources", "Inputs", "Outputs")
README_NAMES = ("README.md", "Readme.md", "readme.md", "README.MD")


# ---------------------------------------------------------------------------
# generation
# ---------------------------------------------------------------------------
def _esc(name: str) -> str:
    return name.replace("_", "\\_")


def _anchor(kind: str, name: str) -> str:
    return f"<a name=\"{kind}_{name}\"></a> [{_esc(name)}](#{kind}\\_{_esc(name)})"


def _source_text(m: Module, file: str, node: A.Attribute) -> str:
    for f in m.files:
        if f.path == file:
            toks = f.tokens[node.expr_tok_start:node.expr_tok_end]
            return "".join(t.raw_before + t.text for t in toks).strip()
    return ""


def _json(value) -> str:
    return json.dumps(value, indent=2, ensure_ascii=False)


def _code_cell(text: str) -> str:
    if "\n" in text:
        return "<pre>" + text.replace("\n", "<br>") + "</pre>"
    return f"`{text}`"


def type_cell(m: Module, var: Variable) -> str:
    attrs = var.block.body.attributes if var.block else {}
    if "type" in attrs:
        text = _source_text(m, var.file, attrs["type"])
        # terraform-docs keeps a multi-line type expression as written
        return _code_cell(text)
    if var.has_default:
        d = var.default
        if isinstance(d, bool):
            return "`bool`"
        if isinstance(d, (int, float)):
            return "`number`"
        if isinstance(d, str):
            return "`string`"
        if isinstance(d, list):
            return "`list`"
        if isinstance(d, dict):
            return "`map`"
    return "`any`"


def default_cell(var: Variable) -> str:
    if not var.has_default:
        return "n/a"
    if var.default is typeexpr.UNKNOWN or _holds_unknown(var.default):
        return "n/a"           # not a literal (itself an error: "Variables not allowed")
    return _code_cell(_json(var.default))


def _holds_unknown(value) -> bool:
    if value is typeexpr.UNKNOWN:
        return True
    if isinstance(value, dict):
        return any(_holds_unknown(v) for v in value.values())
    if isinstance(value, (list, tuple)):
        return any(_holds_unknown(v) for v in value)
    return False


def leading_comment(m: Module, decl) -> Optional[str]:
    """terraform-docs falls back to the ``#``/``//`` comment lines directly above
    a variable or output block when it has no ``description``."""
    if decl.block is None:
        return None
    for f in m.files:
        if f.path != decl.file:
            continue
        lines: list[str] = []
        i = decl.block.tok_start - 1
        while i >= 0 and f.tokens[i].type is T.COMMENT and f.tokens[i].text.endswith("\n"):
            text = f.tokens[i].text.strip()
            text = text[2:] if text.startswith("//") else text[1:]
            lines.append(text.strip())
            i -= 1
        if lines:
            return " ".join(reversed(lines))
    return None


def described(m: Module, decl) -> Optional[str]:
    return decl.description or leading_comment(m, decl)


def doc_resource_name(r) -> str:
    """terraform-docs prints <provider>_<type without provider prefix>.<name>."""
    prov = r.implied_provider
    short = r.type[len(prov) + 1:] if r.type.startswith(prov + "_") else r.type
    return f"{prov}_{short}.{r.name}"


def _desc(text: Optional[str]) -> str:
    if not text:
        return "n/a"
    text = text.strip().replace("\n", "<br>").replace("|", "\\|")
    # underscores are escaped outside `code spans`
    parts = text.split("`")
    for i in range(0, len(parts), 2):
        parts[i] = parts[i].replace("_", "\\_")
    return "`".join(parts)


def render(m: Module, lock: Optional[LockFile] = None) -> str:
    out: list[str] = []

    out.append("## Requirements\n")
    rows = [(_anchor("requirement", "terraform"), text) for text, _ in m.required_versions[:1]]
    rows += [(_anchor("requirement", n), r.constraint_text) for n, r in sorted(m.required_providers.items())
             if r.constraint_text]
    if rows:
        out.append("| Name | Version |\n|------|---------|")
        out.extend(f"| {a} | {b} |" for a, b in rows)
    else:
        out.append("No requirements.")

    out.append("\n## Providers\n")
    used = sorted(m.used_provider_names())
    if used:
        out.append("| Name | Version |\n|------|---------|")
        for name in used:
            req = m.required_providers.get(name)
            ver = "n/a"
            fqn = req.fqn if req else f"registry.terraform.io/hashicorp/{name}"
            if lock and fqn in lock.providers and lock.providers[fqn].version:
                ver = str(lock.providers[fqn].version)
            elif req and req.constraint_text:
                ver = req.constraint_text
            out.append(f"| {_anchor('provider', name)} | {ver} |")
    else:
        out.append("No providers.")

    out.append("\n## Modules\n")
    if m.module_calls:
        out.append("| Name | Source | Version |\n|------|--------|---------|")
        for name, mc in sorted(m.module_calls.items()):
            out.append(f"| {_anchor('module', name)} | {mc.source or 'n/a'} | {mc.version or 'n/a'} |")
    else:
        out.append("No modules.")

    out.append("\n## Resources\n")
    if m.resources:
        out.append("| Name | Type |\n|------|------|")
        ordered = sorted(m.resources.values(), key=lambda r: (r.mode == "data", doc_resource_name(r)))
        for r in ordered:
            req = m.required_providers.get(r.implied_provider)
            ns_type = req.source if req and req.source and req.source.count("/") == 1 \
                else f"hashicorp/{r.implied_provider}"
            ver = "latest"
            if req and req.constraints is not None and req.constraints.is_exact():
                ver = req.constraint_text.strip().lstrip("= ")
            kind = "data-sources" if r.mode == "data" else "resources"
            prov = r.implied_provider
            short = r.type[len(prov) + 1:] if r.type.startswith(prov + "_") else r.type
            url = f"https://registry.terraform.io/providers/{ns_type}/{ver}/docs/{kind}/{short}"
            out.append(f"| [{doc_resource_name(r)}]({url}) | {'data source' if r.mode == 'data' else 'resource'} |")
    else:
        out.append("No resources.")

    out.append("\n## Inputs\n")
    if m.variables:
        out.append("| Name | Description | Type | Default | Required |\n"
                   "|------|-------------|------|---------|:--------:|")
        for name, var in sorted(m.variables.items()):
            out.append(f"| {_anchor('input', name)} | {_desc(described(m, var))} | {type_cell(m, var)} | "
                       f"{default_cell(var)} | {'yes' if var.required else 'no'} |")
    else:
        out.append("No inputs.")

    out.append("\n## Outputs\n")
    if m.outputs:
        out.append("| Name | Description |\n|------|-------------|")
        for name, o in sorted(m.outputs.items()):
            out.append(f"| {_anchor('output', name)} | {_desc(described(m, o))} |")
    else:
        out.append("No outputs.")
    return "\n".join(out) + "\n"


# ---------------------------------------------------------------------------
# parsing an existing README
# ---------------------------------------------------------------------------
@dataclass
class DocTable:
    heading_line: int
    header: list[str]
    rows: list[tuple[int, list[str]]] = field(default_factory=list)   # (line number, cells)


def _split_row(line: str) -> list[str]:
    cells, cur, i = [], [], 0
    body = line.strip()
    if body.startswith("|"):
        body = body[1:]
    if body.endswith("|") and not body.endswith("\\|"):
        body = body[:-1]
    while i < len(body):
        if body[i] == "\\" and i + 1 < len(body) and body[i + 1] == "|":
            cur.append("|")
            i += 2
        elif body[i] == "|":
            cells.append("".join(cur).strip())
            cur = []
            i += 1
        else:
            cur.append(body[i])
            i += 1
    cells.append("".join(cur).strip())
    return cells


def parse_readme(text: str) -> dict[str, DocTable]:
    tables: dict[str, DocTable] = {}
    lines = text.splitlines()
    i = 0
    while i < len(lines):
        m = re.match(r"^##\s+(\w+)\s*$", lines[i])
        if m and m.group(1) in SECTIONS and m.group(1) not in tables:
            name, heading = m.group(1), i + 1
            j = i + 1
            while j < len(lines) and not lines[j].lstrip().startswith("|") and not lines[j].startswith("#"):
                j += 1
            if j < len(lines) and lines[j].lstrip().startswith("|"):
                table = DocTable(heading, _split_row(lines[j]))
                j += 1
                if j < len(lines) and re.match(r"^\s*\|[\s:|-]+\|\s*$", lines[j]):
                    j += 1
                while j < len(lines) and lines[j].lstrip().startswith("|"):
                    table.rows.append((j + 1, _split_row(lines[j])))
                    j += 1
                tables[name] = table
            else:
                tables[name] = DocTable(heading, [])
            i = j
        else:
            i += 1
    return tables


_LINK = re.compile(r"\[([^\]]+)\]\([^)]*\)")


def _cell_name(cell: str) -> str:
    cell = re.sub(r"<a [^>]*></a>", "", cell).strip()
    m = _LINK.search(cell)
    if m:
        cell = m.group(1)
    return cell.replace("\\_", "_").strip()


def _norm_code(cell: str) -> str:
    cell = cell.strip()
    if cell.startswith("<pre>") and cell.endswith("</pre>"):
        cell = cell[5:-6].replace("<br>", "\n").replace("<br/>", "\n")
    cell = cell.strip("`")
    return re.sub(r"\s+", "", cell)


def _norm_text(cell: str) -> str:
    return re.sub(r"\s+", " ", cell.replace("<br>", " ").replace("\\_", "_").replace("\\|", "|")).strip()


def _default_equal(doc_cell: str, var: Variable) -> bool:
    want = default_cell(var)
    if _norm_code(doc_cell) == _norm_code(want):
        return True
    if not var.has_default:
        return doc_cell.strip() in ("n/a", "")
    raw = doc_cell.strip()
    if raw.startswith("<pre>"):
        raw = raw[5:-6].replace("<br>", "\n")
    raw = raw.strip("`")
    try:
        return json.loads(raw) == var.default
    except (ValueError, TypeError):
        return False


def check_readme(m: Module, readme_path: str, display: Optional[str] = None) -> list[Diagnostic]:
    disp = display or readme_path
    with open(readme_path, "r", encoding="utf-8") as fh:
        tables = parse_readme(fh.read())
    diags: list[Diagnostic] = []

    def add(line: int, summary: str, detail: str = "", rule: str = "docs-drift") -> None:
        diags.append(Diagnostic(WARNING, summary, detail, disp, line, 0, rule))

    if not any(t.header for t in tables.values()):
        return diags          # README has no generated tables at all: nothing to compare

    # ---- Inputs
    t = tables.get("Inputs")
    if t is not None and t.header:
        cols = {h.lower(): i for i, h in enumerate(t.header)}
        seen = set()
        for line, cells in t.rows:
            name = _cell_name(cells[0])
            seen.add(name)
            var = m.variables.get(name)
            if var is None:
                add(line, f"README documents input \"{name}\" which is not declared",
                    "No variable block with this name exists in the module.", "docs-unknown-input")
                continue

            def cell(col: str) -> Optional[str]:
                i = cols.get(col)
                return cells[i] if i is not None and i < len(cells) else None

            c = cell("description")
            if c is not None and _norm_text(c) != _norm_text(_desc(described(m, var))):
                add(line, f"input \"{name}\": description differs from {var.file}:{var.line}",
                    f"README: {_norm_text(c)!r}; code: {_norm_text(_desc(described(m, var)))!r}")
            c = cell("type")
            if c is not None and _norm_code(c) != _norm_code(type_cell(m, var)):
                add(line, f"input \"{name}\": type differs from {var.file}:{var.line}",
                    f"README: {c}; code: {type_cell(m, var)}")
            c = cell("default")
            if c is not None and not _default_equal(c, var):
                add(line, f"input \"{name}\": default differs from {var.file}:{var.line}",
                    f"README: {c}; code: {default_cell(var)}")
            c = cell("required")
            if c is not None and c.strip().lower() in ("yes", "no") and (c.strip().lower() == "yes") != var.required:
                add(line, f"input \"{name}\": README says required={c.strip()}, code says "
                          f"{'yes' if var.required else 'no'}", f"See {var.file}:{var.line}.")
        for name, var in sorted(m.variables.items()):
            if name not in seen:
                add(t.heading_line, f"input \"{name}\" is not documented in the README",
                    f"Declared at {var.file}:{var.line}.", "docs-missing-input")

    # ---- Outputs
    t = tables.get("Outputs")
    if t is not None and t.header:
        seen = set()
        for line, cells in t.rows:
            name = _cell_name(cells[0])
            seen.add(name)
            o = m.outputs.get(name)
            if o is None:
                add(line, f"README documents output \"{name}\" which is not declared", "", "docs-unknown-output")
            elif len(cells) > 1 and _norm_text(cells[1]) != _norm_text(_desc(described(m, o))):
                add(line, f"output \"{name}\": description differs from {o.file}:{o.line}",
                    f"README: {_norm_text(cells[1])!r}; code: {_norm_text(_desc(described(m, o)))!r}")
        for name, o in sorted(m.outputs.items()):
            if name not in seen:
                add(t.heading_line, f"output \"{name}\" is not documented in the README",
                    f"Declared at {o.file}:{o.line}.", "docs-missing-output")
    elif t is not None and m.outputs:
        add(t.heading_line, "README says the module has no outputs",
            f"{len(m.outputs)} output(s) are declared: {', '.join(sorted(m.outputs))}.", "docs-missing-output")

    # ---- Modules
    t = tables.get("Modules")
    if t is not None:
        seen = set()
        for line, cells in t.rows:
            name = _cell_name(cells[0])
            seen.add(name)
            mc = m.module_calls.get(name)
            if mc is None:
                add(line, f"README documents module \"{name}\" which is not called", "", "docs-unknown-module")
                continue
            if len(cells) > 1 and cells[1].strip() != (mc.source or "n/a"):
                add(line, f"module \"{name}\": source differs from {mc.file}:{mc.line}",
                    f"README: {cells[1]}; code: {mc.source}")
            if len(cells) > 2 and cells[2].strip() != (mc.version or "n/a"):
                add(line, f"module \"{name}\": version differs from {mc.file}:{mc.line}",
                    f"README: {cells[2]}; code: {mc.version or 'n/a'}")
        for name, mc in sorted(m.module_calls.items()):
            if name not in seen:
                add(t.heading_line, f"module \"{name}\" is not documented in the README",
                    f"Called at {mc.file}:{mc.line}.", "docs-missing-module")

    # ---- Resources
    t = tables.get("Resources")
    if t is not None:
        seen = set()
        for line, cells in t.rows:
            name = _cell_name(cells[0])
            kind = cells[1].strip() if len(cells) > 1 else ""
            by_doc_name = {(r.mode == "data", doc_resource_name(r)): a for a, r in m.resources.items()}
            addr = by_doc_name.get((kind == "data source", name),
                                   ("data." if kind == "data source" else "") + name)
            seen.add(addr)
            if addr not in m.resources:
                add(line, f"README documents {kind or 'resource'} \"{name}\" which is not declared", "",
                    "docs-unknown-resource")
        for addr, r in sorted(m.resources.items()):
            if addr not in seen:
                add(t.heading_line, f"{'data source' if r.mode == 'data' else 'resource'} "
                    f"\"{r.type}.{r.name}\" is not documented in the README",
                    f"Declared at {r.file}:{r.line}.", "docs-missing-resource")

    # ---- Providers (names only: the version column depends on whether a lock file
    # existed when the table was generated, "n/a" otherwise)
    t = tables.get("Providers")
    if t is not None:
        used = m.used_provider_names()
        seen = set()
        for line, cells in t.rows:
            name = _cell_name(cells[0])
            seen.add(name)
            if name not in used:
                add(line, f"README lists provider \"{name}\" which no resource or data source uses", "",
                    "docs-unknown-provider")
        for name in sorted(used):
            if name not in seen:
                first = min(used[name], key=lambda r: (r.file, r.line))
                add(t.heading_line, f"provider \"{name}\" is not documented in the README",
                    f"Used by {first.addr} at {first.file}:{first.line}.", "docs-missing-provider")

    # ---- Requirements
    t = tables.get("Requirements")
    if t is not None:
        seen = set()
        for line, cells in t.rows:
            name = _cell_name(cells[0])
            seen.add(name)
            ver = cells[1].strip() if len(cells) > 1 else ""
            if name == "terraform":
                want = m.required_versions[0][0] if m.required_versions else None
            else:
                req = m.required_providers.get(name)
                if req is None:
                    add(line, f"README lists requirement \"{name}\" which is not in required_providers", "",
                        "docs-unknown-requirement")
                    continue
                want = req.constraint_text
            if (want or "n/a") != ver:
                add(line, f"requirement \"{name}\": README says {ver}, code says {want or 'n/a'}")
        wanted = (["terraform"] if m.required_versions else []) + sorted(
            n for n, r in m.required_providers.items() if r.constraint_text)
        for name in wanted:
            if name not in seen:
                add(t.heading_line, f"requirement \"{name}\" is not documented in the README", "",
                    "docs-missing-requirement")
    return diags


BEGIN_MARK = "<!-- BEGIN_TF_DOCS -->"
END_MARK = "<!-- END_TF_DOCS -->"


class InjectRefused(ValueError):
    """The README has no markers and no single, unambiguous generated region."""


_HEADING = re.compile(r"^##\s+(\w+)\s*$")


def _generated_sections(lines: list[str]) -> list[tuple[int, int, int]]:
    """``(heading, generated_end, section_end)`` line indices of every section that
    terraform-docs evidently wrote: a ``## Requirements``-style heading followed,
    after blank lines only, by a table whose first header cell is ``Name`` or by the
    one-line ``No inputs.`` sentence.  ``generated_end`` is one past that table or
    sentence; ``section_end`` is the next heading of any level (or end of file).  A
    heading of the same name over hand-written prose is not in the list."""
    found = []
    for i, ln in enumerate(lines):
        m = _HEADING.match(ln)
        if not m or m.group(1) not in SECTIONS:
            continue
        j = i + 1
        while j < len(lines) and lines[j].strip() == "":
            j += 1
        if j >= len(lines):
            continue
        if re.match(r"^No [a-z ]+\.\s*$", lines[j]):
            gen_end = j + 1
        elif lines[j].lstrip().startswith("|") and _split_row(lines[j])[:1] == ["Name"]:
            gen_end = j
            while gen_end < len(lines) and lines[gen_end].lstrip().startswith("|"):
                gen_end += 1
        else:
            continue
        sec_end = gen_end
        while sec_end < len(lines) and not lines[sec_end].startswith("#"):
            sec_end += 1
        found.append((i, gen_end, sec_end))
    return found


def inject(readme_text: str, tables: str) -> str:
    """Put freshly rendered tables into a README.

    * Between ``<!-- BEGIN_TF_DOCS -->`` / ``<!-- END_TF_DOCS -->`` when the README has
      them -- terraform-docs' "inject" mode.
    * Without markers terraform-docs only ever appends.  The reference's READMEs were
      made by pasting its output, without markers, so one more case is handled: when
      the README holds exactly ONE unbroken run of generated sections (see
      ``_generated_sections``; nothing but blank lines between one section's table and
      the next section's heading) that run is replaced, and whatever follows its last
      table is kept.  Sections named "Requirements", "Resources", ... that hold
      hand-written prose are never touched.
    * No generated section at all: the tables are appended between markers, as
      terraform-docs does.
    * Generated sections split up by hand-written text: ``InjectRefused`` -- there is
      no single region to replace, and guessing would delete somebody's prose.
    """
    tables = tables.rstrip("\n") + "\n"
    if BEGIN_MARK in readme_text and END_MARK in readme_text:
        head, rest = readme_text.split(BEGIN_MARK, 1)
        _, tail = rest.split(END_MARK, 1)
        return f"{head}{BEGIN_MARK}\n{tables}{END_MARK}{tail}"
    lines = readme_text.split("\n")
    secs = _generated_sections(lines)
    if not secs:
        sep = "" if readme_text.endswith("\n\n") or not readme_text else ("\n" if readme_text.endswith("\n") else "\n\n")
        return f"{readme_text}{sep}{BEGIN_MARK}\n{tables}{END_MARK}\n"
    for (_, gen_end, sec_end), (nxt, _, _) in zip(secs, secs[1:]):
        if sec_end != nxt or any(ln.strip() for ln in lines[gen_end:sec_end]):
            at = next((k for k in range(gen_end, nxt) if lines[k].strip()), gen_end)
            raise InjectRefused(
                f"generated sections are interrupted by other text at line {at + 1}; put "
                f"{BEGIN_MARK} and {END_MARK} around the region to regenerate")
    first, end = secs[0][0], secs[-1][1]
    out = lines[:first] + tables.rstrip("\n").split("\n") + lines[end:]
    text = "\n".join(out)
    if readme_text.endswith("\n") and not text.endswith("\n"):
        text += "\n"
    return text


def find_readme(directory: str) -> Optional[str]:
    for n in README_NAMES:
        p = os.path.join(directory, n)
        if os.path.isfile(p):
            return p
    return None
