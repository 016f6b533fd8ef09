"""Lint rules: findings that ``terraform validate`` accepts but a reviewer would
not.  SURVEY.md section 4 lists the defects of this kind that exist in the reference as
shipped (unused variables, an always-true ``count`` guard, a copy-pasted role
attachment, an output guarded by the wrong flag ...); each rule below is the
general form of one of them, or of a check from the community linter ``tflint``
(its rule name is given where one corresponds).  The reference has no CI
(CONTRIBUTING.md:56), so these are meant to run as a pre-merge hook.

Every rule is a function ``rule(ctx) -> iterable[Diagnostic]`` registered with
``@rule(id, severity, summary)``.  Rules never need network access, provider
schemas or the Terraform binary.
"""
from __future__ import annotations

import os
import re
from dataclasses import dataclass
from typing import Callable, Iterable, Iterator, Optional

from ..config.model import Module, Resource
from ..config.validate import ValidateResult, Validator
from ..hcl import ast as A
from ..hcl.diagnostics import Diagnostic, NOTICE, WARNING
from ..hcl.tokens import T

_SNAKE = re.compile(r"^[a-z][a-z0-9_]*$")
_SECRETISH = re.compile(r"(secret|password|passwd|token|private_key|shared_key|access_key|client_certificate)",
                        re.I)


@dataclass
class LintContext:
    result: ValidateResult
    children: dict[str, Optional[Module]]

    @property
    def m(self) -> Module:
        return self.result.module

    @property
    def v(self) -> Validator:
        return self.result.validator


@dataclass
class Rule:
    id: str
    severity: str
    summary: str
    fn: Callable[[LintContext], Iterable[Diagnostic]]
    tflint: str = ""


RULES: list[Rule] = []


def rule(rule_id: str, severity: str, summary: str, tflint: str = ""):
    def wrap(fn):
        RULES.append(Rule(rule_id, severity, summary, fn, tflint))
        return fn
    return wrap


def _d(r_id: str, sev: str, summary: str, detail: str, file: str, node) -> Diagnostic:
    if isinstance(node, A.Node):
        return Diagnostic(sev, summary, detail, file, node.pos.line, node.pos.col, r_id)
    return Diagnostic(sev, summary, detail, file, int(node), 0, r_id)


def _all_exprs(m: Module) -> list[tuple[str, A.Node]]:
    """Every syntax node of the module with its file, flattened once and kept
    on the (read-only) module so that each rule does not walk the trees again."""
    nodes = getattr(m, "_flat_nodes", None)
    if nodes is None:
        nodes = m._flat_nodes = [(f.path, node) for f in m.files for node in f.body.walk()]
    return nodes


def _expr_text(m: Module, file: str, attr: A.Attribute) -> str:
    for f in m.files:
        if f.path == file:
            toks = [t for t in f.tokens[attr.expr_tok_start:attr.expr_tok_end]
                    if t.type not in (T.COMMENT, T.NEWLINE)]
            return " ".join(t.text for t in toks)
    return ""


# ---------------------------------------------------------------------------
@rule("unused-variable", WARNING, "variable is declared but never used", "terraform_unused_declarations")
def unused_variable(c: LintContext):
    for name, var in c.m.variables.items():
        if name not in c.v.used_vars:
            yield _d("unused-variable", WARNING, f"variable \"{name}\" is declared but not used",
                     "Nothing in this module refers to it; callers can set it without any effect.",
                     var.file, var.line)


@rule("unused-local", WARNING, "local value is declared but never used", "terraform_unused_declarations")
def unused_local(c: LintContext):
    for name, loc in c.m.locals.items():
        if name not in c.v.used_locals:
            yield _d("unused-local", WARNING, f"local.{name} is declared but not used",
                     "The expression is evaluated and then discarded.", loc.file, loc.line)


@rule("unused-data-source", WARNING, "data source is declared but never read", "terraform_unused_declarations")
def unused_data(c: LintContext):
    for addr, r in c.m.resources.items():
        if r.mode == "data" and addr not in c.v.used_resources:
            yield _d("unused-data-source", WARNING, f"{addr} is declared but not used",
                     "The data source is read on every plan (an API call) and the result is never referenced.",
                     r.file, r.line)


@rule("module-unpinned", WARNING, "module source is not pinned to a version",
      "terraform_module_version / terraform_module_pinned_source")
def module_unpinned(c: LintContext):
    for name, mc in c.m.module_calls.items():
        if mc.source is None or mc.is_local:
            continue
        if mc.is_registry and mc.version is None:
            yield _d("module-unpinned", WARNING, f"module \"{name}\" has no version constraint",
                     f"Registry module \"{mc.source}\" is fetched at whatever version is newest on the day "
                     "\"terraform init\" runs; add a \"version\" argument.", mc.file, mc.line)
        elif not mc.is_registry and "?ref=" not in mc.source and "git" in mc.source:
            yield _d("module-unpinned", WARNING, f"module \"{name}\" git source has no ref",
                     f"\"{mc.source}\" tracks the default branch; append ?ref=<tag or commit>.", mc.file, mc.line)


@rule("required-provider-missing", WARNING, "provider is used but not declared in required_providers",
      "terraform_required_providers")
def required_provider_missing(c: LintContext):
    used = c.m.used_provider_names()
    configured = {p.name: p for p in c.m.providers}
    for name in sorted(set(used) | set(configured)):
        if name in c.m.required_providers or name == "terraform":
            continue
        if name in used:
            first = used[name][0]
            file, line, why = first.file, first.line, f"{first.addr} needs it"
        else:
            p = configured[name]
            file, line, why = p.file, p.line, f"a provider \"{name}\" block configures it"
        yield _d("required-provider-missing", WARNING,
                 f"provider \"{name}\" is not declared in required_providers",
                 f"{why}, so Terraform assumes registry.terraform.io/hashicorp/{name} at any version. "
                 "Declare its source and a version constraint in terraform { required_providers { ... } }.",
                 file, line)


@rule("required-provider-unused", WARNING, "provider is declared in required_providers but never used",
      "terraform_unused_required_providers")
def required_provider_unused(c: LintContext):
    used = set(c.m.used_provider_names()) | {p.name for p in c.m.providers}
    passed_down: set[str] = set()
    for mc in c.m.module_calls.values():
        prov = mc.block.body.attributes.get("providers") if mc.block else None
        if prov is not None:
            for node in prov.expr.walk():
                if isinstance(node, A.ScopeTraversal):
                    passed_down.add(node.root)
    for name, req in c.m.required_providers.items():
        if name not in used and name not in passed_down:
            yield _d("required-provider-unused", WARNING,
                     f"provider \"{name}\" is required but no resource, data source or provider block uses it",
                     f"\"terraform init\" still downloads {req.fqn} and the lock file still pins it.",
                     req.file, req.line)


@rule("provider-version-missing", WARNING, "required_providers entry has no version constraint",
      "terraform_required_providers")
def provider_version_missing(c: LintContext):
    for name, req in c.m.required_providers.items():
        if req.constraint_text is None:
            yield _d("provider-version-missing", WARNING,
                     f"provider \"{name}\" has no version constraint",
                     "Any future release, including breaking major versions, will be accepted.",
                     req.file, req.line)


@rule("required-version-missing", WARNING, "module does not constrain the Terraform version",
      "terraform_required_version")
def required_version_missing(c: LintContext):
    if not c.m.required_versions and c.m.files:
        f = next((f for f in c.m.files if os.path.basename(f.path) in ("versions.tf", "main.tf")), c.m.files[0])
        yield _d("required-version-missing", WARNING, "terraform \"required_version\" is not set",
                 "Add terraform { required_version = \">= x.y.z\" } so unsupported CLI versions fail early.",
                 f.path, 1)


@rule("undocumented", NOTICE, "variable or output has no description",
      "terraform_documented_variables / terraform_documented_outputs")
def undocumented(c: LintContext):
    for name, var in c.m.variables.items():
        if not var.description:
            yield _d("undocumented", NOTICE, f"variable \"{name}\" has no description", "", var.file, var.line)
    for name, out in c.m.outputs.items():
        if not out.description:
            yield _d("undocumented", NOTICE, f"output \"{name}\" has no description", "", out.file, out.line)


@rule("untyped-variable", NOTICE, "variable has no type constraint", "terraform_typed_variables")
def untyped_variable(c: LintContext):
    for name, var in c.m.variables.items():
        if var.type is None and var.block is not None and "type" not in var.block.body.attributes:
            yield _d("untyped-variable", NOTICE, f"variable \"{name}\" has no type", "", var.file, var.line)


@rule("naming-convention", NOTICE, "block name is not snake_case", "terraform_naming_convention")
def naming_convention(c: LintContext):
    def bad(kind: str, name: str, file: str, line: int):
        return _d("naming-convention", NOTICE, f"{kind} name \"{name}\" is not snake_case",
                  "Names with dashes cannot be told apart from subtraction at a glance and break the "
                  "type_name convention of Terraform addresses.", file, line)
    for name, d in c.m.variables.items():
        if not _SNAKE.match(name):
            yield bad("variable", name, d.file, d.line)
    for name, d in c.m.outputs.items():
        if not _SNAKE.match(name):
            yield bad("output", name, d.file, d.line)
    for name, d in c.m.locals.items():
        if not _SNAKE.match(name):
            yield bad("local value", name, d.file, d.line)
    for name, d in c.m.module_calls.items():
        if not _SNAKE.match(name):
            yield bad("module", name, d.file, d.line)
    for r in c.m.resources.values():
        if not _SNAKE.match(r.name):
            yield bad("data source" if r.mode == "data" else "resource", r.name, r.file, r.line)


@rule("deprecated-index", WARNING, "legacy dot index syntax", "terraform_deprecated_index")
def deprecated_index(c: LintContext):
    for file, node in _all_exprs(c.m):
        steps = getattr(node, "steps", None)
        if isinstance(node, (A.ScopeTraversal, A.RelativeTraversal, A.Splat)) and steps:
            for s in steps:
                if s.legacy:
                    yield Diagnostic(WARNING, "List items should be accessed using square brackets",
                                     f"\".{s.value}\" is HCL1 index syntax; write \"[{s.value}]\".",
                                     file, s.line, s.col, "deprecated-index")


@rule("legacy-splat", NOTICE, "attribute-only splat expression (.*)", "terraform_deprecated_index")
def legacy_splat(c: LintContext):
    for file, node in _all_exprs(c.m):
        if isinstance(node, A.Splat) and not node.full:
            yield _d("legacy-splat", NOTICE, "Legacy splat expression",
                     "\".*\" only supports attribute access after it and behaves differently on single "
                     "values; the full splat \"[*]\" is preferred.", file, node)


@rule("comment-syntax", NOTICE, "// comment", "terraform_comment_syntax")
def comment_syntax(c: LintContext):
    for f in c.m.files:
        for t in f.tokens:
            if t.type is T.COMMENT and t.text.startswith("//"):
                yield Diagnostic(NOTICE, "Single line comments should begin with #", "", f.path, t.line, t.col,
                                 "comment-syntax")


@rule("standard-module-structure", NOTICE, "module does not follow the standard file layout",
      "terraform_standard_module_structure")
def standard_module_structure(c: LintContext):
    names = {os.path.basename(f.path): f for f in c.m.files}
    if not names:
        return
    base = os.path.dirname(c.m.files[0].path)
    for want in ("main.tf", "variables.tf", "outputs.tf"):
        if want not in names:
            yield Diagnostic(NOTICE, f"Module should include a {want} file",
                             "Terraform's standard module structure expects main.tf, variables.tf and outputs.tf.",
                             os.path.join(base, want), 0, 0, "standard-module-structure")
    for name, var in c.m.variables.items():
        if os.path.basename(var.file) != "variables.tf":
            yield _d("standard-module-structure", NOTICE,
                     f"variable \"{name}\" should be moved from {os.path.basename(var.file)} to variables.tf", "",
                     var.file, var.line)
    for name, out in c.m.outputs.items():
        if os.path.basename(out.file) != "outputs.tf":
            yield _d("standard-module-structure", NOTICE,
                     f"output \"{name}\" should be moved from {os.path.basename(out.file)} to outputs.tf", "",
                     out.file, out.line)


@rule("eof-newline", NOTICE, "file does not end with a newline")
def eof_newline(c: LintContext):
    seen = set()
    for f in c.m.files:
        seen.add(f.path)
        if f.src and not f.src.endswith("\n"):
            yield Diagnostic(NOTICE, "No newline at end of file",
                             "POSIX text files end with a newline; tools that concatenate or diff files "
                             "mis-handle the last line otherwise.", f.path, f.src.count("\n") + 1, 0, "eof-newline")
    try:
        extra = sorted(n for n in os.listdir(c.m.dir) if n.endswith(".tfvars"))
    except OSError:
        extra = []
    base = os.path.dirname(c.m.files[0].path) if c.m.files else c.m.dir
    for n in extra:
        try:
            with open(os.path.join(c.m.dir, n), "r", encoding="utf-8", newline="") as fh:
                src = fh.read()
        except (OSError, UnicodeDecodeError):
            continue                      # validate reports the unreadable file
        if src and not src.endswith("\n"):
            yield Diagnostic(NOTICE, "No newline at end of file", "", os.path.join(base, n),
                             src.count("\n") + 1, 0, "eof-newline")


def _whole_object_ref(c: LintContext, e: A.Node) -> Optional[str]:
    """Address when ``e`` names a whole single-instance resource / data source /
    module (an object), not one of its attributes."""
    if not isinstance(e, A.ScopeTraversal) or any(s.kind != "attr" for s in e.steps):
        return None
    if e.root == "data" and len(e.steps) == 2:
        r = c.m.resources.get(f"data.{e.steps[0].value}.{e.steps[1].value}")
    elif e.root == "module" and len(e.steps) == 1:
        mc = c.m.module_calls.get(e.steps[0].value)
        return e.dotted() if mc and not (mc.has_count or mc.has_for_each) else None
    elif e.root not in ("var", "local", "data", "module", "path", "each", "count", "self", "terraform") \
            and len(e.steps) == 1:
        r = c.m.resources.get(f"{e.root}.{e.steps[0].value}")
    else:
        return None
    if r is not None and not (r.has_count or r.has_for_each):
        return r.addr
    return None


@rule("length-of-object", WARNING, "length() of a whole resource object is a constant")
def length_of_object(c: LintContext):
    for file, node in _all_exprs(c.m):
        if isinstance(node, A.FunctionCall) and node.name == "length" and len(node.args) == 1:
            addr = _whole_object_ref(c, node.args[0])
            if addr:
                yield _d("length-of-object", WARNING, f"length({addr}) counts the object's attributes",
                         f"{addr} is a single object, so its length is the number of attributes in its schema "
                         "-- a constant that is never 0. A condition built on it cannot do what it appears to; "
                         "take the length of a list attribute instead (for example .ids).", file, node)


_META = {"count", "for_each", "depends_on", "provider", "lifecycle"}


@rule("duplicate-resource-arguments", WARNING, "two resources of one type have identical arguments")
def duplicate_resource_arguments(c: LintContext):
    by_type: dict[tuple[str, str], list[Resource]] = {}
    for r in c.m.resources.values():
        by_type.setdefault((r.mode, r.type), []).append(r)
    for group in by_type.values():
        seen: dict[tuple, Resource] = {}
        for r in sorted(group, key=lambda r: (r.file, r.line)):
            attrs = r.block.body.attributes if r.block else {}
            sig = tuple(sorted((n, _expr_text(c.m, r.file, a)) for n, a in attrs.items() if n not in _META))
            if not sig or not any("." in text for _, text in sig):
                continue       # nothing but constants: too little to judge
            if r.block.body.blocks:
                continue
            if sig in seen:
                first = seen[sig]
                yield _d("duplicate-resource-arguments", WARNING,
                         f"{r.addr} has exactly the same arguments as {first.addr}",
                         f"Both blocks create the same thing ({first.file}:{first.line}). When the names "
                         "differ but the arguments do not, one of them was usually meant to point elsewhere.",
                         r.file, r.line)
            else:
                seen[sig] = r


def _flag_of(e: Optional[A.Node]) -> Optional[str]:
    """Name of the single ``var.<flag>`` a condition depends on."""
    if e is None:
        return None
    names = {n.steps[0].value for n in e.walk()
             if isinstance(n, A.ScopeTraversal) and n.root == "var" and n.steps}
    return next(iter(names)) if len(names) == 1 else None


@rule("guard-mismatch", WARNING, "conditional reads a resource that is switched by a different flag")
def guard_mismatch(c: LintContext):
    switched: dict[str, str] = {}
    for addr, r in c.m.resources.items():
        if r.has_count and r.block is not None:
            ce = r.block.body.attributes["count"].expr
            if isinstance(ce, A.Conditional):
                flag = _flag_of(ce.cond)
                if flag:
                    switched[addr] = flag
    for file, node in _all_exprs(c.m):
        if not isinstance(node, A.Conditional):
            continue
        guard = _flag_of(node.cond)
        if not guard:
            continue
        for branch in (node.true, node.false):
            for ref in branch.walk():
                if isinstance(ref, A.ScopeTraversal) and len(ref.steps) >= 2:
                    addr = (f"data.{ref.steps[0].value}.{ref.steps[1].value}" if ref.root == "data"
                            else f"{ref.root}.{ref.steps[0].value}")
                    flag = switched.get(addr)
                    if flag and flag != guard:
                        yield _d("guard-mismatch", WARNING,
                                 f"{addr} exists only when var.{flag} is set, but this expression is guarded by "
                                 f"var.{guard}",
                                 f"With var.{guard} on and var.{flag} off the index is out of range; with the "
                                 "flags the other way round the value is hidden although it exists.", file, ref)


@rule("secret-in-command", WARNING, "credential interpolated into a local-exec command line")
def secret_in_command(c: LintContext):
    for r in c.m.resources.values():
        if r.block is None:
            continue
        for blk in r.block.body.walk():
            if isinstance(blk, A.Block) and blk.type == "provisioner" and blk.label_values == ["local-exec"]:
                cmd = blk.body.attributes.get("command")
                if cmd is None:
                    continue
                for ref in cmd.expr.walk():
                    if isinstance(ref, A.ScopeTraversal) and ref.steps:
                        last = ref.steps[-1]
                        if last.kind == "attr" and _SECRETISH.search(str(last.value)):
                            yield _d("secret-in-command", WARNING,
                                     f"{ref.dotted()} is placed on a command line",
                                     "Command-line arguments are visible in the process table and in Terraform's "
                                     "own output; pass secrets through the provisioner's \"environment\" map or a "
                                     "file instead.", r.file, ref)


_PAIRS = (("cpu", "gpu"), ("public", "private"), ("primary", "secondary"), ("source", "destination"),
          ("ingress", "egress"), ("blue", "green"), ("reader", "writer"))


def _tokens(name: str) -> set[str]:
    return set(re.split(r"[_\-.]+", name.lower()))


def _sides(name: str) -> set[tuple[int, int]]:
    """(pair index, side) for each pair of which ``name`` carries exactly one side."""
    toks = _tokens(name)
    out = set()
    for i, (a, b) in enumerate(_PAIRS):
        if (a in toks) != (b in toks):
            out.add((i, 0 if a in toks else 1))
    return out


def _opposites(owner: str, subtree: A.Node) -> Iterator[tuple[A.ScopeTraversal, str, str]]:
    mine = _sides(owner)
    if not mine:
        return
    for ref in subtree.walk():
        if not isinstance(ref, A.ScopeTraversal):
            continue
        for step in ref.steps:
            if step.kind != "attr":
                continue
            for (i, side) in _sides(str(step.value)):
                if (i, 1 - side) in mine:
                    yield ref, _PAIRS[i][1 - side], _PAIRS[i][side]
                    break
            else:
                continue
            break


@rule("counterpart-mismatch", NOTICE, "a cpu/gpu (public/private, ...) block refers to its counterpart's value")
def counterpart_mismatch(c: LintContext):
    def note(file: str, owner: str, kind: str, ref: A.ScopeTraversal, own: str, other: str):
        return _d("counterpart-mismatch", NOTICE,
                  f"{kind} \"{owner}\" is the {own} one but reads {ref.dotted()}",
                  f"Names pair up as {own}/{other}; a {own} block reading a {other} value is often a copy-paste "
                  "slip. Ignore if intended.", file, ref)
    for r in c.m.resources.values():
        if r.block is not None:
            for ref, own, other in _opposites(r.name, r.block.body):
                yield note(r.file, r.name, "data source" if r.mode == "data" else "resource", ref, own, other)
    for name, o in c.m.outputs.items():
        if o.block is not None:
            for ref, own, other in _opposites(name, o.block.body):
                yield note(o.file, name, "output", ref, own, other)
    for file, node in _all_exprs(c.m):
        if isinstance(node, A.ObjectCons):
            for key, value in node.items:
                k = key.literal_key()
                if k:
                    for ref, own, other in _opposites(k, value):
                        yield note(file, k, "object key", ref, own, other)


@rule("duplicate-object-key", WARNING, "object constructor sets the same key twice", "terraform_map_duplicate_keys")
def duplicate_object_key(c: LintContext):
    for file, node in _all_exprs(c.m):
        if isinstance(node, A.ObjectCons):
            seen: dict[str, A.Node] = {}
            for key, _ in node.items:
                k = key.literal_key()
                if k is None:
                    continue
                if k in seen:
                    yield _d("duplicate-object-key", WARNING, f"key \"{k}\" is set twice in one object",
                             f"The later value silently replaces the one at line {seen[k].pos.line}.", file, key)
                else:
                    seen[k] = key


@rule("empty-list-equality", WARNING, "comparison with [] is always false", "terraform_empty_list_equality")
def empty_list_equality(c: LintContext):
    for file, node in _all_exprs(c.m):
        if isinstance(node, A.Binary) and node.op in ("==", "!="):
            for side in (node.lhs, node.rhs):
                if isinstance(side, A.TupleCons) and not side.items:
                    yield _d("empty-list-equality", WARNING, "Comparing a collection with an empty list",
                             "[] is an empty tuple, which never equals a list(...) value, so this is always "
                             f"{'false' if node.op == '==' else 'true'}; use length(x) == 0 instead.", file, node)
                    break


@rule("deprecated-lookup", NOTICE, "lookup() with two arguments", "terraform_deprecated_lookup")
def deprecated_lookup(c: LintContext):
    for file, node in _all_exprs(c.m):
        if isinstance(node, A.FunctionCall) and node.name == "lookup" and len(node.args) == 2 \
                and not node.expand_final:
            yield _d("deprecated-lookup", NOTICE, "lookup() without a default",
                     "With two arguments lookup(m, k) is just m[k]; the index syntax is preferred.", file, node)


@rule("sensitive-output", WARNING, "output exposes a sensitive variable without being marked sensitive")
def sensitive_output(c: LintContext):
    secret = {n for n, v in c.m.variables.items() if v.sensitive}
    if not secret:
        return
    for name, o in c.m.outputs.items():
        if o.sensitive or o.block is None or "value" not in o.block.body.attributes:
            continue
        for ref in o.block.body.attributes["value"].expr.walk():
            if isinstance(ref, A.ScopeTraversal) and ref.root == "var" and ref.steps \
                    and ref.steps[0].value in secret:
                yield _d("sensitive-output", WARNING,
                         f"output \"{name}\" is built from sensitive var.{ref.steps[0].value}",
                         "Terraform stops the plan with \"Output refers to sensitive values\" unless the output "
                         "sets sensitive = true.", o.file, ref)
                break


@rule("provider-in-module", NOTICE, "provider configuration inside a module that other modules call")
def provider_in_module(c: LintContext):
    # A module with its own provider blocks cannot be used with count, for_each
    # or depends_on by its callers.  Flag the callee when we can see a caller.
    for name, child in c.children.items():
        if child is None or not child.providers:
            continue
        mc = c.m.module_calls[name]
        p = child.providers[0]
        attrs = mc.block.body.attributes if mc.block else {}
        if any(a in attrs for a in ("count", "for_each", "depends_on")):
            continue            # validate already reports that combination as an error
        yield _d("provider-in-module", NOTICE,
                 f"module \"{name}\" ({mc.source}) contains its own provider \"{p.name}\" configuration",
                 "A module with provider blocks is a legacy module: callers cannot use count, for_each or "
                 "depends_on with it, and removing the call later orphans its resources' provider "
                 f"(declared at {p.file}:{p.line}).", mc.file, mc.line)


@rule("local-exec", NOTICE, "provisioner shells out to a local CLI")
def local_exec(c: LintContext):
    for r in c.m.resources.values():
        if r.block is None:
            continue
        for blk in r.block.body.walk():
            if isinstance(blk, A.Block) and blk.type == "provisioner" and blk.label_values == ["local-exec"]:
                cmd = blk.body.attributes.get("command")
                text = _expr_text(c.m, r.file, cmd) if cmd is not None else ""
                m_ = re.search(r'"\s*([A-Za-z0-9_.-]+)', text)
                tool = m_.group(1) if m_ else "a command"
                yield _d("local-exec", NOTICE, f"{r.addr} runs \"{tool}\" on the operator's machine",
                         "The module only works where that CLI is installed and logged in, the step is invisible "
                         "to \"terraform plan\", and it is not undone by \"terraform destroy\".", r.file, blk)


@rule("helper-script", WARNING, "shell helper beside the module is unreferenced, has an inert shebang or evals input")
def helper_script(c: LintContext):
    """Shell scripts shipped next to ``.tf`` files are meant for ``external`` data
    sources or ``local-exec`` (the reference ships eks/setup-kube-config.sh)."""
    try:
        scripts = sorted(n for n in os.listdir(c.m.dir) if n.endswith((".sh", ".bash")))
    except OSError:
        return
    if not scripts:
        return
    base = os.path.dirname(c.m.files[0].path) if c.m.files else c.m.dir
    hcl_text = "\n".join(f.src for f in c.m.files)
    for name in scripts:
        try:
            with open(os.path.join(c.m.dir, name), "r", encoding="utf-8", errors="replace") as fh:
                lines = fh.read().split("\n")
        except OSError:
            continue
        disp = os.path.join(base, name)
        if name not in hcl_text:
            yield Diagnostic(WARNING, f"{name} is not referenced by any .tf file of this module",
                             "No external data source, provisioner, file() or templatefile() call names it; it is "
                             "dead code or undocumented manual tooling.", disp, 1, 0, "helper-script")
        shebang = next((i for i, ln in enumerate(lines) if ln.startswith("#!")), None)
        if shebang is not None and shebang > 0:
            yield Diagnostic(WARNING, f"{name}: the #! line is on line {shebang + 1}, not line 1",
                             "The kernel only honours an interpreter line that is the very first line; here the "
                             "script runs under whatever shell the caller happens to use.", disp, shebang + 1, 0,
                             "helper-script")
        for i, ln in enumerate(lines):
            if re.search(r"(^|[;&|]\s*)eval\s", ln.strip()):
                yield Diagnostic(WARNING, f"{name}: eval of generated text",
                                 "eval executes whatever the expansion produces; with input that arrives on stdin "
                                 "(the external data source protocol) that is command injection unless every value "
                                 "is shell-quoted.", disp, i + 1, 0, "helper-script")


# ---------------------------------------------------------------------------
def run_rules(result: ValidateResult, only: Optional[set[str]] = None,
              skip: Optional[set[str]] = None) -> list[Diagnostic]:
    # make sure every local child the module calls has been loaded
    for mc in result.module.module_calls.values():
        result.validator.child(mc)
    ctx = LintContext(result, dict(result.validator.children))
    out: list[Diagnostic] = []
    for r in RULES:
        if only and r.id not in only:
            continue
        if skip and r.id in skip:
            continue
        out.extend(r.fn(ctx))
    ignored = _ignore_comments(result.module)
    out = [d for d in out if not _is_ignored(d, ignored)]
    out.sort(key=lambda d: (d.file, d.line, d.col, d.rule))
    return out


_IGNORE_RE = re.compile(r"^(?:#|//)\s*tfcheck-ignore:\s*([A-Za-z0-9_,\- ]+)")


def _ignore_comments(m: Module) -> dict[tuple[str, int], set[str]]:
    """``# tfcheck-ignore: rule-a, rule-b`` silences those rules (or ``all``) on
    its own line and on the line that follows it."""
    table: dict[tuple[str, int], set[str]] = {}
    for f in m.files:
        for t in f.tokens:
            if t.type is T.COMMENT:
                hit = _IGNORE_RE.match(t.text)
                if hit:
                    rules = {r.strip() for r in hit.group(1).split(",") if r.strip()}
                    for line in (t.line, t.line + 1):
                        table.setdefault((f.path, line), set()).update(rules)
    return table


def _is_ignored(d: Diagnostic, table: dict[tuple[str, int], set[str]]) -> bool:
    rules = table.get((d.file, d.line))
    return bool(rules) and (d.rule in rules or "all" in rules)
