"""Module model: what one Terraform module directory declares.

A Terraform *module* is all ``*.tf`` files of one directory read as a single
configuration.  The reference has six of them: ``eks/``, ``gke/``, ``aks/`` and
one ``examples/cnpack`` below each.  This loader restates the
structural rules of Terraform's configuration loader (top-level block schema,
label counts, name syntax, duplicate declarations, meta-arguments) -- the part
of ``terraform validate`` that needs no provider plugin.  Provider schemas are
NOT available offline, so arguments *inside* resource/data/provider bodies are
never judged; DESIGN.md lists what that leaves unchecked.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Optional

from ..hcl import ast as A
from ..hcl.diagnostics import Diagnostic, ERROR, WARNING, has_errors
from ..hcl.parser import parse_path
from ..hcl.scanner import is_valid_identifier
from ..hcl.tokens import Token
from . import tfversion, typeexpr
from .tfversion import Restated
from .versions import Constraints, try_parse_constraints

_BAD_NAME = ("A name must start with a letter or underscore and may contain only letters, digits, "
             "underscores, and dashes.")

# top-level block types -> number of labels
TOP_LEVEL_BLOCKS = {
    "terraform": 0, "provider": 1, "variable": 1, "locals": 0, "output": 1, "module": 1,
    "resource": 2, "data": 2, "moved": 0, "import": 0, "check": 1, "removed": 0, "ephemeral": 2,
}
RESERVED_VARIABLE_NAMES = {"source", "version", "providers", "count", "for_each", "lifecycle",
                           "depends_on", "locals"}
VARIABLE_ARGS = {"description", "default", "type", "sensitive", "nullable", "ephemeral"}
OUTPUT_ARGS = {"description", "value", "sensitive", "depends_on", "ephemeral"}
LIFECYCLE_ARGS = {"create_before_destroy", "prevent_destroy", "ignore_changes", "replace_triggered_by"}


@dataclass
class SourceFile:
    path: str                 # display path
    src: str
    body: A.Body
    tokens: list[Token]


@dataclass
class Decl:
    """Anything declared by a block or attribute, with where it was declared."""
    name: str
    file: str
    line: int
    block: Optional[A.Block] = None
    attr: Optional[A.Attribute] = None


@dataclass
class Variable(Decl):
    type: Optional[typeexpr.Type] = None
    has_default: bool = False
    default: object = None
    description: Optional[str] = None
    sensitive: bool = False
    nullable: Optional[bool] = None

    @property
    def required(self) -> bool:
        return not self.has_default


@dataclass
class Output(Decl):
    description: Optional[str] = None
    sensitive: bool = False


@dataclass
class Resource(Decl):
    mode: str = "managed"     # managed | data | ephemeral
    type: str = ""
    has_count: bool = False
    has_for_each: bool = False
    provider_ref: Optional[str] = None      # provider = google-beta / aws.west
    scope: Optional[str] = None             # name of the check block a nested data source lives in

    @property
    def addr(self) -> str:
        return resource_addr(self.mode, self.type, self.name)

    @property
    def implied_provider(self) -> str:
        if self.provider_ref:
            return self.provider_ref.split(".")[0]
        return self.type.split("_", 1)[0]


def resource_addr(mode: str, typ: str, name: str) -> str:
    return ("" if mode == "managed" else mode + ".") + f"{typ}.{name}"


@dataclass
class ModuleCall(Decl):
    source: Optional[str] = None
    version: Optional[str] = None
    has_count: bool = False
    has_for_each: bool = False

    @property
    def is_local(self) -> bool:
        return bool(self.source) and (self.source.startswith("./") or self.source.startswith("../")
                                      or self.source in (".", ".."))

    @property
    def is_registry(self) -> bool:
        s = self.source or ""
        if self.is_local or "::" in s or s.startswith(("git@", "http://", "https://", "github.com/",
                                                       "bitbucket.org/", "/")):
            return False
        parts = s.split("//", 1)[0].split("/")
        return len(parts) in (3, 4)


@dataclass
class ProviderConfig(Decl):
    alias: Optional[str] = None


@dataclass
class RequiredProvider(Decl):
    source: Optional[str] = None
    constraint_text: Optional[str] = None
    constraints: Optional[Constraints] = None
    aliases: tuple = ()       # configuration_aliases = [aws.src, aws.dst] -> ("src", "dst")

    @property
    def display(self) -> str:
        """``hashicorp/aws``: the address the way Terraform prints it in messages."""
        host, _, rest = self.fqn.partition("/")
        return rest if host == "registry.terraform.io" else self.fqn

    @property
    def fqn(self) -> str:
        """registry.terraform.io/<namespace>/<type>"""
        src = self.source or f"hashicorp/{self.name}"
        parts = src.split("/")
        if len(parts) == 1:
            parts = ["hashicorp"] + parts
        if len(parts) == 2:
            parts = ["registry.terraform.io"] + parts
        return "/".join(parts).lower()     # provider addresses compare case-insensitively


@dataclass
class Module:
    dir: str
    files: list[SourceFile] = field(default_factory=list)
    variables: dict[str, Variable] = field(default_factory=dict)
    locals: dict[str, Decl] = field(default_factory=dict)
    outputs: dict[str, Output] = field(default_factory=dict)
    resources: dict[str, Resource] = field(default_factory=dict)     # keyed by addr
    module_calls: dict[str, ModuleCall] = field(default_factory=dict)
    providers: list[ProviderConfig] = field(default_factory=list)
    required_providers: dict[str, RequiredProvider] = field(default_factory=dict)
    required_versions: list[tuple[str, Decl]] = field(default_factory=list)
    diags: list[Diagnostic] = field(default_factory=list)
    broken_files: set = field(default_factory=set)       # files with syntax errors
    display_root: Optional[str] = None                   # paths in messages are relative to this
    restated: Restated = field(default_factory=lambda: Restated(tfversion.NEWEST, "default"))
    experiments: set = field(default_factory=set)        # active language experiments

    def used_provider_names(self) -> dict[str, list[Resource]]:
        out: dict[str, list[Resource]] = {}
        for r in self.resources.values():
            out.setdefault(r.implied_provider, []).append(r)
        return out


# ---------------------------------------------------------------------------
class _Loader:
    def __init__(self, directory: str, display_root: Optional[str] = None, restated: Optional[Restated] = None):
        self.m = Module(directory)
        self.m.display_root = display_root
        self.root = display_root
        self.forced = restated

    @property
    def v(self) -> Restated:
        return self.m.restated

    def needs(self, feature: str, what: str) -> str:
        """The sentence appended to Terraform's own wording when something is refused
        only because of the release being restated."""
        return (f"\n\n({what} needs Terraform {tfversion.text(tfversion.SINCE[feature])} or later; this run "
                f"restates {self.v.describe()}.)")

    def sniff_required_versions(self) -> list[tuple[str, str]]:
        found = []
        for f in self.m.files:
            for item in f.body.items:
                if isinstance(item, A.Block) and item.type == "terraform":
                    attr = item.body.attributes.get("required_version")
                    val = typeexpr.literal_value(attr.expr) if attr is not None else None
                    if isinstance(val, str):
                        found.append((val, f"{f.path}:{attr.pos.line}"))
        return found

    def diag(self, sev: str, summary: str, detail: str, file: str, node, rule: str) -> None:
        line = node.pos.line if isinstance(node, A.Node) else int(node)
        col = node.pos.col if isinstance(node, A.Node) else 0
        self.m.diags.append(Diagnostic(sev, summary, detail, file, line, col, rule))

    def display(self, path: str) -> str:
        if self.root:
            return os.path.relpath(path, self.root)
        return path

    def load(self) -> Module:
        try:
            names = sorted(os.listdir(self.m.dir))
        except OSError as e:
            self.m.diags.append(Diagnostic(ERROR, "Failed to read module directory",
                                           f"Module directory {self.m.dir} does not exist or cannot be read: {e}",
                                           self.m.dir, rule="module-dir"))
            return self.m
        tf = [n for n in names if n.endswith(".tf") and not n.startswith((".", "~", "#"))]
        for name in tf:
            path = os.path.join(self.m.dir, name)
            disp = self.display(path)
            src, body, diags, tokens = parse_path(path, disp)
            self.m.diags.extend(diags)
            self.m.files.append(SourceFile(disp, src, body, tokens))
            if has_errors(diags):
                self.m.broken_files.add(disp)
        sniffed = self.sniff_required_versions()
        self.m.restated = self.forced or tfversion.from_required_version(sniffed)
        if self.forced is not None and self.forced.requested is not None:
            # the release named on the command line must be one the module accepts, exactly as
            # a Terraform binary of that release would refuse the configuration
            asked = self.forced.requested
            for text, where in sniffed:
                cons, _ = try_parse_constraints(text)
                if cons is not None and not tfversion.series_admitted(asked, [cons]):
                    file, _, line = where.rpartition(":")
                    self.m.diags.append(Diagnostic(
                        ERROR, "Unsupported Terraform Core version",
                        f"This configuration does not support Terraform version {tfversion.text(asked)}. To proceed, "
                        "either choose another supported Terraform version or update this version constraint. "
                        "Version constraints are normally set for good reason, so updating the constraint may lead "
                        f"to other errors or unexpected behavior.\n\n(required_version = \"{text}\")",
                        file, int(line), 0, "terraform-version"))
        # experiments apply to the whole module, whichever file happens to be read first
        for f in self.m.files:
            if f.path not in self.m.broken_files:
                for b in f.body.blocks_of("terraform"):
                    for it in self._experiment_keywords(b):
                        if it.root == "module_variable_optional_attrs" and not self.v.has("optional-attributes"):
                            self.m.experiments.add(it.root)
        for f in self.m.files:
            # a file that does not parse contributes no declarations: what the
            # recovering parser salvaged would only produce follow-on noise
            if f.path not in self.m.broken_files:
                self.top_level(f)
        return self.m

    # -- top level -----------------------------------------------------------
    def top_level(self, f: SourceFile) -> None:
        for item in f.body.items:
            if isinstance(item, A.Attribute):
                self.diag(ERROR, "Unsupported argument",
                          f"An argument named \"{item.name}\" is not expected here.", f.path, item, "structure")
                continue
            want = TOP_LEVEL_BLOCKS.get(item.type)
            feature = f"block:{item.type}"
            if want is not None and feature in tfversion.SINCE and not self.v.has(feature):
                self.diag(ERROR, "Unsupported block type",
                          f"Blocks of type \"{item.type}\" are not expected here."
                          + self.needs(feature, f"A \"{item.type}\" block"), f.path, item, "terraform-version")
                continue
            if want is None:
                self.diag(ERROR, "Unsupported block type",
                          f"Blocks of type \"{item.type}\" are not expected here.", f.path, item, "structure")
                continue
            if not self.check_labels(f, item, want):
                continue
            getattr(self, "block_" + item.type, self.block_other)(f, item)

    def check_labels(self, f: SourceFile, b: A.Block, want: int) -> bool:
        want_names = "type, name" if want == 2 else "name"
        if len(b.labels) < want:
            missing = "type" if want == 2 and not b.labels else "name"
            self.diag(ERROR, f"Missing {missing} for {b.type}",
                      f"All {b.type} blocks must have {want} labels ({want_names}).", f.path, b, "structure")
            return False
        if len(b.labels) > want:
            detail = (f"Only {want} labels ({want_names}) are expected for {b.type} blocks." if want
                      else f"No labels are expected for {b.type} blocks.")
            self.diag(ERROR, f"Extraneous label for {b.type}", detail, f.path, b, "structure")
            return False
        return True

    def check_name(self, f: SourceFile, b: A.Block, value: str, what: str) -> bool:
        if not is_valid_identifier(value):
            self.diag(ERROR, f"Invalid {what}", _BAD_NAME, f.path, b, "structure")
            return False
        return True

    def check_args(self, f: SourceFile, b: A.Block, allowed: set[str], blocks: dict[str, int],
                   gates: str = "") -> None:
        """``gates``: prefix under which tfversion.SINCE lists the arguments and nested
        blocks of this block type that newer releases added (``"variable"`` ->
        ``variable:nullable`` ...)."""
        def too_new(name: str) -> Optional[str]:
            feature = f"{gates}:{name}"
            return feature if gates and feature in tfversion.SINCE and not self.v.has(feature) else None

        for item in b.body.items:
            if isinstance(item, A.Attribute):
                if item.name in allowed and too_new(item.name):
                    self.diag(ERROR, "Unsupported argument",
                              f"An argument named \"{item.name}\" is not expected here."
                              + self.needs(too_new(item.name), f"The \"{item.name}\" argument"),
                              f.path, item, "terraform-version")
                elif item.name not in allowed:
                    self.diag(ERROR, "Unsupported argument",
                              f"An argument named \"{item.name}\" is not expected here.", f.path, item, "structure")
            elif item.type in blocks and too_new(item.type):
                self.diag(ERROR, "Unsupported block type",
                          f"Blocks of type \"{item.type}\" are not expected here."
                          + self.needs(too_new(item.type), f"A \"{item.type}\" block here"),
                          f.path, item, "terraform-version")
            elif item.type not in blocks:
                self.diag(ERROR, "Unsupported block type",
                          f"Blocks of type \"{item.type}\" are not expected here.", f.path, item, "structure")
            elif len(item.labels) != blocks[item.type]:
                self.diag(ERROR, f"Wrong number of labels for {item.type}",
                          f"{item.type} blocks take {blocks[item.type]} label(s).", f.path, item, "structure")
        seen: dict[str, A.Attribute] = {}
        for item in b.body.items:
            if isinstance(item, A.Attribute):
                if item.name in seen:
                    self.diag(ERROR, "Attribute redefined",
                              f"The argument \"{item.name}\" was already set at {f.path}:{seen[item.name].pos.line}. "
                              "Each argument may be set only once.", f.path, item, "structure")
                seen.setdefault(item.name, item)

    def duplicate(self, f: SourceFile, b: A.Node, table: dict, key: str, summary: str, detail: str) -> bool:
        if key in table:
            prev = table[key]
            self.diag(ERROR, summary, detail.format(prev=f"{prev.file}:{prev.line}"), f.path, b, "duplicate")
            return True
        return False

    def literal(self, f: SourceFile, attr: A.Attribute, kind: type, what: str):
        v = typeexpr.literal_value(attr.expr)
        if v is typeexpr.UNKNOWN:
            self.no_variables(f, attr.expr)
            return None
        if kind is bool and isinstance(v, str) and v in ("true", "false"):
            v = v == "true"
        if not isinstance(v, kind) or (kind is not bool and isinstance(v, bool)):
            self.diag(ERROR, f"Invalid {what}", f"A {kind.__name__ if kind is not str else 'string'} "
                      f"value is required for \"{attr.name}\".", f.path, attr.expr, "structure")
            return None
        return v

    def no_variables(self, f: SourceFile, expr: A.Expr) -> None:
        for node in expr.walk():
            if isinstance(node, A.ScopeTraversal):
                self.diag(ERROR, "Variables not allowed", "Variables may not be used here.",
                          f.path, node, "structure")
                return
            if isinstance(node, A.FunctionCall):
                self.diag(ERROR, "Function calls not allowed", "Functions may not be called here.",
                          f.path, node, "structure")
                return

    @staticmethod
    def _free_references(expr: A.Expr) -> list[A.ScopeTraversal]:
        """References in ``expr`` that are not for-expression iterator names."""
        out: list[A.ScopeTraversal] = []

        def visit(n: A.Node, bound: frozenset) -> None:
            if isinstance(n, A.ScopeTraversal):
                if n.root not in bound:
                    out.append(n)
                return
            if isinstance(n, (A.ForExpr, A.TemplateFor)):
                visit(n.coll, bound)
                inner = bound | {x for x in (n.key_var, n.val_var) if x}
                for ch in n.children():
                    if ch is not n.coll:
                        visit(ch, inner)
                return
            for ch in n.children():
                visit(ch, bound)

        visit(expr, frozenset())
        return out

    # -- block handlers ----------------------------------------------------
    def block_other(self, f: SourceFile, b: A.Block) -> None:
        pass

    def require(self, f: SourceFile, b: A.Block, *names: str) -> None:
        attrs = b.body.attributes
        for n in names:
            if n not in attrs:
                self.diag(ERROR, "Missing required argument",
                          f"The argument \"{n}\" is required, but no definition was found.", f.path, b, "structure")

    def condition_block(self, f: SourceFile, b: A.Block) -> None:
        """precondition / postcondition / validation / assert"""
        self.check_args(f, b, {"condition", "error_message"}, {})
        self.require(f, b, "condition", "error_message")

    def block_moved(self, f: SourceFile, b: A.Block) -> None:
        self.check_args(f, b, {"from", "to"}, {})
        self.require(f, b, "from", "to")

    def block_import(self, f: SourceFile, b: A.Block) -> None:
        self.check_args(f, b, {"to", "id", "provider", "for_each"}, {}, gates="import")
        self.require(f, b, "to", "id")

    def block_removed(self, f: SourceFile, b: A.Block) -> None:
        self.check_args(f, b, {"from"}, {"lifecycle": 0, "provisioner": 1}, gates="removed")
        self.require(f, b, "from")

    def block_check(self, f: SourceFile, b: A.Block) -> None:
        if not self.check_name(f, b, b.labels[0].value, "check block name"):
            return
        self.check_args(f, b, set(), {"assert": 0, "data": 2})
        asserts = b.body.blocks_of("assert")
        if not asserts:
            self.diag(ERROR, "Invalid check block", "Check blocks must have at least one assert block.",
                      f.path, b, "structure")
        for a in asserts:
            self.condition_block(f, a)
        nested = [d for d in b.body.blocks_of("data") if len(d.labels) == 2]
        if len(nested) > 1:
            self.diag(ERROR, "Multiple nested data resources",
                      "Check blocks can only contain one nested data resource.", f.path, b, "structure")
        for d in nested[:1]:
            # the scoped data source shares the module's data-source namespace, but only
            # its own check block may refer to it
            self._resource(f, d, "data", scope=b.labels[0].value)

    def block_variable(self, f: SourceFile, b: A.Block) -> None:
        name = b.labels[0].value
        if not self.check_name(f, b, name, "variable name"):
            return
        if name in RESERVED_VARIABLE_NAMES:
            self.diag(ERROR, "Invalid variable name",
                      f"The variable name \"{name}\" is reserved due to its special meaning inside module blocks.",
                      f.path, b, "structure")
        self.check_args(f, b, VARIABLE_ARGS, {"validation": 0}, gates="variable")
        for vb in b.body.blocks_of("validation"):
            self.condition_block(f, vb)
            cond = vb.body.attributes.get("condition")
            if cond is not None:
                refs = [n for n in cond.expr.walk() if isinstance(n, A.ScopeTraversal)]
                own = any(r.root == "var" and r.steps and r.steps[0].value == name for r in refs)
                if not own:
                    self.diag(ERROR, "Invalid variable validation condition",
                              f"The condition for variable \"{name}\" must refer to var.{name} in order to test "
                              "incoming values.", f.path, cond.expr, "structure")
            if not self.v.has("validation-refers-to-other-objects"):
                # until 1.9 a validation rule could see nothing but its own variable
                for arg, what in (("condition", "condition"), ("error_message", "error message")):
                    attr = vb.body.attributes.get(arg)
                    for r in self._free_references(attr.expr) if attr is not None else ():
                        if not (r.root == "var" and r.steps and r.steps[0].value == name):
                            self.diag(ERROR, "Invalid reference in variable validation",
                                      f"The {what} for variable \"{name}\" can only refer to the variable itself, "
                                      f"using var.{name}."
                                      + self.needs("validation-refers-to-other-objects",
                                                   "A validation rule that refers to other objects"),
                                      f.path, r, "terraform-version")
        if self.duplicate(f, b, self.m.variables, name, "Duplicate variable declaration",
                          f"A variable named \"{name}\" was already declared at {{prev}}. "
                          "Variable names must be unique within a module."):
            return
        v = Variable(name, f.path, b.pos.line, block=b)
        attrs = b.body.attributes
        if "type" in attrs:
            try:
                v.type = self.quoted_type(f, attrs["type"].expr) or typeexpr.parse_type(attrs["type"].expr)
            except typeexpr.TypeError_ as e:
                self.diag(ERROR, "Invalid type specification", str(e), f.path, e.node, "type")
            self.optional_attributes(f, b, attrs["type"].expr)
        if "description" in attrs:
            v.description = self.literal(f, attrs["description"], str, "description")
        if "sensitive" in attrs:
            v.sensitive = bool(self.literal(f, attrs["sensitive"], bool, "sensitive flag"))
        if "nullable" in attrs:
            v.nullable = self.literal(f, attrs["nullable"], bool, "nullable flag")
        if "default" in attrs:
            v.has_default = True
            self.no_variables(f, attrs["default"].expr)
            v.default = typeexpr.literal_value(attrs["default"].expr)
            if v.type is not None:
                err = typeexpr.conformance_error(v.default, v.type)
                if err:
                    self.diag(ERROR, "Invalid default value for variable",
                              f"This default value is not compatible with the variable's type constraint: {err}.",
                              f.path, attrs["default"].expr, "type")
            if v.default is None and v.nullable is False:
                self.diag(ERROR, "Invalid default value for variable",
                          "A null default value is not valid when nullable=false.",
                          f.path, attrs["default"].expr, "type")
        self.m.variables[name] = v

    def quoted_type(self, f: SourceFile, expr: A.Expr) -> Optional[typeexpr.Type]:
        """``type = "string"`` / ``"list"`` / ``"map"``: the 0.11 spelling.  0.12-0.14 read it
        with a deprecation warning, 0.15 turned it into an error (``fmt`` rewrites it either
        way: fmt/rewrites.py).  Returns the type when the old spelling was understood."""
        word = expr.literal_value() if isinstance(expr, A.Template) and not expr.heredoc else None
        if word is None:
            return None
        better = {"string": "", "list": " and write list(string) instead to explicitly indicate that the list "
                  "elements are strings", "map": " and write map(string) instead to explicitly indicate that the "
                  "map elements are strings"}
        if word not in better:
            self.diag(ERROR, "Invalid legacy variable type hint",
                      "To provide a full type expression, remove the surrounding quotes and give the type "
                      "expression directly.", f.path, expr, "type")
            return typeexpr.Type("any")
        if self.v.release >= tfversion.LEGACY_FUNCTIONS_REMOVED:        # the same clean-up release, 0.15
            self.diag(ERROR, "Invalid quoted type constraints",
                      "Terraform 0.11 and earlier required type constraints to be given in quotes, but that form "
                      "is now deprecated and will be removed in a future version of Terraform. Remove the quotes "
                      f"around \"{word}\"{better[word]}.", f.path, expr, "type")
        else:
            self.diag(WARNING, "Quoted type constraints are deprecated",
                      "Terraform 0.11 and earlier required type constraints to be given in quotes, but that form "
                      "is now deprecated and will be removed in a future version of Terraform. To silence this "
                      f"warning, remove the quotes around \"{word}\"{better[word]}.", f.path, expr, "deprecated")
        return typeexpr.Type(word) if word == "string" else typeexpr.Type(word, elem=typeexpr.Type("string"))

    def block_locals(self, f: SourceFile, b: A.Block) -> None:
        for item in b.body.items:
            if isinstance(item, A.Block):
                self.diag(ERROR, "Unsupported block type",
                          f"Blocks of type \"{item.type}\" are not expected here.", f.path, item, "structure")
                continue
            if not is_valid_identifier(item.name):
                self.diag(ERROR, "Invalid local value name", _BAD_NAME, f.path, item, "structure")
                continue
            if self.duplicate(f, item, self.m.locals, item.name, "Duplicate local value definition",
                              f"A local value named \"{item.name}\" was already defined at {{prev}}. "
                              "Local value names must be unique within a module."):
                continue
            self.m.locals[item.name] = Decl(item.name, f.path, item.pos.line, attr=item)

    def block_output(self, f: SourceFile, b: A.Block) -> None:
        name = b.labels[0].value
        if not self.check_name(f, b, name, "output name"):
            return
        self.check_args(f, b, OUTPUT_ARGS, {"precondition": 0}, gates="output")
        for pb in b.body.blocks_of("precondition"):
            self.condition_block(f, pb)
        attrs = b.body.attributes
        if "value" not in attrs:
            self.diag(ERROR, "Missing required argument",
                      "The argument \"value\" is required, but no definition was found.", f.path, b, "structure")
        if self.duplicate(f, b, self.m.outputs, name, "Duplicate output definition",
                          f"An output named \"{name}\" was already defined at {{prev}}. "
                          "Output names must be unique within a module."):
            return
        o = Output(name, f.path, b.pos.line, block=b)
        if "description" in attrs:
            o.description = self.literal(f, attrs["description"], str, "description")
        if "sensitive" in attrs:
            o.sensitive = bool(self.literal(f, attrs["sensitive"], bool, "sensitive flag"))
        self.m.outputs[name] = o

    def _count_for_each(self, f: SourceFile, b: A.Block) -> tuple[bool, bool]:
        attrs = b.body.attributes
        has_count, has_each = "count" in attrs, "for_each" in attrs
        if has_count:
            v = typeexpr.literal_value(attrs["count"].expr)
            ok = v is typeexpr.UNKNOWN or (isinstance(v, (int, float)) and not isinstance(v, bool)) \
                or (isinstance(v, str) and v.isdigit())
            if v is None:
                self.diag(ERROR, "Invalid count argument",
                          "The given \"count\" argument value is null. An integer is required.",
                          f.path, attrs["count"].expr, "type")
            elif not ok:
                self.diag(ERROR, "Incorrect value type", "Invalid expression value: number required.",
                          f.path, attrs["count"].expr, "type")
            elif isinstance(v, (int, float)) and (v < 0 or v != int(v)):
                self.diag(ERROR, "Invalid count argument",
                          "The given \"count\" argument value is unsuitable: must be a whole number that is "
                          "greater than or equal to zero.", f.path, attrs["count"].expr, "type")
        if has_each:
            e = attrs["for_each"].expr
            if isinstance(e, A.TupleCons) or (isinstance(e, A.ForExpr) and e.key_expr is None):
                self.diag(ERROR, "Invalid for_each argument",
                          "The given \"for_each\" argument value is unsuitable: the \"for_each\" argument must be "
                          "a map, or set of strings, and you have provided a value of type tuple. Wrap it in "
                          "toset(...) to use its elements as keys.", f.path, e, "type")
            elif isinstance(e, A.Literal) or (isinstance(e, A.Template) and e.literal_value() is not None):
                self.diag(ERROR, "Invalid for_each argument",
                          "The given \"for_each\" argument value is unsuitable: the \"for_each\" argument must be "
                          "a map, or set of strings.", f.path, e, "type")
        if has_count and has_each:
            self.diag(ERROR, "Invalid combination of \"count\" and \"for_each\"",
                      "The \"count\" and \"for_each\" meta-arguments are mutually-exclusive, only one should "
                      "be used to be explicit about how many instances are to be created.",
                      f.path, attrs["for_each"], "structure")
        return has_count, has_each

    def block_module(self, f: SourceFile, b: A.Block) -> None:
        name = b.labels[0].value
        if not self.check_name(f, b, name, "module instance name"):
            return
        if self.duplicate(f, b, self.m.module_calls, name, "Duplicate module call",
                          f"A module call named \"{name}\" was already defined at {{prev}}. "
                          "Module calls must have unique names within a module."):
            return
        mc = ModuleCall(name, f.path, b.pos.line, block=b)
        attrs = b.body.attributes
        mc.has_count, mc.has_for_each = self._count_for_each(f, b)
        if "source" not in attrs:
            self.diag(ERROR, "Missing required argument",
                      "The argument \"source\" is required, but no definition was found.", f.path, b, "structure")
        else:
            mc.source = self.literal(f, attrs["source"], str, "module source")
        if "version" in attrs:
            mc.version = self.literal(f, attrs["version"], str, "version constraint")
            if mc.version is not None:
                cons, err = try_parse_constraints(mc.version)
                if cons is None:
                    self.diag(ERROR, "Invalid version constraint",
                              f"This string does not use correct version constraint syntax: {err}.",
                              f.path, attrs["version"].expr, "version")
                if mc.source is not None and not mc.is_registry:
                    self.diag(ERROR, "Invalid registry module source address",
                              f"Failed to parse module registry address: \"{mc.source}\" is not a module registry "
                              "address.\n\nTerraform assumed that you intended a module registry source address "
                              "because you also set the argument \"version\", which applies only to registry modules.",
                              f.path, attrs["source"].expr, "version")
        for blk in b.body.blocks:
            if blk.type in ("lifecycle", "locals", "provider", "dynamic"):
                self.diag(ERROR, "Reserved block type name in module block" if blk.type != "dynamic"
                          else "Unsupported block type",
                          f"The block type name \"{blk.type}\" is reserved for use by Terraform in a future version."
                          if blk.type != "dynamic" else
                          f"Blocks of type \"{blk.type}\" are not expected here.", f.path, blk, "structure")
            else:
                self.diag(ERROR, "Unsupported block type",
                          f"Blocks of type \"{blk.type}\" are not expected here.", f.path, blk, "structure")
        self.m.module_calls[name] = mc

    def _resource(self, f: SourceFile, b: A.Block, mode: str, scope: Optional[str] = None) -> None:
        typ, name = b.labels[0].value, b.labels[1].value
        word = {"data": "data source", "ephemeral": "ephemeral resource"}.get(mode, "resource")
        if not self.check_name(f, b, typ, f"{word} type name" if mode != "data" else "data source name"):
            return
        if not self.check_name(f, b, name, f"{word} name"):
            return
        r = Resource(name, f.path, b.pos.line, block=b, mode=mode, type=typ, scope=scope)
        kind = {"data": "data", "ephemeral": "ephemeral"}.get(mode, "resource")
        if self.duplicate(f, b, self.m.resources, r.addr,
                          f"Duplicate {kind} \"{typ}\" configuration",
                          f"A {typ} {word} named \"{name}\" was already declared at {{prev}}. "
                          f"Resource names must be unique per type in each module."):
            return
        attrs = b.body.attributes
        if scope is not None:
            for meta in ("count", "for_each"):
                if meta in attrs:
                    self.diag(ERROR, f"Invalid \"{meta}\" attribute",
                              "The \"count\" and \"for_each\" meta-arguments are not supported within nested "
                              "data blocks.", f.path, attrs[meta], "structure")
            for blk in b.body.blocks_of("lifecycle"):
                self.diag(ERROR, "Invalid lifecycle block",
                          "Nested data blocks do not support \"lifecycle\" blocks as the lifecycle is managed "
                          "by the containing block.", f.path, blk, "structure")
        else:
            r.has_count, r.has_for_each = self._count_for_each(f, b)
        if "provider" in attrs:
            e = attrs["provider"].expr
            if isinstance(e, A.ScopeTraversal) and len(e.steps) <= 1 and all(s.kind == "attr" for s in e.steps):
                r.provider_ref = e.dotted()
            elif isinstance(e, A.Template) and e.literal_value() is not None:
                r.provider_ref = e.literal_value()
                self.diag(WARNING, "Quoted references are deprecated",
                          "In this context, references are expected literally rather than in quotes. "
                          "Terraform 0.11 and earlier required quotes, but quoted references are now "
                          "deprecated and will be removed in a future version of Terraform. Remove the "
                          "quotes surrounding this reference to silence this warning.", f.path, e, "deprecated")
            else:
                self.diag(ERROR, "Invalid provider configuration reference",
                          "The provider argument requires a provider type name, optionally followed by a "
                          "period and then a configuration alias.", f.path, e, "structure")
        seen_lifecycle = False
        for blk in b.body.blocks:
            if blk.type == "lifecycle":
                if seen_lifecycle:
                    self.diag(ERROR, "Duplicate lifecycle block",
                              "This resource already has a lifecycle block.", f.path, blk, "structure")
                seen_lifecycle = True
                self.check_args(f, blk, LIFECYCLE_ARGS if mode == "managed" else set(),
                                {"precondition": 0, "postcondition": 0}, gates="lifecycle")
                for cb in blk.body.blocks:
                    if cb.type in ("precondition", "postcondition"):
                        self.condition_block(f, cb)
                ic = blk.body.attributes.get("ignore_changes")
                if ic is not None and not isinstance(ic.expr, A.TupleCons) and not (
                        isinstance(ic.expr, A.ScopeTraversal) and ic.expr.root == "all" and not ic.expr.steps):
                    # neither the keyword ``all`` nor a list: Terraform asks HCL for a static list
                    self.diag(ERROR, "Invalid expression", "A static list expression is required.",
                              f.path, ic.expr, "structure")
            elif blk.type in ("provisioner", "connection") and mode != "managed":
                self.diag(ERROR, "Unsupported block type" if blk.type == "connection" or mode == "ephemeral"
                          else "Data resources do not have provisioners",
                          "Data resources do not have provisioners." if blk.type == "provisioner" and mode == "data"
                          else f"Blocks of type \"{blk.type}\" are not expected here.", f.path, blk, "structure")
            elif blk.type == "provisioner":
                if len(blk.labels) != 1:
                    self.diag(ERROR, "Missing type for provisioner" if not blk.labels
                              else "Extraneous label for provisioner",
                              "All provisioner blocks must have 1 labels (type).", f.path, blk, "structure")
            elif blk.type == "dynamic":
                self.check_dynamic(f, blk)
        self._walk_dynamic(f, b.body)
        self.m.resources[r.addr] = r

    def _walk_dynamic(self, f: SourceFile, body: A.Body) -> None:
        for blk in body.blocks:
            if blk.type == "dynamic":
                self.check_dynamic(f, blk)
            self._walk_dynamic(f, blk.body)

    def check_dynamic(self, f: SourceFile, blk: A.Block) -> None:
        if getattr(blk, "_dyn_checked", False):
            return
        blk._dyn_checked = True
        if len(blk.labels) != 1:
            self.diag(ERROR, "Invalid dynamic block", "A dynamic block must have exactly one label: "
                      "the type of block to generate.", f.path, blk, "structure")
        attrs = blk.body.attributes
        if "for_each" not in attrs:
            self.diag(ERROR, "Missing required argument",
                      "The argument \"for_each\" is required, but no definition was found.", f.path, blk, "structure")
        for a in attrs:
            if a not in ("for_each", "iterator", "labels"):
                self.diag(ERROR, "Unsupported argument",
                          f"An argument named \"{a}\" is not expected here.", f.path, attrs[a], "structure")
        content = blk.body.blocks_of("content")
        if not content:
            self.diag(ERROR, "Missing dynamic content block",
                      "A dynamic block must have a nested block of type \"content\" to describe the body of each "
                      "generated block.", f.path, blk, "structure")
        elif len(content) > 1:
            self.diag(ERROR, "Multiple content blocks",
                      "Only one nested content block is allowed for each dynamic block.", f.path, content[1], "structure")
        for other in blk.body.blocks:
            if other.type != "content":
                self.diag(ERROR, "Unsupported block type",
                          f"Blocks of type \"{other.type}\" are not expected here.", f.path, other, "structure")

    def block_resource(self, f: SourceFile, b: A.Block) -> None:
        self._resource(f, b, "managed")

    def block_data(self, f: SourceFile, b: A.Block) -> None:
        self._resource(f, b, "data")

    def block_ephemeral(self, f: SourceFile, b: A.Block) -> None:
        self._resource(f, b, "ephemeral")

    def block_provider(self, f: SourceFile, b: A.Block) -> None:
        name = b.labels[0].value
        if not self.check_name(f, b, name, "provider configuration name"):
            return
        p = ProviderConfig(name, f.path, b.pos.line, block=b)
        attrs = b.body.attributes
        if "alias" in attrs:
            p.alias = self.literal(f, attrs["alias"], str, "provider configuration alias")
            if p.alias is not None and not is_valid_identifier(p.alias):
                self.diag(ERROR, "Invalid provider configuration alias",
                          f"An alias must be a valid name. {_BAD_NAME}", f.path, attrs["alias"].expr, "structure")
        if "version" in attrs:
            self.diag(WARNING, "Version constraints inside provider configuration blocks are deprecated",
                      "Terraform 0.13 and earlier allowed provider version constraints inside the provider "
                      "configuration block, but that is now deprecated and will be removed in a future version "
                      "of Terraform. To silence this warning, move the provider version constraint into the "
                      "required_providers block.", f.path, attrs["version"], "deprecated")
        for prev in self.m.providers:
            if prev.name == name and prev.alias == p.alias:
                self.diag(ERROR, "Duplicate provider configuration",
                          f"A {'default (non-aliased)' if p.alias is None else 'aliased'} provider configuration "
                          f"for \"{name}\" was already given at {prev.file}:{prev.line}. If multiple configurations "
                          "are required, set the \"alias\" argument for alternative configurations.",
                          f.path, b, "duplicate")
                return
        self.m.providers.append(p)

    def block_terraform(self, f: SourceFile, b: A.Block) -> None:
        self.check_args(f, b, {"required_version", "experiments", "language"},
                        {"required_providers": 0, "backend": 1, "cloud": 0, "provider_meta": 1},
                        gates="terraform")
        self.experiments(f, b)
        attrs = b.body.attributes
        if "required_version" in attrs:
            text = self.literal(f, attrs["required_version"], str, "version constraint")
            if text is not None:
                cons, err = try_parse_constraints(text)
                if cons is None:
                    self.diag(ERROR, "Invalid version constraint",
                              f"This string does not use correct version constraint syntax: {err}.",
                              f.path, attrs["required_version"].expr, "version")
                else:
                    self.m.required_versions.append(
                        (text, Decl("required_version", f.path, attrs["required_version"].pos.line,
                                    attr=attrs["required_version"])))
        backends = b.body.blocks_of("backend")
        if len(backends) > 1:
            self.diag(ERROR, "Duplicate backend configuration",
                      f"A module may have only one backend configuration. The backend was previously configured at "
                      f"{f.path}:{backends[0].pos.line}.", f.path, backends[1], "duplicate")
        for be in backends:
            for node in be.body.walk():
                if isinstance(node, A.ScopeTraversal):
                    self.diag(ERROR, "Variables not allowed", "Variables may not be used here.",
                              f.path, node, "structure")
                elif isinstance(node, A.FunctionCall):
                    self.diag(ERROR, "Function calls not allowed", "Functions may not be called here.",
                              f.path, node, "structure")
        rp_blocks = b.body.blocks_of("required_providers")
        for rp in rp_blocks:
            self.required_providers(f, rp)

    def experiments(self, f: SourceFile, b: A.Block) -> None:
        """``experiments = [name, ...]``.  Only the one experiment that changes what
        this checker accepts is judged: optional object attributes, an experiment from
        0.14 until 1.3 made the (changed) feature part of the language."""
        for it in self._experiment_keywords(b):
            if it.root != "module_variable_optional_attrs":
                continue
            if self.v.has("optional-attributes"):
                self.diag(ERROR, "Experiment has concluded",
                          "Experiment \"module_variable_optional_attrs\" is no longer available. The final "
                          "feature corresponding to this experiment differs from the experimental form and is "
                          "available in the Terraform language from Terraform v1.3.0 onwards.",
                          f.path, it, "terraform-version")
            else:
                self.diag(WARNING, "Experimental feature \"module_variable_optional_attrs\" is active",
                          "Experimental features are subject to breaking changes in future minor or patch "
                          "releases, based on feedback.\n\nIf you have feedback on the design of this feature, "
                          "please open a GitHub issue to discuss it.", f.path, it, "terraform-version")

    @staticmethod
    def _experiment_keywords(b: A.Block) -> list[A.ScopeTraversal]:
        attr = b.body.attributes.get("experiments")
        if attr is None or not isinstance(attr.expr, A.TupleCons):
            return []
        return [it for it in attr.expr.items if isinstance(it, A.ScopeTraversal) and not it.steps]

    def optional_attributes(self, f: SourceFile, b: A.Block, type_expr: A.Expr) -> None:
        """Before 1.3 ``optional(...)`` needed the experiment, and took no default."""
        if self.v.has("optional-attributes"):
            return
        calls = [n for n in type_expr.walk() if isinstance(n, A.FunctionCall) and n.name == "optional"]
        if not calls:
            return
        if "module_variable_optional_attrs" not in self.m.experiments:
            self.diag(ERROR, "Optional object type attributes are experimental",
                      "This feature is currently an opt-in experiment, subject to change in future releases "
                      "based on feedback.\n\nActivate the feature for this module by adding "
                      "module_variable_optional_attrs to the list of active experiments."
                      + self.needs("optional-attributes", "optional() without the experiment"),
                      f.path, b, "terraform-version")
            return
        for c in calls:
            if len(c.args) > 1:
                self.diag(ERROR, "Invalid type specification",
                          "Optional attribute modifier expects only one argument: the attribute type."
                          + self.needs("optional-attributes", "A default value inside optional()"),
                          f.path, c, "terraform-version")

    def required_providers(self, f: SourceFile, rp: A.Block) -> None:
        for blk in rp.body.blocks:
            self.diag(ERROR, "Unsupported block type",
                      f"Blocks of type \"{blk.type}\" are not expected here.", f.path, blk, "structure")
        for item in rp.body.items:
            if not isinstance(item, A.Attribute):
                continue
            name = item.name
            if not is_valid_identifier(name):
                self.diag(ERROR, "Invalid provider local name", _BAD_NAME, f.path, item, "structure")
                continue
            if name in self.m.required_providers:
                prev = self.m.required_providers[name]
                self.diag(ERROR, "Duplicate required providers configuration",
                          f"A module may have only one required providers configuration. The required providers "
                          f"were previously configured at {prev.file}:{prev.line}.", f.path, item, "duplicate")
                continue
            req = RequiredProvider(name, f.path, item.pos.line, attr=item)
            val = typeexpr.literal_value(item.expr)
            if isinstance(val, str):
                req.constraint_text = val              # legacy: aws = "~> 4.0"
            elif isinstance(val, dict):
                aliases_known = self.v.has("required_providers:configuration_aliases")
                for k in val:
                    if k == "configuration_aliases" and not aliases_known:
                        self.diag(ERROR, "Invalid required_providers object",
                                  "required_providers objects can only contain \"version\" and \"source\" "
                                  "attributes. To configure a provider, use a \"provider\" block."
                                  + self.needs("required_providers:configuration_aliases",
                                               "The \"configuration_aliases\" attribute"),
                                  f.path, item.expr, "terraform-version")
                    elif k not in ("source", "version", "configuration_aliases"):
                        self.diag(ERROR, "Invalid required_providers object",
                                  "required_providers objects can only contain \"version\", \"source\" and "
                                  "\"configuration_aliases\" attributes. To configure a provider, use a "
                                  "\"provider\" block.", f.path, item.expr, "structure")
                if isinstance(val.get("source"), str):
                    req.source = val["source"]
                    parts = req.source.split("/")
                    if not 1 <= len(parts) <= 3 or any(not p for p in parts):
                        self.diag(ERROR, "Invalid provider source string",
                                  "The \"source\" attribute must be in the format \"[hostname/][namespace/]name\"",
                                  f.path, item.expr, "structure")
                elif "source" in val and val["source"] is not typeexpr.UNKNOWN:
                    self.diag(ERROR, "Invalid source", "Source must be specified as a string.",
                              f.path, item.expr, "structure")
                if isinstance(val.get("version"), str):
                    req.constraint_text = val["version"]
                wanted = item.expr.get("configuration_aliases") if isinstance(item.expr, A.ObjectCons) else None
                if isinstance(wanted, A.TupleCons):
                    req.aliases = tuple(t.steps[0].value for t in wanted.items
                                        if isinstance(t, A.ScopeTraversal) and t.root == name
                                        and len(t.steps) == 1 and t.steps[0].kind == "attr")
            else:
                self.diag(ERROR, "Invalid required_providers object",
                          "required_providers entries must be strings or objects.", f.path, item.expr, "structure")
                continue
            if req.constraint_text is not None:
                req.constraints, err = try_parse_constraints(req.constraint_text)
                if req.constraints is None:
                    self.diag(ERROR, "Invalid version constraint",
                              f"This string does not use correct version constraint syntax: {err}.",
                              f.path, item.expr, "version")
            self.m.required_providers[name] = req


_MODULE_CACHE: dict[tuple, Module] = {}


def _signature(directory: str) -> tuple:
    try:
        names = sorted(n for n in os.listdir(directory) if n.endswith(".tf"))
    except OSError:
        return ()
    sig = []
    for n in names:
        try:
            st = os.stat(os.path.join(directory, n))
        except OSError:                      # dangling symlink: reported when the file is read
            sig.append((n, 0, -1))
            continue
        sig.append((n, st.st_mtime_ns, st.st_size))
    return tuple(sig)


def load_module(directory: str, display_root: Optional[str] = None,
                restated: Optional[Restated] = None) -> Module:
    """Parse every ``*.tf`` file of ``directory`` and build its Module.  The
    result is shared (and must be treated as read-only) until a file changes.

    ``restated`` fixes the Terraform release whose rules apply (a child module is
    read under its caller's); without it the module's own ``required_version``
    decides -- see ``tfversion``."""
    key = (os.path.abspath(directory), display_root, restated, _signature(directory))
    m = _MODULE_CACHE.get(key)
    if m is None:
        if len(_MODULE_CACHE) > 1024:
            _MODULE_CACHE.clear()
        m = _MODULE_CACHE[key] = _Loader(directory, display_root, restated).load()
    return m
