"""``validate``: the schema-free part of ``terraform validate``.

BASELINE.json config 2 is "``terraform validate`` on eks/"; the real command needs
the Terraform binary, six provider plugins and three registry modules, none of
which exist offline.  What CAN be decided from the ``.tf`` text alone
is restated here from Terraform's documented behaviour:

* every reference resolves -- ``var.*``, ``local.*``, ``module.*``,
  ``data.<type>.<name>``, ``<type>.<name>``, and the contextual objects
  ``count``/``each``/``self``/``path``/``terraform`` are used where they exist
  (the reference leans on these everywhere, e.g. ``random_string.amp[count.index]``
  at eks/examples/cnpack/aws-amp.tf:19);
* instance keys agree with ``count``/``for_each``
  (``azurerm_resource_group.holoscan[0]`` at aks/main.tf:22 needs the ``count``
  at aks/main.tf:10);
* calls name real built-in functions with a possible number of arguments;
* a module call whose source is a local path (``source = "../.."``,
  eks/examples/cnpack/main.tf:36) passes only declared variables, passes all
  required ones, and reads only declared outputs;
* ``depends_on`` holds static references; deprecated 0.11 forms are warned about.

NOT covered, and never claimed: anything that needs a provider schema (argument
names and types inside resource/data/provider blocks) or a registry module's
interface.  The summary line this command prints says so every time.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field, replace
from typing import Optional

from ..hcl import ast as A
from ..hcl.diagnostics import Diagnostic, ERROR, WARNING, has_errors
from ..hcl.parser import parse_path
from . import functions, kinds, tfversion, typeexpr
from .model import Module, ModuleCall, Resource, SourceFile, load_module, resource_addr

_PATH_ATTRS = {"module", "root", "cwd"}
_TERRAFORM_ATTRS = {"workspace", "applying"}


@dataclass
class Ctx:
    file: str
    count: bool = False           # count.index is available
    each: bool = False            # each.key / each.value are available
    self_ok: bool = False         # "self" is available (provisioner, connection, postcondition)
    names: frozenset = frozenset()    # for-expression / dynamic-block iterator names in scope
    where: str = "module"         # wording: which kind of block we are in
    check: Optional[str] = None   # inside this check block: its nested data source is visible
    destroy: bool = False         # inside a destroy-time provisioner

    def with_names(self, *names: Optional[str]) -> "Ctx":
        return replace(self, names=self.names | {n for n in names if n})

    def but(self, **kw) -> "Ctx":
        return replace(self, **kw)


@dataclass
class Validator:
    m: Module
    children: dict[str, Optional[Module]] = field(default_factory=dict)
    diags: list[Diagnostic] = field(default_factory=list)
    used_vars: set = field(default_factory=set)
    used_locals: set = field(default_factory=set)
    used_resources: set = field(default_factory=set)
    used_modules: set = field(default_factory=set)
    edges: dict = field(default_factory=dict)       # who refers to whom, for cycle detection
    _owner: Optional[str] = None

    # -- plumbing ------------------------------------------------------------
    def diag(self, sev: str, summary: str, detail: str, ctx: Ctx, node, rule: str) -> None:
        if isinstance(node, A.Node):
            line, col = node.pos.line, node.pos.col
        elif isinstance(node, A.Step):
            line, col = node.line, node.col
        else:
            line, col = int(node), 0
        self.diags.append(Diagnostic(sev, summary, detail, ctx.file, line, col, rule))

    def child(self, mc: ModuleCall) -> Optional[Module]:
        if not mc.is_local:
            return None
        if mc.name not in self.children:
            path = os.path.normpath(os.path.join(self.m.dir, mc.source))
            self.children[mc.name] = (load_module(path, self.m.display_root, self.m.restated)
                                      if os.path.isdir(path) else None)
        return self.children[mc.name]

    # -- entry ---------------------------------------------------------------
    def run(self) -> list[Diagnostic]:
        for f in self.m.files:
            if f.path in self.m.broken_files:
                continue
            for item in f.body.items:
                if isinstance(item, A.Block):
                    handler = getattr(self, "visit_" + item.type, None)
                    if handler:
                        handler(f, item)
        self.cycles()
        return self.diags

    def cycles(self) -> None:
        """Dependency cycles among local values, resources and data sources
        (module calls are left out: Terraform orders them per input/output, and
        a registry module's internals are not visible offline)."""
        WHITE, GREY, BLACK = 0, 1, 2
        colour: dict[str, int] = {}
        reported: set[frozenset] = set()

        def where(addr: str) -> tuple[str, int]:
            if addr.startswith("local."):
                d = self.m.locals[addr[6:]]
            else:
                d = self.m.resources[addr]
            return d.file, d.line

        def visit(node: str, path: list[str]) -> None:
            colour[node] = GREY
            path.append(node)
            for nxt in sorted(self.edges.get(node, ())):
                if nxt == node:
                    continue                      # reported as self-reference
                c = colour.get(nxt, WHITE)
                if c == GREY:
                    cyc = path[path.index(nxt):]
                    key = frozenset(cyc)
                    if key not in reported:
                        reported.add(key)
                        file, line = where(cyc[0])
                        self.diags.append(Diagnostic(ERROR, "Cycle: " + ", ".join(sorted(cyc)),
                                                     "These objects refer to each other in a circle, so no order "
                                                     "exists in which Terraform could evaluate them.",
                                                     file, line, 0, "cycle"))
                elif c == WHITE:
                    visit(nxt, path)
            path.pop()
            colour[node] = BLACK

        for node in sorted(self.edges):
            if colour.get(node, WHITE) == WHITE:
                visit(node, [])

    # -- blocks --------------------------------------------------------------
    def visit_variable(self, f: SourceFile, b: A.Block) -> None:
        ctx = Ctx(f.path, where="variable")
        for name, attr in b.body.attributes.items():
            if name == "type":
                continue            # a type expression, not a value expression
            self.calls_only(attr.expr, ctx)
        for blk in b.body.blocks_of("validation"):
            for attr in blk.body.attributes.values():
                self.expr(attr.expr, ctx)

    def visit_locals(self, f: SourceFile, b: A.Block) -> None:
        ctx = Ctx(f.path, where="locals")
        for item in b.body.items:
            if isinstance(item, A.Attribute):
                self._owner = f"local.{item.name}"
                try:
                    self.attr_value(item, ctx)
                finally:
                    self._owner = None

    def visit_output(self, f: SourceFile, b: A.Block) -> None:
        ctx = Ctx(f.path, where="output")
        for name, attr in b.body.attributes.items():
            if name == "depends_on":
                self.depends_on(attr, ctx)
            else:
                self.attr_value(attr, ctx)
        for blk in b.body.blocks_of("precondition"):
            self.body(blk.body, ctx)

    def visit_provider(self, f: SourceFile, b: A.Block) -> None:
        ctx = Ctx(f.path, where="provider")
        self.body(b.body, ctx, skip={"alias", "version"})

    def visit_check(self, f: SourceFile, b: A.Block) -> None:
        name = b.labels[0].value if b.labels else None
        ctx = Ctx(f.path, where="check", check=name)
        for blk in b.body.blocks:
            if blk.type == "data":
                self._resource(f, blk, "data", ctx)
            else:
                self.body(blk.body, ctx)

    def visit_module(self, f: SourceFile, b: A.Block) -> None:
        mc = self.m.module_calls.get(b.labels[0].value) if b.labels else None
        if mc is None or mc.block is not b:
            return
        base = Ctx(f.path, where="module")
        ctx = base.but(count=mc.has_count, each=mc.has_for_each)
        child = self.child(mc)
        if mc.is_local and child is None:
            self.diag(ERROR, "Unreadable module directory",
                      f"Unable to evaluate directory symlink or the directory \"{mc.source}\" does not exist.",
                      base, b, "module-source")
        if child is not None and child.providers:
            hard = [a for a in ("count", "for_each", "depends_on") if a in b.body.attributes]
            if hard:
                p = child.providers[0]
                self.diag(ERROR, "Module is incompatible with count, for_each, and depends_on",
                          f"The module at module.{mc.name} is a legacy module which contains its own local provider "
                          f"configurations, and so calls to it may not use the {', '.join(hard)} argument"
                          f"{'s' if len(hard) > 1 else ''}. (provider \"{p.name}\" is configured at "
                          f"{p.file}:{p.line}.)", base, b.body.attributes[hard[0]], "legacy-module")
        passed: set[str] = set()
        for name, attr in b.body.attributes.items():
            if name in ("source", "version"):
                continue
            if name in ("count", "for_each"):
                self.attr_value(attr, base)      # cannot refer to its own count/each
                continue
            if name == "depends_on":
                self.depends_on(attr, base)
                continue
            if name == "providers":
                continue                          # provider references, not values
            passed.add(name)
            self.attr_value(attr, ctx)
            if child is not None and name not in child.variables:
                self.diag(ERROR, "Unsupported argument",
                          f"An argument named \"{name}\" is not expected here.", base, attr, "module-arguments")
        if child is not None:
            handed = set()
            pattr = b.body.attributes.get("providers")
            if pattr is not None and isinstance(pattr.expr, A.ObjectCons):
                for key, _ in pattr.expr.items:
                    w = key.wrapped
                    handed.add(key.bare_name or (w.dotted() if isinstance(w, A.ScopeTraversal) else ""))
            for req in child.required_providers.values():
                for alias in req.aliases:
                    local = f"{req.name}.{alias}"
                    if local not in handed:
                        self.diag(ERROR, "Missing required provider configuration",
                                  f"The child module requires an additional configuration for provider {req.display}, "
                                  f"with the local name \"{local}\".\n\nRefer to the module's documentation to "
                                  "understand the intended purpose of this additional provider configuration, and then "
                                  f"add an entry for {local} in the \"providers\" meta-argument in the module block to "
                                  "choose which provider configuration the module should use for that purpose.",
                                  base, b, "module-arguments")
            for vname, var in child.variables.items():
                if var.required and vname not in passed:
                    self.diag(ERROR, "Missing required argument",
                              f"The argument \"{vname}\" is required, but no definition was found.",
                              base, b, "module-arguments")

    def visit_resource(self, f: SourceFile, b: A.Block) -> None:
        self._resource(f, b, "managed")

    def visit_data(self, f: SourceFile, b: A.Block) -> None:
        self._resource(f, b, "data")

    def visit_ephemeral(self, f: SourceFile, b: A.Block) -> None:
        self._resource(f, b, "ephemeral")

    def _resource(self, f: SourceFile, b: A.Block, mode: str, outer: Optional[Ctx] = None) -> None:
        if len(b.labels) != 2:
            return
        addr = resource_addr(mode, b.labels[0].value, b.labels[1].value)
        r = self.m.resources.get(addr)
        if r is None or r.block is not b:
            return
        base = (outer or Ctx(f.path)).but(where="resource" if mode == "managed" else mode)
        ctx = base.but(count=r.has_count, each=r.has_for_each)
        self._owner = addr
        try:
            self._resource_body(b, r, base, ctx)
        finally:
            self._owner = None

    def _resource_body(self, b: A.Block, r: Resource, base: Ctx, ctx: Ctx) -> None:
        for name, attr in b.body.attributes.items():
            if name in ("count", "for_each"):
                self.attr_value(attr, base)
            elif name == "depends_on":
                self.depends_on(attr, base)
            elif name == "provider":
                continue
            else:
                self.attr_value(attr, ctx)
        for blk in b.body.blocks:
            if blk.type == "lifecycle":
                self.lifecycle(blk, ctx)
            elif blk.type == "provisioner":
                self.provisioner(blk, ctx)
            elif blk.type == "connection":
                # a resource-level connection is shared by every provisioner; it is held to
                # the destroy-time rule as soon as one of them runs at destroy time
                destroy = any(self._when(p) == "destroy" for p in b.body.blocks_of("provisioner"))
                self.body(blk.body, ctx.but(self_ok=True, where="provisioner", destroy=destroy))
            else:
                self.nested(blk, ctx)

    _KEYWORDS = {"when": ("create", "destroy"), "on_failure": ("continue", "fail")}

    @staticmethod
    def _keyword(e: A.Expr) -> Optional[str]:
        """``destroy`` or the deprecated ``"destroy"``: the bare word, else None."""
        if isinstance(e, A.ScopeTraversal) and not e.steps:
            return e.root
        if isinstance(e, A.Template) and not e.heredoc:
            return e.literal_value()
        return None

    def _when(self, blk: A.Block) -> Optional[str]:
        attr = blk.body.attributes.get("when")
        return self._keyword(attr.expr) if attr is not None else None

    def provisioner(self, blk: A.Block, ctx: Ctx) -> None:
        """``when`` and ``on_failure`` take bare keywords, not expressions; everything else
        goes to the provisioner plugin and may use ``self``."""
        inner = ctx.but(self_ok=True, where="provisioner", destroy=self._when(blk) == "destroy")
        for item in blk.body.items:
            if isinstance(item, A.Block):
                if item.type == "connection":
                    self.body(item.body, inner)
                else:
                    self.nested(item, inner)
                continue
            allowed = self._KEYWORDS.get(item.name)
            if allowed is None:
                self.attr_value(item, inner)
                continue
            if isinstance(item.expr, A.Template) and self._keyword(item.expr) is not None:
                self.diag(WARNING, "Quoted keywords are deprecated",
                          "In this context, keywords are expected literally rather than in quotes. Terraform 0.11 "
                          "and earlier required quotes, but quoted keywords are now deprecated and will be removed "
                          "in a future version of Terraform. Remove the quotes surrounding this keyword to silence "
                          "this warning.", ctx, item.expr, "deprecated")
            if self._keyword(item.expr) not in allowed:
                self.diag(ERROR, f"Invalid \"{item.name}\" keyword",
                          f"The \"{item.name}\" argument requires one of the following keywords: "
                          f"{allowed[0]} or {allowed[1]}.", ctx, item.expr, "structure")

    def lifecycle(self, blk: A.Block, ctx: Ctx) -> None:
        for name, attr in blk.body.attributes.items():
            if name == "ignore_changes":
                e = attr.expr
                if isinstance(e, A.TupleCons):
                    for it in e.items:
                        if isinstance(it, A.Template):
                            self.diag(WARNING, "Quoted references are deprecated",
                                      "In this context, references are expected literally rather than in quotes. "
                                      "Remove the quotes surrounding this reference to silence this warning.",
                                      ctx, it, "deprecated")
                continue                      # attribute names of the resource itself
            if name == "replace_triggered_by":
                if isinstance(attr.expr, A.TupleCons):
                    for it in attr.expr.items:
                        self.expr(it, ctx)
                continue
            self.calls_only(attr.expr, ctx)
        for cond in blk.body.blocks:
            self.body(cond.body, ctx.but(self_ok=cond.type == "postcondition"))

    def nested(self, blk: A.Block, ctx: Ctx) -> None:
        """A nested configuration block, possibly ``dynamic``."""
        if blk.type == "dynamic" and blk.labels:
            attrs = blk.body.attributes
            iterator = blk.labels[0].value
            it = attrs.get("iterator")
            if it is not None:
                if isinstance(it.expr, A.ScopeTraversal) and not it.expr.steps:
                    iterator = it.expr.root
                else:
                    self.diag(ERROR, "Invalid dynamic iterator name", "Dynamic iterator must be a single variable name.",
                              ctx, it.expr, "structure")
            if "for_each" in attrs:
                self.attr_value(attrs["for_each"], ctx)
            inner = ctx.with_names(iterator)
            if "labels" in attrs:
                self.attr_value(attrs["labels"], inner)
            for content in blk.body.blocks_of("content"):
                self.body(content.body, inner)
            return
        self.body(blk.body, ctx)

    def body(self, body: A.Body, ctx: Ctx, skip: set = frozenset()) -> None:
        for item in body.items:
            if isinstance(item, A.Attribute):
                if item.name not in skip:
                    self.attr_value(item, ctx)
            else:
                self.nested(item, ctx)

    # -- expressions -----------------------------------------------------------
    def attr_value(self, attr: A.Attribute, ctx: Ctx) -> None:
        self.expr(attr.expr, ctx)

    def calls_only(self, e: A.Expr, ctx: Ctx) -> None:
        for node in e.walk():
            if isinstance(node, A.FunctionCall):
                self.call(node, ctx)

    def depends_on(self, attr: A.Attribute, ctx: Ctx) -> None:
        e = attr.expr
        if not isinstance(e, A.TupleCons):
            self.diag(ERROR, "Invalid expression",
                      "A static list expression is required.", ctx, e, "depends-on")
            return
        for it in e.items:
            if isinstance(it, A.Template) and it.literal_value() is not None:
                self.diag(WARNING, "Quoted references are deprecated",
                          "In this context, references are expected literally rather than in quotes. "
                          "Terraform 0.11 and earlier required quotes, but quoted references are now "
                          "deprecated and will be removed in a future version of Terraform. Remove the "
                          "quotes surrounding this reference to silence this warning.", ctx, it, "deprecated")
                continue
            if not isinstance(it, A.ScopeTraversal):
                self.diag(ERROR, "Invalid expression",
                          "A single static variable reference is required: only attribute access and "
                          "indexing with constant keys. No calculations, function calls, template "
                          "expressions, etc are allowed here.", ctx, it, "depends-on")
                continue
            self.reference(it, ctx, keyed=self._has_key(it))

    def expr(self, e: A.Node, ctx: Ctx, keyed: bool = False) -> None:
        """Check ``e`` and everything below it.  ``keyed`` tells a resource
        reference that the enclosing expression indexes it."""
        if isinstance(e, A.ScopeTraversal):
            self.reference(e, ctx, keyed or self._has_key(e))
        elif isinstance(e, A.Index):
            self.expr(e.coll, ctx, keyed=True)
            self.expr(e.key, ctx)
        elif isinstance(e, A.Splat):
            self.expr(e.source, ctx, keyed=True)       # [*] / .* stand for every key
        elif isinstance(e, A.RelativeTraversal):
            self.expr(e.source, ctx, keyed=keyed)
        elif isinstance(e, A.FunctionCall):
            self.call(e, ctx)
            for a in e.args:
                self.expr(a, ctx)
        elif isinstance(e, A.ForExpr):
            self.expr(e.coll, ctx)
            inner = ctx.with_names(e.key_var, e.val_var)
            for part in (e.key_expr, e.val_expr, e.cond):
                if part is not None:
                    self.expr(part, inner)
        elif isinstance(e, A.TemplateFor):
            self.expr(e.coll, ctx)
            inner = ctx.with_names(e.key_var, e.val_var)
            for p in e.body:
                if isinstance(p, A.Node):
                    self.expr(p, inner)
        elif isinstance(e, A.Template):
            if e.is_single_interpolation():
                self.diag(WARNING, "Interpolation-only expressions are deprecated",
                          "Terraform 0.11 and earlier required all non-constant expressions to be provided "
                          "via interpolation syntax, but this pattern is now deprecated. To silence this "
                          "warning, remove the \"${ sequence from the start and the }\" sequence from the end "
                          "of this expression, leaving just the inner expression.", ctx, e, "deprecated")
            for p in e.parts:
                if isinstance(p, A.Node):
                    self.expr(p, ctx)
        elif isinstance(e, A.ObjectKey):
            if e.bare_name is None:
                w = e.wrapped
                if isinstance(w, A.ScopeTraversal) and w.steps and not e.forced_expr:
                    self.diag(ERROR, "Ambiguous attribute key",
                              "If this expression is intended to be a reference, wrap it in parentheses. If it's "
                              "instead intended as a literal name containing periods, wrap it in quotes to create "
                              "a string literal.", ctx, w, "syntax")
                else:
                    self.expr(w, ctx)
        else:
            if isinstance(e, A.Conditional):
                clash = kinds.conditional_mismatch(e)
                if clash:
                    self.diag(ERROR, "Inconsistent conditional result types",
                              "The true and false result expressions must have consistent types. The 'true' value "
                              f"is {clash[0]}, but the 'false' value is {clash[1]}.", ctx, e, "type")
            elif isinstance(e, (A.Binary, A.Unary)):
                bad = kinds.operand_problem(e)
                if bad:
                    side, operand, need = bad
                    self.diag(ERROR, "Invalid operand",
                              f"Unsuitable value for {side + ' ' if side else ''}operand: {need} required.",
                              ctx, operand, "type")
            for ch in e.children():
                self.expr(ch, ctx)

    @staticmethod
    def _has_key(t: A.ScopeTraversal) -> bool:
        """Does the traversal index the resource/module it names?"""
        n = 3 if t.root in ("data", "ephemeral") else 2
        return len(t.steps) >= n and t.steps[n - 1].kind == "index"

    def call(self, call: A.FunctionCall, ctx: Ctx) -> None:
        name = call.name
        v = self.m.restated
        if name.startswith("provider::"):
            if not v.has("provider-defined-functions"):
                # older HCL stops at the "::" with a syntax error of its own wording; what is
                # restated here is only that the call is refused
                self.diag(ERROR, "Provider-defined functions are not available",
                          f"The call to {name}() uses the provider:: namespace, which needs Terraform "
                          f"{tfversion.text(tfversion.SINCE['provider-defined-functions'])} or later; this run "
                          f"restates {v.describe()}.", ctx, call, "terraform-version")
            return
        bare = name[len("core::"):] if name.startswith("core::") else name
        if name.startswith("core::") and not v.has("provider-defined-functions"):
            self.diag(ERROR, "Call to unknown function", f"There is no function named \"{name}\".\n\n(The core:: "
                      f"namespace needs Terraform 1.8 or later; this run restates {v.describe()}.)",
                      ctx, call, "terraform-version")
            return
        if functions.lookup(name) is not None and not v.has_function(bare):
            self.diag(ERROR, "Call to unknown function",
                      f"There is no function named \"{name}\".\n\n(\"{bare}\" was added in Terraform "
                      f"{tfversion.text(tfversion.FUNCTION_SINCE[bare])}; this run restates {v.describe()}.)",
                      ctx, call, "terraform-version")
            return
        if bare in functions.REMOVED and v.release < tfversion.LEGACY_FUNCTIONS_REMOVED:
            self.diag(WARNING, f"The \"{bare}\" function is deprecated",
                      f"Terraform {v.text} still accepts {bare}(...), and Terraform 0.15 removed it: use "
                      f"{'tolist([...])' if bare == 'list' else 'tomap({...})'} instead.", ctx, call, "deprecated")
            return
        if functions.lookup(name) is None:
            if name in functions.REMOVED:
                self.diag(ERROR, "Error in function call",
                          f"Call to function \"{name}\" failed: {functions.REMOVED[name]}", ctx, call, "function")
            else:
                self.diag(ERROR, "Call to unknown function",
                          f"There is no function named \"{name}\".", ctx, call, "function")
            return
        err = functions.arity_error(name, len(call.args), call.expand_final)
        if err:
            self.diag(ERROR, err[0], err[1], ctx, call, "function")

    # -- references ----------------------------------------------------------
    def reference(self, t: A.ScopeTraversal, ctx: Ctx, keyed: bool) -> None:
        root, steps = t.root, t.steps
        if root in ctx.names:
            return

        def first_attr() -> Optional[str]:
            return steps[0].value if steps and steps[0].kind == "attr" else None

        if ctx.destroy and not (root in ("self", "path", "terraform")
                                or (root, first_attr(), len(steps)) in (("count", "index", 1), ("each", "key", 1))):
            self.diag(ERROR, "Invalid reference from destroy provisioner",
                      "Destroy-time provisioners and their connection configurations may only reference "
                      "attributes of the related resource, via 'self', 'count.index', or 'each.key'.\n\n"
                      "References to other resources during the destroy phase can cause dependency cycles "
                      "and interact poorly with create_before_destroy.", ctx, t, "destroy-provisioner")
            return

        if root == "var":
            name = first_attr()
            if name is None:
                self.diag(ERROR, "Invalid reference", "The \"var\" object cannot be accessed directly. "
                          "Instead, access one of its attributes.", ctx, t, "reference")
            elif name not in self.m.variables:
                self.diag(ERROR, "Reference to undeclared input variable",
                          f"An input variable with the name \"{name}\" has not been declared. This variable "
                          f"can be declared with a variable \"{name}\" {{}} block.", ctx, t, "undeclared-variable")
            else:
                self.used_vars.add(name)
        elif root == "local":
            name = first_attr()
            if name is None:
                self.diag(ERROR, "Invalid reference", "The \"local\" object cannot be accessed directly. "
                          "Instead, access one of its attributes.", ctx, t, "reference")
            elif name not in self.m.locals:
                self.diag(ERROR, "Reference to undeclared local value",
                          f"A local value with the name \"{name}\" has not been declared.",
                          ctx, t, "undeclared-local")
            else:
                self.used_locals.add(name)
                if self._owner == f"local.{name}":
                    self.diag(ERROR, "Self-referencing local value",
                              f"Local value \"{name}\" cannot use its own result as part of its expression.",
                              ctx, t, "self-reference")
                elif self._owner:
                    self.edges.setdefault(self._owner, set()).add(f"local.{name}")
        elif root == "module":
            self.module_ref(t, ctx, keyed)
        elif root in ("data", "ephemeral"):
            if len(steps) < 2 or steps[0].kind != "attr" or steps[1].kind != "attr":
                what = "data source" if root == "data" else "ephemeral resource"
                self.diag(ERROR, "Invalid reference",
                          f"The \"{root}\" object must be followed by two attribute names: the {what} type "
                          "and the resource name.", ctx, t, "reference")
                return
            self.resource_ref(t, ctx, root, steps[0].value, steps[1].value, steps[2:], keyed)
        elif root == "count":
            name = first_attr()
            if not ctx.count:
                self.diag(ERROR, "Reference to \"count\" in non-counted context",
                          "The \"count\" object can only be used in \"module\", \"resource\", and \"data\" blocks, "
                          "and only when the \"count\" argument is set.", ctx, t, "count-context")
            elif name != "index":
                self.diag(ERROR, "Invalid \"count\" attribute",
                          f"The \"count\" object does not have an attribute named \"{name}\". The only "
                          "supported attribute is count.index, which is the index of each instance of a "
                          "resource block that has the \"count\" argument set.", ctx, t, "count-context")
        elif root == "each":
            name = first_attr()
            if not ctx.each:
                self.diag(ERROR, "Reference to \"each\" in context without for_each",
                          "The \"each\" object can be used only in \"module\" or \"resource\" blocks, and only "
                          "when the \"for_each\" argument is set.", ctx, t, "each-context")
            elif name not in ("key", "value"):
                self.diag(ERROR, "Invalid \"each\" attribute",
                          f"The \"each\" object does not have an attribute named \"{name}\". The supported "
                          "attributes are each.key and each.value, the current key and value pair of the "
                          "\"for_each\" attribute set.", ctx, t, "each-context")
        elif root == "self":
            if not ctx.self_ok:
                self.diag(ERROR, "Invalid \"self\" reference",
                          "The \"self\" object is not available in this context. This object can be used only "
                          "in resource provisioner, connection, and postcondition blocks.", ctx, t, "self-context")
        elif root == "path":
            name = first_attr()
            if name not in _PATH_ATTRS:
                self.diag(ERROR, "Invalid \"path\" attribute",
                          f"The \"path\" object does not have an attribute named \"{name}\". The supported "
                          "attributes are path.module, path.root, and path.cwd.", ctx, t, "reference")
        elif root == "terraform":
            name = first_attr()
            if name == "applying" and not self.m.restated.has("terraform.applying"):
                name = None
            if name not in _TERRAFORM_ATTRS:
                self.diag(ERROR, "Invalid \"terraform\" attribute",
                          f"The \"terraform\" object does not have an attribute named \"{name}\". The only "
                          "supported attribute is terraform.workspace, the name of the currently-selected "
                          "workspace.", ctx, t, "reference")
        elif root in ("resource",) and len(steps) >= 2:
            self.resource_ref(t, ctx, "managed", steps[0].value, steps[1].value, steps[2:], keyed)
        else:
            name = first_attr()
            if name is None:
                self.diag(ERROR, "Invalid reference",
                          "A reference to a resource type must be followed by at least one attribute access, "
                          "specifying the resource name.", ctx, t, "reference")
                return
            self.resource_ref(t, ctx, "managed", root, name, steps[1:], keyed)

    def resource_ref(self, t: A.ScopeTraversal, ctx: Ctx, mode: str, typ: str, name: str,
                     rest: list[A.Step], keyed: bool) -> None:
        addr = resource_addr(mode, typ, name)
        r: Optional[Resource] = self.m.resources.get(addr)
        if r is None:
            self.diag(ERROR, "Reference to undeclared resource",
                      f"A {mode} resource \"{typ}\" \"{name}\" has not been declared in the root module.",
                      ctx, t, "undeclared-resource")
            return
        if r.scope is not None and r.scope != ctx.check:
            self.diag(ERROR, "Reference to scoped resource",
                      f"The referenced {mode} resource \"{typ}\" \"{name}\" is not available from this context.",
                      ctx, t, "undeclared-resource")
            return
        self.used_resources.add(addr)
        if self._owner and self._owner != addr:
            self.edges.setdefault(self._owner, set()).add(addr)
        if self._owner == addr and not ctx.self_ok:
            self.diag(ERROR, "Self-referential block",
                      f"Configuration for {addr} may not refer to itself.", ctx, t, "self-reference")
        multi = r.has_count or r.has_for_each
        has_key = bool(rest) and rest[0].kind == "index"
        if not rest and keyed:
            has_key = True                      # aws_x.y[count.index] / aws_x.y[*]
        if has_key and not multi:
            self.diag(ERROR, "Unexpected resource instance key",
                      f"Because {addr} does not have \"count\" or \"for_each\" set, references to it must not "
                      "include an index key. Remove the bracketed index to refer to the single instance of "
                      "this resource.", ctx, t, "instance-key")
        elif multi and rest and rest[0].kind == "attr":
            which = "count" if r.has_count else "for_each"
            hint = (f"{addr}[count.index]" if r.has_count else f"{addr}[each.key]")
            self.diag(ERROR, "Missing resource instance key",
                      f"Because {addr} has \"{which}\" set, its attributes must be accessed on specific "
                      f"instances.\n\nFor example, to correlate with indices of a referring resource, use:\n"
                      f"    {hint}", ctx, t, "instance-key")

    def module_ref(self, t: A.ScopeTraversal, ctx: Ctx, keyed: bool) -> None:
        steps = t.steps
        if not steps or steps[0].kind != "attr":
            self.diag(ERROR, "Invalid reference", "The \"module\" object cannot be accessed directly. "
                      "Instead, access one of its attributes.", ctx, t, "reference")
            return
        name = steps[0].value
        mc = self.m.module_calls.get(name)
        if mc is None:
            self.diag(ERROR, "Reference to undeclared module",
                      f"No module call named \"{name}\" is declared in the root module.",
                      ctx, t, "undeclared-module")
            return
        self.used_modules.add(name)
        rest = steps[1:]
        multi = mc.has_count or mc.has_for_each
        has_key = bool(rest) and rest[0].kind == "index"
        if has_key:
            rest = rest[1:]
        elif not rest and keyed:
            has_key = True
        if has_key and not multi:
            self.diag(ERROR, "Unexpected module instance key",
                      f"Because module.{name} does not have \"count\" or \"for_each\" set, references to it "
                      "must not include an index key.", ctx, t, "instance-key")
            return
        if multi and not has_key and rest and rest[0].kind == "attr":
            which = "count" if mc.has_count else "for_each"
            self.diag(ERROR, "Missing module instance key",
                      f"Because module.{name} has \"{which}\" set, its outputs must be accessed on specific "
                      f"instances, for example module.{name}[0].{rest[0].value}.", ctx, t, "instance-key")
            return
        child = self.child(mc)
        if child is not None and rest and rest[0].kind == "attr" and rest[0].value not in child.outputs:
            self.diag(ERROR, "Unsupported attribute",
                      f"This object does not have an attribute named \"{rest[0].value}\".",
                      ctx, rest[0], "module-output")


# ---------------------------------------------------------------------------
def check_tfvars(m: Module) -> list[Diagnostic]:
    """``terraform.tfvars`` / ``*.auto.tfvars`` beside the module: values must be
    literals, should name declared variables and must fit their types."""
    diags: list[Diagnostic] = []
    try:
        names = sorted(os.listdir(m.dir))
    except OSError:
        return diags
    for name in names:
        if not (name == "terraform.tfvars" or name.endswith(".auto.tfvars")):
            continue
        path = os.path.join(m.dir, name)
        disp = os.path.join(os.path.dirname(m.files[0].path), name) if m.files else path
        _, body, pd, _ = parse_path(path, disp)
        diags.extend(pd)
        for item in body.items:
            if isinstance(item, A.Block):
                diags.append(Diagnostic(ERROR, "Unexpected block",
                                        "Variable definition files may contain only arguments.",
                                        disp, item.pos.line, item.pos.col, "tfvars"))
                continue
            var = m.variables.get(item.name)
            if var is None:
                diags.append(Diagnostic(WARNING, "Value for undeclared variable",
                                        f"The root module does not declare a variable named \"{item.name}\" but "
                                        f"a value was found in file \"{name}\". If you meant to use this value, "
                                        "add a \"variable\" block to the configuration.",
                                        disp, item.pos.line, item.pos.col, "tfvars"))
                continue
            val = typeexpr.literal_value(item.expr)
            for node in item.expr.walk():
                if isinstance(node, (A.ScopeTraversal, A.FunctionCall)):
                    diags.append(Diagnostic(ERROR, "Variables not allowed" if isinstance(node, A.ScopeTraversal)
                                            else "Function calls not allowed",
                                            "Variables may not be used here." if isinstance(node, A.ScopeTraversal)
                                            else "Functions may not be called here.",
                                            disp, node.pos.line, node.pos.col, "tfvars"))
                    break
            if var.type is not None:
                err = typeexpr.conformance_error(val, var.type)
                if err:
                    diags.append(Diagnostic(ERROR, "Invalid value for input variable",
                                            f"The given value is not suitable for var.{item.name} declared at "
                                            f"{var.file}:{var.line}: {err}.",
                                            disp, item.pos.line, item.pos.col, "tfvars"))
    return diags


@dataclass
class ValidateResult:
    module: Module
    validator: Validator
    diags: list[Diagnostic]

    @property
    def valid(self) -> bool:
        return not has_errors(self.diags)


def validate_module(directory: str, display_root: Optional[str] = None,
                    terraform_version: Optional[str] = None) -> ValidateResult:
    """``terraform_version`` ("1.5", "1.5.7"): the release whose rules apply; by default
    the lowest one the module's ``required_version`` admits (``tfversion``)."""
    restated = tfversion.from_flag(terraform_version) if terraform_version else None
    m = load_module(directory, display_root, restated)
    v = Validator(m)
    diags = list(m.diags)
    if m.broken_files:
        # like Terraform: nothing beyond the syntax errors is reported until
        # every file parses, since missing declarations would cascade
        diags = [d for d in diags if d.rule == "syntax"]
        diags.sort(key=lambda d: (d.file, d.line, d.col))
        return ValidateResult(m, v, diags)
    diags.extend(v.run())
    diags.extend(check_tfvars(m))
    diags.sort(key=lambda d: (d.file, d.line, d.col))
    return ValidateResult(m, v, diags)
