"""Seeded generator of synthetic Terraform module trees.

/root/reference does not exist on the measurement box and its sources may not
be copied, so ``bench.py``, ``smoke()`` and the property tests need a corpus of
their own.  ``write_tree`` produces module directories with the same *shape* as
the reference's (a cluster module with ~30 variables, a few
resources with nested blocks and object-valued arguments, a Helm release,
outputs, ``versions.tf``, a lock file, terraform-docs tables in a README, and an
``examples/`` child that calls the parent through ``source = "../.."``) -- but
every name and value is made up here.

The generated text is canonical by construction (the generator does its own
alignment).  ``defects`` then plants a chosen number of known problems, each
of which one specific check must report; tests use that to verify the checks
find exactly what was planted and nothing else.
"""
from __future__ import annotations

import os
import random
import re
from dataclasses import dataclass, field

_WORDS = ("alpha bravo cedar delta ember fjord grove haven iris jade kelp lumen mesa noble onyx "
          "pearl quartz ridge slate tidal umber vale wren xenon yarrow zephyr").split()
_KINDS = ("string", "number", "bool", "list(string)", "map(string)")


def _aligned(pairs: list[tuple[str, str]], indent: int) -> list[str]:
    """Lines ``name = value`` with the equals signs of the group aligned."""
    width = max(len(k) for k, _ in pairs)
    pad = " " * indent
    return [f"{pad}{k}{' ' * (width - len(k))} = {v}" for k, v in pairs]


@dataclass
class Planted:
    kind: str          # which check must fire
    file: str
    note: str = ""


@dataclass
class SynthTree:
    root: str
    modules: list[str] = field(default_factory=list)
    planted: list[Planted] = field(default_factory=list)
    bytes: int = 0
    files: int = 0


class _Gen:
    def __init__(self, rng: random.Random, n_vars: int, n_resources: int):
        self.rng = rng
        self.n_vars = n_vars
        self.n_resources = n_resources
        self.vars: list[tuple[str, str]] = []

    def name(self, k: int = 2) -> str:
        return "_".join(self.rng.choice(_WORDS) for _ in range(k))

    def default_for(self, kind: str) -> str:
        r = self.rng
        if kind == "string":
            return f"\"{self.name(1)}-{r.randint(1, 99)}\""
        if kind == "number":
            return str(r.randint(1, 512))
        if kind == "bool":
            return r.choice(("true", "false"))
        if kind == "list(string)":
            return "[" + ", ".join(f"\"{self.name(1)}\"" for _ in range(r.randint(0, 3))) + "]"
        return "{}"

    def variables_tf(self) -> str:
        out: list[str] = ["# Inputs of the synthetic cluster module", ""]
        seen = set()
        while len(self.vars) < self.n_vars:
            n = self.name(self.rng.randint(2, 3))
            if n in seen:
                continue
            seen.add(n)
            kind = self.rng.choice(_KINDS)
            self.vars.append((n, kind))
            pairs = [("type", kind), ("default", self.default_for(kind)),
                     ("description", f"\"Synthetic input {len(self.vars)}: {n.replace('_', ' ')}\"")]
            self.rng.shuffle(pairs)
            out.append(f"variable \"{n}\" {{")
            out.extend(_aligned(pairs, 2))
            out.append("}")
            out.append("")
        out.append("variable \"cluster_name\" {")
        out.extend(_aligned([("type", "string"), ("description", "\"Name of the cluster (required)\"")], 2))
        out.append("}")
        self.vars.append(("cluster_name", "string"))
        return "\n".join(out) + "\n"

    def var_ref(self, kind: str | None = None) -> str:
        pool = [n for n, k in self.vars if kind is None or k == kind] or ["cluster_name"]
        return "var." + self.rng.choice(pool)

    def main_tf(self) -> tuple[str, list[str]]:
        r = self.rng
        out: list[str] = ["/***************************", "Synthetic cluster", "***************************/"]
        names: list[str] = []
        for i in range(self.n_resources):
            rname = self.name(2)
            while rname in names:
                rname = self.name(2)
            names.append(rname)
            pairs = [("name", f"\"tf-${{var.cluster_name}}-{i}\""),
                     ("region", self.var_ref("string")),
                     ("node_count", self.var_ref("number")),
                     ("enabled", f"{self.var_ref('bool')} ? true : false"),
                     ("labels", f"merge({self.var_ref('map(string)')}, {{ managed_by = \"synthetic\" }})")]
            if i:
                pairs.append(("parent_id", f"synth_cluster.{names[r.randrange(i)]}.id"))
            out.append(f"resource \"synth_cluster\" \"{rname}\" {{")
            out.extend(_aligned(pairs, 2))
            out.append("")
            out.append("  node_pool {")
            out.extend(_aligned([("machine_type", self.var_ref("string")),
                                 ("zones", f"concat({self.var_ref('list(string)')}, [\"zone-{i}\"])"),
                                 ("disk_size_gb", str(r.choice((100, 256, 512))))], 4))
            out.append("  }")
            out.append("")
            out.append("  tags = {")
            out.extend(_aligned([("group", "\"synthetic\""), ("index", str(i))], 4))
            out.append("  }")
            out.append("}")
            out.append("")
        body = "\n".join(out)
        used = set(re.findall(r"var\.([A-Za-z0-9_]+)", body))
        rest = [n for n, _ in self.vars if n not in used]
        if rest:
            # every declared input is consumed somewhere, as in a tidy module
            out.append("resource \"synth_settings\" \"rest\" {")
            out.append("  values = {")
            out.extend(_aligned([(n, f"var.{n}") for n in rest], 4))
            out.append("  }")
            out.append("}")
            out.append("")
        out.append("data \"synth_zones\" \"available\" {}")
        out.append("")
        out.append("locals {")
        out.extend(_aligned([("cluster_ids", "[" + ", ".join(f"synth_cluster.{n}.id" for n in names) + "]"),
                             ("zone_names", "data.synth_zones.available.names")], 2))
        out.append("}")
        out.append("")
        out.append("resource \"helm_release\" \"operator\" {")
        out.extend(_aligned([("count", "length(local.cluster_ids) > 0 ? 1 : 0"), ("name", "\"synthetic-operator\""),
                             ("chart", "\"operator\""), ("namespace", self.var_ref("string")),
                             ("atomic", "true")], 2))
        out.append("")
        out.append("  set {")
        out.extend(_aligned([("name", "\"zones\""), ("value", "join(\",\", local.zone_names)")], 4))
        out.append("  }")
        out.append("}")
        return "\n".join(out) + "\n", names

    def outputs_tf(self, names: list[str]) -> str:
        out: list[str] = []
        for n in names:
            out.append(f"output \"{n}_id\" {{")
            out.extend(_aligned([("description", f"\"Id of cluster {n.replace('_', ' ')}\""),
                                 ("value", f"synth_cluster.{n}.id")], 2))
            out.append("}")
            out.append("")
        out.append("output \"operator_release\" {")
        out.extend(_aligned([("description", "\"Name of the Helm release, when one was created\""),
                             ("value", "one(helm_release.operator[*].name)")], 2))
        out.append("}")
        return "\n".join(out) + "\n"


VERSIONS_TF = """terraform {
  required_providers {
    synth = {
      source  = "example/synth"
      version = "~> 1.4.0"
    }
    helm = {
      source  = "hashicorp/helm"
      version = "~> 2.9.0"
    }
  }

  required_version = ">= 1.3.0"
}
"""

_H1 = "h1:" + "A" * 43 + "="
_ZH = "zh:" + "0" * 64

LOCK_HCL = f"""# This file is maintained automatically by "terraform init".
# Manual edits may be lost in future updates.

provider "registry.terraform.io/example/synth" {{
  version     = "1.4.2"
  constraints = "~> 1.4.0"
  hashes = [
    "{_H1}",
    "{_ZH}",
  ]
}}

provider "registry.terraform.io/hashicorp/helm" {{
  version     = "2.9.0"
  constraints = "~> 2.9.0"
  hashes = [
    "{_H1}",
    "{_ZH}",
  ]
}}
"""


def _write(tree: SynthTree, rel: str, text: str) -> None:
    path = os.path.join(tree.root, rel)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w", encoding="utf-8", newline="") as fh:
        fh.write(text)
    tree.bytes += len(text.encode("utf-8"))
    tree.files += 1


def write_tree(root: str, *, seed: int = 0, modules: int = 3, n_vars: int = 30, n_resources: int = 6,
               defects: int = 0, with_readme: bool = True) -> SynthTree:
    """Write ``modules`` cluster modules (each with an ``examples/basic`` child)
    under ``root``; plant ``defects`` known problems per module."""
    from .config.lockfile import load_lockfile
    from .config.model import load_module
    from .docs.tfdocs import render

    rng = random.Random(seed)
    tree = SynthTree(root)
    for mi in range(modules):
        mod = f"cloud{mi}"
        g = _Gen(rng, n_vars, n_resources)
        files = {"variables.tf": g.variables_tf()}
        main, names = g.main_tf()
        files["main.tf"] = main
        files["outputs.tf"] = g.outputs_tf(names)
        files["versions.tf"] = VERSIONS_TF
        example = {
            "main.tf": "\n".join(["module \"cluster\" {"] + _aligned(
                [("source", "\"../..\""), ("cluster_name", "var.cluster_name")], 2) + ["}", ""]),
            "variables.tf": "\n".join(["variable \"cluster_name\" {"] + _aligned(
                [("type", "string"), ("default", "\"example\""),
                 ("description", "\"Name passed to the cluster module\"")], 2) + ["}", ""]),
            "outputs.tf": "\n".join(["output \"operator_release\" {"] + _aligned(
                [("description", "\"Release name reported by the cluster module\""),
                 ("value", "module.cluster.operator_release")], 2) + ["}", ""]),
            "versions.tf": "terraform {\n  required_version = \">= 1.3.0\"\n}\n",
        }

        planted = _plant(rng, files, g, defects, mod)
        tree.planted.extend(planted)
        for name, text in files.items():
            _write(tree, f"{mod}/{name}", text)
        _write(tree, f"{mod}/.terraform.lock.hcl", LOCK_HCL)
        _write(tree, f"{mod}/terraform.tfvars", "# cluster_name = \"demo\"\n")
        for name, text in example.items():
            _write(tree, f"{mod}/examples/basic/{name}", text)
        _write(tree, f"{mod}/examples/basic/.terraform.lock.hcl", LOCK_HCL)
        if with_readme:
            for rel in (mod, f"{mod}/examples/basic"):
                d = os.path.join(root, rel)
                m = load_module(d, root)
                _write(tree, f"{rel}/README.md", f"# {rel}\n\nSynthetic module.\n\n" + render(m, load_lockfile(d, root)))
        tree.modules.extend([mod, f"{mod}/examples/basic"])
    return tree


# kinds of planted defect -> the rule / check expected to report it
DEFECT_KINDS = ("fmt-align", "fmt-tab", "fmt-trailing-space", "unused-variable", "undeclared-variable",
                "unknown-function", "unused-local", "length-of-object")


def _plant(rng: random.Random, files: dict[str, str], g: _Gen, count: int, mod: str) -> list[Planted]:
    planted: list[Planted] = []
    kinds = list(DEFECT_KINDS)
    for i in range(count):
        kind = kinds[i % len(kinds)]
        if kind == "fmt-align":
            # squeeze the padding out of one aligned line in outputs.tf
            lines = files["outputs.tf"].split("\n")
            idx = next(j for j, ln in enumerate(lines) if ln.startswith("  value") and "  =" in ln)
            lines[idx] = "  value = " + lines[idx].split("=", 1)[1].strip()
            files["outputs.tf"] = "\n".join(lines)
            planted.append(Planted("fmt", f"{mod}/outputs.tf", "alignment"))
        elif kind == "fmt-tab":
            files["versions.tf"] = files["versions.tf"].replace("      source  = \"example/synth\"",
                                                                "\tsource  = \"example/synth\"", 1)
            planted.append(Planted("fmt", f"{mod}/versions.tf", "tab"))
        elif kind == "fmt-trailing-space":
            files["variables.tf"] = files["variables.tf"].replace(
                "variable \"cluster_name\" {", "variable \"cluster_name\" { ", 1)
            planted.append(Planted("fmt", f"{mod}/variables.tf", "trailing space"))
        elif kind == "unused-variable":
            files["variables.tf"] += "\n".join(["", "variable \"never_read\" {"] + _aligned(
                [("type", "string"), ("default", "\"\""), ("description", "\"planted\"")], 2) + ["}", ""])
            planted.append(Planted("unused-variable", f"{mod}/variables.tf"))
        elif kind == "undeclared-variable":
            files["outputs.tf"] += "\noutput \"ghost\" {\n  description = \"planted\"\n" \
                                   "  value       = var.does_not_exist\n}\n"
            planted.append(Planted("undeclared-variable", f"{mod}/outputs.tf"))
        elif kind == "unknown-function":
            files["outputs.tf"] += "\noutput \"shout\" {\n  description = \"planted\"\n" \
                                   "  value       = uppercase(var.cluster_name)\n}\n"
            planted.append(Planted("function", f"{mod}/outputs.tf"))
        elif kind == "unused-local":
            files["main.tf"] += "\nlocals {\n  planted_unused = 42\n}\n"
            planted.append(Planted("unused-local", f"{mod}/main.tf"))
        elif kind == "length-of-object":
            files["main.tf"] += "\nresource \"synth_gate\" \"planted\" {\n" \
                                "  count = length(data.synth_zones.available) > 0 ? 1 : 0\n}\n"
            planted.append(Planted("length-of-object", f"{mod}/main.tf"))
    return planted
