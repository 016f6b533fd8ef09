"""Dependency lock file parsing and its agreement with required_providers."""
from nvidia_terraform_modules_b200.config.lockfile import check_lock_consistency, load_lockfile
from nvidia_terraform_modules_b200.config.model import load_module

H1 = "h1:" + "A" * 43 + "="
ZH = "zh:" + "0" * 64


def lock_block(fqn, version, constraints=None, hashes=(H1, ZH)):
    lines = [f'provider "{fqn}" {{', f'  version     = "{version}"']
    if constraints:
        lines.append(f'  constraints = "{constraints}"')
    lines.append("  hashes = [")
    lines.extend(f'    "{h}",' for h in hashes)
    lines.append("  ]")
    lines.append("}")
    return "\n".join(lines) + "\n"


CONFIG = ('terraform {\n  required_providers {\n    aws = {\n      source  = "hashicorp/aws"\n'
          '      version = "~>4.45.0"\n    }\n    azapi = {\n      source  = "Azure/azapi"\n'
          '      version = "~>1.4.0"\n    }\n  }\n}\nresource "aws_a" "x" {}\nresource "helm_release" "y" {}\n'
          'resource "azapi_resource" "z" {}\n')


def findings(module_dir, lock_text, config=CONFIG, extra=None):
    files = {"main.tf": config, ".terraform.lock.hcl": lock_text}
    files.update(extra or {})
    d = module_dir(files)
    m = load_module(d)
    return [(x.rule, x.severity) for x in check_lock_consistency(m, load_lockfile(d))]


def test_parse(module_dir):
    d = module_dir({".terraform.lock.hcl": "# header\n" + lock_block("registry.terraform.io/hashicorp/aws", "4.45.0",
                                                                       ">= 3.72.0, ~> 4.45.0")})
    lock = load_lockfile(d)
    p = lock.providers["registry.terraform.io/hashicorp/aws"]
    assert str(p.version) == "4.45.0" and p.constraints_text == ">= 3.72.0, ~> 4.45.0" and len(p.hashes) == 2
    assert lock.diags == []


def test_consistent_lock_is_quiet(module_dir):
    text = (lock_block("registry.terraform.io/hashicorp/aws", "4.45.0", ">= 3.72.0, ~> 4.45.0")
            + lock_block("registry.terraform.io/hashicorp/helm", "2.9.0")
            + lock_block("registry.terraform.io/azure/azapi", "1.4.0", "~> 1.4.0"))
    assert findings(module_dir, text) == []


def test_missing_and_out_of_range_and_unused(module_dir):
    text = (lock_block("registry.terraform.io/hashicorp/aws", "4.46.1", "~> 4.46.0")
            + lock_block("registry.terraform.io/azure/azapi", "1.4.0", "~> 1.4.0")
            + lock_block("registry.terraform.io/hashicorp/external", "2.3.1"))
    assert sorted(findings(module_dir, text)) == [("lock-missing", "error"), ("lock-unused", "notice"),
                                                  ("lock-version", "error")]


def test_recorded_constraints_drift_is_a_notice(module_dir):
    text = (lock_block("registry.terraform.io/hashicorp/aws", "4.45.3", "~> 4.0")
            + lock_block("registry.terraform.io/hashicorp/helm", "2.9.0")
            + lock_block("registry.terraform.io/azure/azapi", "1.4.0", "~> 1.4.0"))
    assert findings(module_dir, text) == [("lock-constraints", "notice")]


def test_malformed_lock_entries(module_dir):
    text = ('provider "hashicorp/aws" {\n  version = "4.45.0"\n}\n'
            'provider "registry.terraform.io/hashicorp/helm" {\n  version = "two"\n  hashes = ["md5:abc"]\n}\n'
            'frob = 1\n')
    d = module_dir({".terraform.lock.hcl": text})
    got = [x.summary for x in load_lockfile(d).diags]
    assert got == ["Invalid provider source address", "Invalid provider version number", "Invalid provider hash",
                   "Unsupported argument"]


def test_child_module_needs_are_included(module_dir):
    text = (lock_block("registry.terraform.io/hashicorp/aws", "4.45.0", "~> 4.45.0")
            + lock_block("registry.terraform.io/hashicorp/helm", "2.9.0")
            + lock_block("registry.terraform.io/azure/azapi", "1.4.0", "~> 1.4.0"))
    d = module_dir({"main.tf": 'module "c" {\n  source = "./c"\n}\n', "c/main.tf": CONFIG,
                    ".terraform.lock.hcl": text})
    root, child = load_module(d), load_module(d + "/c")
    assert [x.rule for x in check_lock_consistency(root, load_lockfile(d), (child,))] == []
    assert {x.rule for x in check_lock_consistency(root, load_lockfile(d))} == {"lock-unused"}


def test_lock_without_hashes_is_flagged(module_dir):
    text = lock_block("registry.terraform.io/hashicorp/null", "3.2.1", hashes=())
    assert ("lock-hashes", "warning") in findings(module_dir, text, config='resource "null_resource" "n" {}\n')
