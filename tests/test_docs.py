"""README tables: generation in terraform-docs' layout and drift detection."""
import pytest

from nvidia_terraform_modules_b200.config.model import load_module
from nvidia_terraform_modules_b200.docs.tfdocs import (BEGIN_MARK, END_MARK, InjectRefused, check_readme, inject,
                                                       parse_readme, render)

FILES = {
    "versions.tf": 'terraform {\n  required_version = ">= 1.2.4"\n  required_providers {\n    aws = {\n'
                   '      source  = "hashicorp/aws"\n      version = "~>4.45.0"\n    }\n    pinned = {\n'
                   '      source  = "Acme/pinned"\n      version = "1.2.3"\n    }\n    loose = {\n'
                   '      source = "acme/loose"\n    }\n  }\n}\n',
    "variables.tf": '''variable "cluster_name" {
  type        = string
  description = "Name of the vpc_id | cluster, see `a_b`"
}
variable "sizes" {
  default = [1, 2]
}
// Comment used as description
variable "vpc" {
  type = object({
    vpc_id     = string
    subnet_ids = list(string)
  })
  default = null
}
variable "flag" {
  type    = bool
  default = true
}
''',
    "main.tf": 'resource "aws_eks_thing" "a" {}\ndata "aws_ami" "l" {}\nresource "pinned_x" "p" {}\n'
               'data "aws_b" "beta" {\n  provider = aws-beta\n}\n'
               'module "vpc" {\n  source  = "terraform-aws-modules/vpc/aws"\n  version = "4.0.2"\n}\n'
               'module "local" {\n  source = "../.."\n}\n',
    "outputs.tf": '# where it lives\noutput "endpoint" {\n  value = 1\n}\noutput "ca" {\n  description = "CA data"\n'
                  '  value = 2\n}\n',
}


def module(module_dir, readme=None):
    files = dict(FILES)
    if readme is not None:
        files["README.md"] = readme
    d = module_dir(files)
    return d, load_module(d)


def test_render_rows(module_dir):
    _, m = module(module_dir)
    text = render(m)
    rows = set(text.splitlines())
    assert '| <a name="requirement_terraform"></a> [terraform](#requirement\\_terraform) | >= 1.2.4 |' in rows
    assert '| <a name="requirement_aws"></a> [aws](#requirement\\_aws) | ~>4.45.0 |' in rows
    assert not any("requirement_loose" in r for r in rows)         # no constraint, no row
    assert '| <a name="module_vpc"></a> [vpc](#module\\_vpc) | terraform-aws-modules/vpc/aws | 4.0.2 |' in rows
    assert '| <a name="module_local"></a> [local](#module\\_local) | ../.. | n/a |' in rows
    assert ('| [aws_eks_thing.a](https://registry.terraform.io/providers/hashicorp/aws/latest/docs/resources/eks_thing)'
            ' | resource |') in rows
    assert ('| [pinned_x.p](https://registry.terraform.io/providers/Acme/pinned/1.2.3/docs/resources/x) | resource |'
            ) in rows
    assert ('| [aws-beta_aws_b.beta](https://registry.terraform.io/providers/hashicorp/aws-beta/latest/docs/'
            'data-sources/aws_b) | data source |') in rows
    assert ('| <a name="input_cluster_name"></a> [cluster\\_name](#input\\_cluster\\_name) | '
            'Name of the vpc\\_id \\| cluster, see `a_b` | `string` | n/a | yes |') in rows
    assert ('| <a name="input_sizes"></a> [sizes](#input\\_sizes) | n/a | `list` | '
            '<pre>[<br>  1,<br>  2<br>]</pre> | no |') in rows
    assert ('| <a name="input_vpc"></a> [vpc](#input\\_vpc) | Comment used as description | '
            '<pre>object({<br>    vpc_id     = string<br>    subnet_ids = list(string)<br>  })</pre> | `null` | no |'
            ) in rows
    assert '| <a name="output_endpoint"></a> [endpoint](#output\\_endpoint) | where it lives |' in rows
    assert text.index("## Requirements") < text.index("## Providers") < text.index("## Modules") \
        < text.index("## Resources") < text.index("## Inputs") < text.index("## Outputs")


def test_generated_readme_has_no_drift(module_dir, tmp_path):
    d, m = module(module_dir)
    readme = tmp_path / "README.md"
    readme.write_text("# Title\n\nprose\n\n### Requirements\n\nnot a table\n\n" + render(m))
    assert check_readme(m, str(readme)) == []


def test_every_kind_of_drift_is_reported(module_dir, tmp_path):
    d, m = module(module_dir)
    text = render(m)
    text = text.replace("| `true` | no |", "| `false` | no |")                       # default
    text = text.replace("| `string` | n/a | yes |", "| `number` | n/a | no |")        # type + required
    text = text.replace("| CA data |", "| Certificate |")                             # output description
    text = text.replace("| 4.0.2 |", "| 3.0.0 |")                                     # module version
    text = text.replace("| >= 1.2.4 |", "| >= 1.0 |")                                 # requirement
    lines = [ln for ln in text.splitlines() if "input_sizes" not in ln and "aws_ami.l" not in ln]
    lines.insert(lines.index("## Inputs") + 4,
                 '| <a name="input_ghost"></a> [ghost](#input\\_ghost) | n/a | `any` | n/a | yes |')
    readme = tmp_path / "README.md"
    readme.write_text("\n".join(lines) + "\n")
    got = sorted((x.rule, x.summary.split(":")[0]) for x in check_readme(m, str(readme)))
    assert got == [
        ("docs-drift", 'input "cluster_name"'), ("docs-drift", 'input "cluster_name"'),
        ("docs-drift", 'input "flag"'), ("docs-drift", 'module "vpc"'), ("docs-drift", 'output "ca"'),
        ("docs-drift", 'requirement "terraform"'),
        ("docs-missing-input", 'input "sizes" is not documented in the README'),
        ("docs-missing-resource", 'data source "aws_ami.l" is not documented in the README'),
        ("docs-unknown-input", 'README documents input "ghost" which is not declared'),
    ]


def test_readme_without_tables_is_not_judged(module_dir, tmp_path):
    _, m = module(module_dir)
    readme = tmp_path / "README.md"
    readme.write_text("# Title\n\n## Usage\n\nprose only\n")
    assert check_readme(m, str(readme)) == []


def test_parse_readme_handles_escaped_pipes_and_line_numbers():
    t = parse_readme("x\n## Inputs\n\n| Name | Description |\n|---|---|\n| a | b \\| c |\n\n## Outputs\n\nNo outputs.\n")
    assert t["Inputs"].rows == [(6, ["a", "b | c"])]
    assert t["Outputs"].header == []


def test_inject_replaces_stale_tables_and_keeps_the_prose(module_dir, tmp_path):
    _, m = module(module_dir)
    fresh = render(m)
    stale = fresh.replace("| `true` | no |", "| `false` | no |").replace("| CA data |", "| old |")
    before = "# Title\n\nintro\n\n" + stale + "\n## License\n\nApache\n"
    after = inject(before, fresh)
    assert after == "# Title\n\nintro\n\n" + fresh + "\n## License\n\nApache\n"
    readme = tmp_path / "R.md"
    readme.write_text(after)
    assert check_readme(m, str(readme)) == []
    assert inject(after, fresh) == after                      # idempotent


def test_inject_between_markers_and_append(module_dir):
    _, m = module(module_dir)
    fresh = render(m)
    marked = f"# T\n\n{BEGIN_MARK}\nold junk\n{END_MARK}\n\nfooter\n"
    assert inject(marked, fresh) == f"# T\n\n{BEGIN_MARK}\n{fresh}{END_MARK}\n\nfooter\n"
    # no markers and nothing generated yet: appended between markers, like terraform-docs
    wrapped = f"{BEGIN_MARK}\n{fresh}{END_MARK}\n"
    assert inject("# T\n\nprose\n", fresh) == "# T\n\nprose\n\n" + wrapped
    assert inject("", fresh) == wrapped
    assert inject(inject("# T\n", fresh), fresh) == inject("# T\n", fresh)      # and then idempotent


def test_inject_leaves_hand_written_sections_alone(module_dir):
    """Advisor finding, round 1: a prose "## Requirements" and a "## Usage" between it
    and the generated tables used to be deleted by `docs --write`."""
    _, m = module(module_dir)
    fresh = render(m)
    stale = fresh.replace("| `true` | no |", "| `false` | no |")
    prose = "# T\n\n## Requirements\n\nYou need an AWS account.\n\n## Usage\n\nrun it\n\n"
    tail = "\nSee also the wiki.\n\n## Resources\n\nLinks, not a table.\n"
    after = inject(prose + stale + tail, fresh)
    assert after == prose + fresh + tail
    assert inject(after, fresh) == after


def test_inject_refuses_when_generated_sections_are_split_by_prose(module_dir):
    _, m = module(module_dir)
    fresh = render(m)
    cut = fresh.index("## Inputs")
    split = fresh[:cut] + "## Notes\n\nhand-written, between two generated tables\n\n" + fresh[cut:]
    with pytest.raises(InjectRefused, match="interrupted by other text at line"):
        inject("# T\n\n" + split, fresh)
    loose = fresh[:cut] + "a stray paragraph\n\n" + fresh[cut:]
    with pytest.raises(InjectRefused):
        inject(loose, fresh)


def test_providers_table_is_checked_by_name(module_dir, tmp_path):
    _, m = module(module_dir)
    fresh = render(m)
    readme = tmp_path / "P.md"
    start = fresh.index("## Providers")
    end = fresh.index("## Modules")
    rows = [ln for ln in fresh[start:end].splitlines() if ln.startswith("| <a")]
    assert rows, "the fixture module uses at least one provider"
    doctored = fresh[:start] + "## Providers\n\n| Name | Version |\n|------|---------|\n" \
        "| <a name=\"provider_ghost\"></a> [ghost](#provider\\_ghost) | n/a |\n\n" + fresh[end:]
    readme.write_text(doctored)
    got = sorted((d.rule, d.summary) for d in check_readme(m, str(readme)))
    assert ("docs-unknown-provider", 'README lists provider "ghost" which no resource or data source uses') in got
    assert any(rule == "docs-missing-provider" for rule, _ in got)
    # a version column of "n/a" (README generated before a lock file existed) is not drift
    readme.write_text(fresh.replace(rows[0], rows[0].rsplit("|", 2)[0] + "| n/a |"))
    assert check_readme(m, str(readme)) == []


def test_a_default_that_is_not_a_literal_renders_as_n_a(module_dir):
    """Found by the fuzzer: ``default = [var.a]`` (an error in its own right) made the
    table generator raise inside json.dumps."""
    d = module_dir({"main.tf": 'variable "a" {}\nvariable "v" {\n  default = ([var.a])\n}\n'})
    m = load_module(d, d)
    row = next(ln for ln in render(m).splitlines() if "input_v" in ln)
    assert row.endswith("| `list` | n/a | no |")
