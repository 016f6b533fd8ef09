"""Properties on seeded synthetic module trees (the corpus bench.py and smoke()
use, since the reference cannot travel to the measurement box)."""
import os

import pytest

from nvidia_terraform_modules_b200.check import check_tree, discover_modules
from nvidia_terraform_modules_b200.fmt.command import format_source, run_fmt
from nvidia_terraform_modules_b200.synth import DEFECT_KINDS, write_tree


@pytest.mark.parametrize("seed", [0, 1, 7, 1234])
def test_clean_tree_has_no_findings_at_all(tmp_path, seed):
    tree = write_tree(str(tmp_path), seed=seed, modules=2)
    rep = check_tree(str(tmp_path))
    assert rep.ok and rep.fmt.changed == []
    assert [d.format() for m in rep.modules for d in m.all()] == []
    assert [m.path for m in rep.modules] == sorted(tree.modules, key=lambda p: (p.count(os.sep), p))
    assert rep.bytes_scanned > 0


@pytest.mark.parametrize("seed,defects", [(0, 1), (3, 4), (5, len(DEFECT_KINDS))])
def test_planted_defects_are_found_exactly(tmp_path, seed, defects):
    tree = write_tree(str(tmp_path), seed=seed, modules=2, defects=defects)
    rep = check_tree(str(tmp_path))
    found = sorted((d.rule, d.file) for m in rep.modules for d in m.all())
    found += [("fmt", f) for f in rep.fmt.changed]
    assert sorted(found) == sorted((p.kind, p.file) for p in tree.planted)
    assert not rep.ok


def test_same_seed_same_bytes(tmp_path):
    a = write_tree(str(tmp_path / "a"), seed=42, modules=1)
    b = write_tree(str(tmp_path / "b"), seed=42, modules=1)
    c = write_tree(str(tmp_path / "c"), seed=43, modules=1)
    assert a.bytes == b.bytes and a.files == b.files

    def text(root):
        return open(os.path.join(root, "cloud0", "main.tf")).read()
    assert text(a.root) == text(b.root) != text(c.root)


def test_fixing_with_fmt_clears_the_fmt_findings(tmp_path):
    write_tree(str(tmp_path), seed=2, modules=1, defects=3)     # the three fmt defects
    before = run_fmt([str(tmp_path)], recursive=True, check=True)
    assert len(before.changed) == 3
    run_fmt([str(tmp_path)], recursive=True)                    # rewrite in place
    after = run_fmt([str(tmp_path)], recursive=True, check=True)
    assert after.changed == []


def test_scrambled_white_space_formats_back_to_the_canonical_text(tmp_path):
    """Damage the layout of canonical text without touching tokens; fmt must
    restore the original byte for byte."""
    import random
    import re
    tree = write_tree(str(tmp_path), seed=9, modules=1, n_vars=12, n_resources=3, with_readme=False)
    rng = random.Random(9)
    for rel in ("cloud0/main.tf", "cloud0/variables.tf", "cloud0/outputs.tf", "cloud0/versions.tf"):
        want = open(os.path.join(tree.root, rel)).read()
        lines = []
        for ln in want.split("\n"):
            if ln.strip() and not ln.lstrip().startswith(("/*", "*", "Synthetic")):
                body = ln.strip()
                body = re.sub(r"\s+=\s+", lambda m: " " * rng.randint(0, 3) + "=" + " " * rng.randint(0, 3), body, 1)
                ln = rng.choice(["", " ", "\t", "      "]) + body + rng.choice(["", "  ", "\t"])
            lines.append(ln)
        got, diags = format_source("\n".join(lines), rel)
        assert not diags
        assert got == want, rel


def test_discover_modules_orders_parents_first(tmp_path):
    write_tree(str(tmp_path), seed=0, modules=2, with_readme=False)
    assert discover_modules(str(tmp_path)) == ["cloud0", "cloud1", "cloud0/examples/basic", "cloud1/examples/basic"]
