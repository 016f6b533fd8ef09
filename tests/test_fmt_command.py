"""fmt command: flags, file selection, output and exit codes as `terraform fmt`."""
import io
import os

import pytest

from nvidia_terraform_modules_b200.fmt.command import discover, main, run_fmt

BAD = 'a   = 1\nbb = 2\n'
GOOD = 'a  = 1\nbb = 2\n'


@pytest.fixture
def tree(tmp_path, monkeypatch):
    (tmp_path / "mod" / "examples" / "x").mkdir(parents=True)
    (tmp_path / "mod" / ".terraform").mkdir()
    (tmp_path / "mod" / "main.tf").write_text(BAD)
    (tmp_path / "mod" / "ok.tf").write_text(GOOD)
    (tmp_path / "mod" / "terraform.tfvars").write_text("x=1\n")
    (tmp_path / "mod" / ".terraform.lock.hcl").write_text("provider   \"x\" {}\n")
    (tmp_path / "mod" / "README.md").write_text("a   = 1\n")
    (tmp_path / "mod" / "examples" / "x" / "main.tf").write_text(BAD)
    (tmp_path / "mod" / ".terraform" / "cached.tf").write_text(BAD)
    monkeypatch.chdir(tmp_path)
    return tmp_path


def run(argv, stdin=""):
    out, err = io.StringIO(), io.StringIO()
    code = main(argv, stdout=out, stderr=err, stdin=io.StringIO(stdin))
    return code, out.getvalue(), err.getvalue()


def test_discover_skips_hidden_dirs_lock_files_and_other_suffixes(tree):
    assert discover("mod", recursive=False) == ["mod/main.tf", "mod/ok.tf", "mod/terraform.tfvars"]
    assert discover("mod", recursive=True) == ["mod/examples/x/main.tf", "mod/main.tf", "mod/ok.tf",
                                               "mod/terraform.tfvars"]


def test_check_lists_files_and_exits_3_without_writing(tree):
    code, out, _ = run(["-check", "-recursive", "mod"])
    assert code == 3
    assert out.splitlines() == ["mod/examples/x/main.tf", "mod/main.tf", "mod/terraform.tfvars"]
    assert (tree / "mod" / "main.tf").read_text() == BAD


def test_check_is_clean_exit_0(tree):
    assert run(["-check", "mod/ok.tf"])[0] == 0


def test_default_rewrites_files_in_place(tree):
    code, out, _ = run(["mod"])
    assert code == 0 and "mod/main.tf" in out
    assert (tree / "mod" / "main.tf").read_text() == GOOD
    assert (tree / "mod" / "examples" / "x" / "main.tf").read_text() == BAD   # not recursive
    assert run(["-check", "mod"])[0] == 0


def test_write_false_and_list_false_and_diff(tree):
    code, out, _ = run(["-write=false", "-list=false", "-diff", "mod/main.tf"])
    assert code == 0
    assert out.startswith("--- old/mod/main.tf\n+++ new/mod/main.tf\n@@ -1,2 +1,2 @@\n")
    assert "-a   = 1\n+a  = 1\n" in out
    assert (tree / "mod" / "main.tf").read_text() == BAD


def test_default_target_is_current_directory(tree, monkeypatch):
    monkeypatch.chdir(tree / "mod")
    code, out, _ = run(["-check"])
    assert code == 3 and out.splitlines() == ["main.tf", "terraform.tfvars"]


def test_stdin_to_stdout(tree):
    code, out, _ = run(["-"], stdin=BAD)
    assert (code, out) == (0, GOOD)
    assert run(["-check", "-"], stdin=BAD)[0] == 3
    assert run(["-check", "-"], stdin=GOOD)[0] == 0


def test_syntax_error_exits_2_and_reports_position(tree):
    (tree / "mod" / "broken.tf").write_text("a = (\n")
    code, out, err = run(["-check", "mod"])
    assert code == 2
    assert "mod/broken.tf:2" in err and "Error:" in err
    assert out == ""                      # -check output is buffered and dropped on errors
    code, out, err = run(["-write=false", "mod"])
    assert code == 2 and out.splitlines() == ["mod/main.tf", "mod/terraform.tfvars"]


def test_unknown_flag_and_bad_targets(tree):
    assert run(["-frobnicate"])[0] == 1
    code, _, err = run(["mod/README.md"])
    assert code == 2 and "Only .tf, .tfvars" in err
    code, _, err = run(["nope"])
    assert code == 2 and "No file or directory at nope" in err


def test_result_counts_bytes(tree):
    res = run_fmt(["mod"], recursive=True, check=True)
    assert res.bytes_checked == sum(os.path.getsize(p) for p in
                                    ["mod/examples/x/main.tf", "mod/main.tf", "mod/ok.tf", "mod/terraform.tfvars"])


def test_in_process_diff_equals_the_diff_program_on_white_space_edits():
    """`terraform fmt -diff` shells out to `diff -u`; the fallback used where no such program
    exists must print the same bytes -- hunk ranges (`-3`, `-2,0`), context, and the
    "\\ No newline at end of file" marker included."""
    import random
    from nvidia_terraform_modules_b200.fmt import command
    if not command._DIFF_PROGRAM:
        pytest.skip("no diff program on this machine")
    rng = random.Random(5)
    compared = 0
    for _ in range(150):
        n = rng.randint(1, 30)
        old = [f"name{i} = {rng.randint(0, 99)}\n" for i in range(n)]
        new = list(old)
        for _ in range(rng.randint(1, 4)):
            i = rng.randrange(n)
            new[i] = new[i].replace(" = ", " " * rng.randint(2, 5) + "= ")
        if rng.random() < 0.3:
            old[-1], new[-1] = old[-1].rstrip("\n"), new[-1].rstrip("\n")
        a, b = "".join(old), "".join(new)
        if a == b:
            continue
        compared += 1
        assert command.unified_diff("x.tf", a, b, external=False) == command.unified_diff("x.tf", a, b, external=True)
    assert compared > 100
