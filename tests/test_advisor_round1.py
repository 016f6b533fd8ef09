"""Regression tests for the round-1 advisor findings: each case is the
reproduction the advisor gave, turned into an assertion."""
import os

import pytest

from nvidia_terraform_modules_b200.check import check_tree, clear_caches
from nvidia_terraform_modules_b200.config.validate import validate_module
from nvidia_terraform_modules_b200.fmt.command import run_fmt
from nvidia_terraform_modules_b200.hcl.diagnostics import ERROR


def diags(module_dir, text, **more):
    files = {"main.tf": text}
    files.update(more)
    res = validate_module(module_dir(files))
    return [(d.severity, d.summary, d.line) for d in res.diags]


# ------------------------------------------------------- provisioner keywords
def test_provisioner_when_and_on_failure_are_keywords_not_references(module_dir):
    text = ('resource "null_resource" "n" {\n'
            '  provisioner "local-exec" {\n'
            '    when       = destroy\n'
            '    on_failure = continue\n'
            '    command    = "echo bye ${self.id}"\n'
            '  }\n'
            '  provisioner "local-exec" {\n'
            '    when       = create\n'
            '    on_failure = fail\n'
            '    command    = "echo hi"\n'
            '  }\n'
            '}\n')
    assert diags(module_dir, text) == []


def test_provisioner_keywords_reject_anything_else(module_dir):
    text = ('variable "w" {}\n'
            'resource "null_resource" "n" {\n'
            '  provisioner "local-exec" {\n'
            '    when       = later\n'
            '    on_failure = var.w\n'
            '    command    = "x"\n'
            '  }\n'
            '}\n')
    assert diags(module_dir, text) == [(ERROR, 'Invalid "when" keyword', 4), (ERROR, 'Invalid "on_failure" keyword', 5)]
    res = validate_module(module_dir({"main.tf": text}))
    assert res.diags[0].detail == ('The "when" argument requires one of the following keywords: create or destroy.')
    assert res.diags[1].detail == ('The "on_failure" argument requires one of the following keywords: continue or fail.')


def test_quoted_provisioner_keyword_is_a_deprecation_warning(module_dir):
    text = ('resource "null_resource" "n" {\n  provisioner "local-exec" {\n    when    = "destroy"\n'
            '    command = "x"\n  }\n}\n')
    assert diags(module_dir, text) == [("warning", "Quoted keywords are deprecated", 3)]


def test_destroy_provisioner_may_only_look_at_its_own_resource(module_dir):
    text = ('variable "host" {}\n'
            'resource "null_resource" "n" {\n'
            '  count = 2\n'
            '  connection {\n'
            '    host = var.host\n'
            '  }\n'
            '  provisioner "local-exec" {\n'
            '    when    = destroy\n'
            '    command = "echo ${self.id} ${count.index} ${path.module} ${var.host}"\n'
            '  }\n'
            '}\n')
    got = diags(module_dir, text)
    assert got == [(ERROR, "Invalid reference from destroy provisioner", 5),
                   (ERROR, "Invalid reference from destroy provisioner", 9)]


def test_create_time_provisioner_may_refer_to_anything(module_dir):
    text = ('variable "host" {}\nresource "null_resource" "n" {\n  connection {\n    host = var.host\n  }\n'
            '  provisioner "remote-exec" {\n    inline = [var.host]\n    connection {\n      user = var.host\n    }\n'
            '  }\n}\n')
    assert diags(module_dir, text) == []


# ----------------------------------------- ephemeral resources, scoped data sources
def test_ephemeral_resources_resolve(module_dir):
    text = ('ephemeral "aws_secret" "e" {\n  name = "x"\n}\n'
            'ephemeral "aws_secret" "many" {\n  count = 2\n  name  = "y${count.index}"\n}\n'
            'provider "vault" {\n  token = ephemeral.aws_secret.e.value\n  other = ephemeral.aws_secret.many[0].value\n}\n')
    assert diags(module_dir, text) == []


def test_ephemeral_reference_errors_use_terraforms_wording(module_dir):
    text = ('ephemeral "aws_secret" "e" {\n  name = "x"\n}\n'
            'locals {\n  a = ephemeral.aws_secret.nope.value\n  b = ephemeral.aws_secret.e[0].value\n'
            '  c = ephemeral.aws_secret\n}\n')
    res = validate_module(module_dir({"main.tf": text}))
    assert [(d.summary, d.line) for d in res.diags] == [
        ("Reference to undeclared resource", 5), ("Unexpected resource instance key", 6), ("Invalid reference", 7)]
    assert res.diags[0].detail == ('A ephemeral resource "aws_secret" "nope" has not been declared in the root module.')


def test_data_source_nested_in_a_check_block_is_visible_to_its_asserts_only(module_dir):
    text = ('check "health" {\n'
            '  data "http" "probe" {\n    url = "https://example.invalid"\n  }\n'
            '  assert {\n    condition     = data.http.probe.status_code == 200\n    error_message = "down"\n  }\n'
            '}\n'
            'output "leak" {\n  value = data.http.probe.status_code\n}\n')
    res = validate_module(module_dir({"main.tf": text}))
    assert [(d.summary, d.line) for d in res.diags] == [("Reference to scoped resource", 11)]
    assert res.diags[0].detail == 'The referenced data resource "http" "probe" is not available from this context.'


def test_nested_data_source_takes_no_count_and_collides_with_a_module_level_one(module_dir):
    text = ('data "http" "probe" {\n  url = "a"\n}\n'
            'check "c" {\n  data "http" "probe" {\n    count = 1\n    url   = "b"\n  }\n'
            '  assert {\n    condition     = true\n    error_message = "m"\n  }\n}\n')
    assert diags(module_dir, text) == [(ERROR, 'Duplicate data "http" configuration', 5)]
    assert diags(module_dir, text.replace('"probe" {\n    count', '"other" {\n    count')) == [
        (ERROR, 'Invalid "count" attribute', 6)]


# ------------------------------------------------------------- check: no double visit
def test_root_module_with_sub_directories_is_formatted_once(tmp_path):
    (tmp_path / "sub").mkdir()
    (tmp_path / "main.tf").write_text('variable "a" {\n  default = 1\n}\n')
    (tmp_path / "sub" / "main.tf").write_text('variable "b" {\n  default=1\n}\n')
    clear_caches()
    rep = check_tree(str(tmp_path))
    assert rep.fmt.checked == ["main.tf", "sub/main.tf"]
    assert rep.fmt.changed == ["sub/main.tf"]
    assert rep.to_dict()["counts"]["fmt_differing"] == 1
    assert rep.bytes_scanned == sum(os.path.getsize(tmp_path / p) for p in ("main.tf", "sub/main.tf"))


def test_overlapping_fmt_targets_name_each_file_once(tmp_path, monkeypatch):
    (tmp_path / "sub").mkdir()
    (tmp_path / "sub" / "main.tf").write_text("a=1\n")
    monkeypatch.chdir(tmp_path)
    res = run_fmt([".", "sub", "sub/main.tf"], recursive=True, check=True)
    assert res.checked == ["sub/main.tf"] and res.changed == ["sub/main.tf"]


# ---------------------------------------------------------- unreadable / non-UTF-8 files
def test_a_non_utf8_file_is_one_diagnostic_not_a_traceback(tmp_path):
    (tmp_path / "ok.tf").write_text('variable "a" {}\n')
    (tmp_path / "bad.tf").write_bytes(b'# caf\xe9\nvariable "b" {}\n')
    clear_caches()
    res = validate_module(str(tmp_path), str(tmp_path))
    assert [(d.severity, d.summary, d.file, d.line, d.col) for d in res.diags] == [
        (ERROR, "Invalid character encoding", "bad.tf", 1, 6)]
    assert res.module.broken_files == {"bad.tf"}
    rep = check_tree(str(tmp_path))
    assert not rep.ok and rep.fmt.checked == ["bad.tf", "ok.tf"] and rep.fmt.changed == []


@pytest.mark.skipif(not hasattr(os, "symlink"), reason="needs symlinks")
def test_a_dangling_symlink_is_one_diagnostic_not_a_traceback(tmp_path, monkeypatch):
    (tmp_path / "ok.tf").write_text('variable "a" {}\n')
    os.symlink(str(tmp_path / "gone.tf"), str(tmp_path / "link.tf"))
    clear_caches()
    res = validate_module(str(tmp_path), str(tmp_path))
    assert [(d.summary, d.file) for d in res.diags] == [("Failed to read file", "link.tf")]
    monkeypatch.chdir(tmp_path)
    fmt = run_fmt(["."], check=True)
    assert fmt.exit_code == 2 and [d.summary for d in fmt.diags] == ["Failed to read file"]
