"""Scanner: token kinds, positions and loss-free round trip."""
import pytest

from nvidia_terraform_modules_b200.hcl.scanner import is_valid_identifier, scan
from nvidia_terraform_modules_b200.hcl.tokens import T, render


def kinds(src):
    return [t.type for t in scan(src)]


def test_empty_source_is_just_eof():
    assert kinds("") == [T.EOF]


def test_attribute_tokens_and_positions():
    toks = scan("name = \"x\"\n")
    assert [t.type for t in toks] == [T.IDENT, T.EQUAL, T.OQUOTE, T.QUOTED_LIT, T.CQUOTE, T.NEWLINE, T.EOF]
    assert (toks[1].line, toks[1].col) == (1, 6)
    assert toks[1].spaces_before == 1


def test_identifiers_may_contain_dashes_and_unicode():
    assert [t.text for t in scan("google-beta héllo _x9")[:3]] == ["google-beta", "héllo", "_x9"]
    assert is_valid_identifier("cnpack-log-group")
    assert not is_valid_identifier("9lives")
    assert not is_valid_identifier("")


@pytest.mark.parametrize("text", ["0", "12", "1.5", "1e9", "2.5E-3", "007"])
def test_number_literals(text):
    toks = scan(text)
    assert toks[0].type is T.NUMBER_LIT and toks[0].text == text


def test_legacy_index_is_dot_number_dot():
    assert kinds("a.0.b")[:5] == [T.IDENT, T.DOT, T.NUMBER_LIT, T.DOT, T.IDENT]


def test_operators():
    src = "== != <= >= && || => ... :: ? : ! < > + - * / %"
    want = [T.EQUAL_OP, T.NOT_EQUAL, T.LESS_THAN_EQ, T.GREATER_THAN_EQ, T.AND, T.OR, T.FAT_ARROW, T.ELLIPSIS,
            T.DOUBLE_COLON, T.QUESTION, T.COLON, T.BANG, T.LESS_THAN, T.GREATER_THAN, T.PLUS, T.MINUS, T.STAR,
            T.SLASH, T.PERCENT, T.EOF]
    assert kinds(src) == want


def test_line_comments_swallow_their_newline():
    toks = scan("# one\n// two\na = 1 # three\n")
    comments = [t for t in toks if t.type is T.COMMENT]
    assert [c.text for c in comments] == ["# one\n", "// two\n", "# three\n"]
    assert all(c.is_line_end for c in comments)


def test_block_comment_is_one_token_and_not_a_line_end():
    toks = scan("/* a\n b */\nx = 1\n")
    assert toks[0].type is T.COMMENT and toks[0].text == "/* a\n b */"
    assert not toks[0].is_line_end
    assert toks[1].type is T.NEWLINE


def test_template_interpolation_nests_normal_mode():
    toks = scan('"a${ b("c${d}") }e"')
    assert [t.type for t in toks] == [
        T.OQUOTE, T.QUOTED_LIT, T.TEMPLATE_INTERP, T.IDENT, T.OPAREN, T.OQUOTE, T.QUOTED_LIT,
        T.TEMPLATE_INTERP, T.IDENT, T.TEMPLATE_SEQ_END, T.CQUOTE, T.CPAREN, T.TEMPLATE_SEQ_END,
        T.QUOTED_LIT, T.CQUOTE, T.EOF]


def test_braces_inside_interpolation_do_not_end_it():
    toks = scan('"${ { a = 1 } }"')
    assert [t.type for t in toks].count(T.TEMPLATE_SEQ_END) == 1
    assert [t.type for t in toks].count(T.CBRACE) == 1


def test_escaped_template_introducers_stay_literal():
    toks = scan('"$${a} %%{b} \\"q\\""')
    assert [t.type for t in toks] == [T.OQUOTE, T.QUOTED_LIT, T.CQUOTE, T.EOF]


def test_strip_markers():
    toks = scan('"${~ a ~}"')
    assert toks[1].text == "${~" and toks[3].text == "~}"


def test_heredoc_tokens():
    toks = scan("x = <<-EOT\n  hi ${v}\n  EOT\n")
    assert [t.type for t in toks] == [T.IDENT, T.EQUAL, T.OHEREDOC, T.STRING_LIT, T.TEMPLATE_INTERP, T.IDENT,
                                      T.TEMPLATE_SEQ_END, T.STRING_LIT, T.CHEREDOC, T.NEWLINE, T.EOF]
    assert toks[2].text == "<<-EOT\n"
    assert toks[8].text == "  EOT"      # closing marker keeps its indentation


def test_heredoc_marker_must_stand_alone():
    toks = scan("x = <<EOT\nEOT not yet\nEOT\n")
    lits = [t.text for t in toks if t.type is T.STRING_LIT]
    assert lits == ["EOT not yet\n"]


def test_invalid_character_is_reported_as_token():
    assert T.INVALID in kinds("a = 1 @ 2")
    assert T.INVALID in kinds("a = ${b}")


def test_crlf_newlines():
    toks = scan("a = 1\r\nb = 2\r\n")
    assert [t.text for t in toks if t.type is T.NEWLINE] == ["\r\n", "\r\n"]


SAMPLES = [
    "",
    "a = 1",
    "a\t=\t1  \n\n\n",
    'resource "x" "y" {\n\tname   = "n" # c\n  tags = { a = 1, "b" : 2 }\n}\n',
    'x = <<EOT\n  raw ${ v } %{ if a }b%{ endif }\nEOT\n',
    "/* c */ a = [for k, v in m : k => v...]\n",
]


@pytest.mark.parametrize("src", SAMPLES)
def test_round_trip_is_byte_exact(src):
    assert render(scan(src), keep_raw=True) == src
