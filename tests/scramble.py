"""White-space damage for ``fmt`` recovery tests.

The only ground truth for ``terraform fmt`` that the reference holds is the set of
files its authors left in canonical form (33 of 38; e.g. eks/variables.tf,
gke/outputs.tf, eks/examples/cnpack/aws-pca.tf).  A fixed-point test ("formatting
them changes nothing") is satisfied by a formatter that does nothing at all.  The
stronger statement is: ruin the layout of such a file WITHOUT touching a single
token, and the formatter must give back the authors' bytes exactly.

Two kinds of damage, both driven by the scanner so that nothing inside a heredoc
body, a quoted template or a comment's own text is ever altered:

``damage_lines``   what a careless editor does -- every line's leading indent
                   replaced by a random mix of spaces and tabs, 0-3 blanks either
                   side of every ``=``, the gap before a trailing ``#``/``//``
                   comment collapsed or stretched, blanks appended to code lines;
``damage_tokens``  every gap between two tokens outside templates is re-drawn:
                   0-4 blanks where the tokens were adjacent and may stay so,
                   1-4 where they were separated (so no two tokens fuse).

Neither function knows anything about the canonical layout; they only destroy.
"""
from __future__ import annotations

import random

from nvidia_terraform_modules_b200.hcl.scanner import scan
from nvidia_terraform_modules_b200.hcl.tokens import T, Token, render

# a gap is part of a template's literal text when it follows one of these ...
_TEMPLATE_BEFORE = {T.OQUOTE, T.QUOTED_LIT, T.STRING_LIT, T.OHEREDOC, T.CHEREDOC, T.TEMPLATE_SEQ_END}
# ... or precedes one of these
_TEMPLATE_AFTER = {T.CQUOTE, T.QUOTED_LIT, T.STRING_LIT, T.CHEREDOC}


def _free_gaps(tokens: list[Token]) -> list[int]:
    """Indices of tokens whose preceding white space is layout, not content."""
    free = []
    depth_stack: list[str] = []       # "q" inside quotes/heredoc, "i" inside ${ } / %{ }
    prev = None
    for i, tok in enumerate(tokens):
        in_literal = bool(depth_stack) and depth_stack[-1] == "q"
        ok = (tok.type is not T.EOF and not in_literal
              and (prev is None or prev.type not in _TEMPLATE_BEFORE)
              and tok.type not in _TEMPLATE_AFTER)
        if ok:
            free.append(i)
        if tok.type in (T.OQUOTE, T.OHEREDOC):
            depth_stack.append("q")
        elif tok.type in (T.CQUOTE, T.CHEREDOC):
            depth_stack.pop()
        elif tok.type in (T.TEMPLATE_INTERP, T.TEMPLATE_CONTROL):
            depth_stack.append("i")
        elif tok.type is T.TEMPLATE_SEQ_END:
            depth_stack.pop()
        prev = tok
    return free


def _line_starts(tokens: list[Token]) -> set[int]:
    starts, at_start = set(), True
    for i, tok in enumerate(tokens):
        if at_start:
            starts.add(i)
        at_start = tok.is_line_end
    return starts


def damage_lines(src: str, rng: random.Random) -> str:
    tokens = scan(src)
    free = set(_free_gaps(tokens))
    starts = _line_starts(tokens)
    for i, tok in enumerate(tokens):
        if i not in free:
            continue
        nxt = tokens[i + 1] if i + 1 < len(tokens) else None
        if i in starts:
            if tok.type is T.NEWLINE:
                tok.raw_before = rng.choice(["", "", " ", "\t", "   "])          # blank line
            else:
                tok.raw_before = "".join(rng.choice(" \t") for _ in range(rng.randint(0, 7)))
        elif tok.type is T.EQUAL:
            tok.raw_before = " " * rng.randint(0, 3)
            if nxt is not None and (i + 1) in free and nxt.type is not T.NEWLINE:
                nxt.raw_before = " " * rng.randint(0, 3)
        elif tok.type is T.COMMENT and tok.text.endswith("\n"):
            tok.raw_before = " " * rng.choice([0, 1, 1, 2, 9])                   # trailing comment
        elif tok.type is T.NEWLINE:
            tok.raw_before = rng.choice(["", "", " ", "  ", "\t"])               # trailing blanks
    return render(tokens, keep_raw=True)


def damage_tokens(src: str, rng: random.Random) -> str:
    tokens = scan(src)
    for i in _free_gaps(tokens):
        tok = tokens[i]
        if tok.raw_before == "":
            tok.raw_before = " " * rng.choice([0, 0, 0, 1, 2, 4])
        else:
            tok.raw_before = " " * rng.randint(1, 4)
    return render(tokens, keep_raw=True)


def token_texts(src: str) -> list[tuple[str, str]]:
    return [(t.type.name, t.text) for t in scan(src)]


def recovery_counts(src: str, name: str, seeds: int, base_seed: int = 0) -> dict:
    """How many of ``seeds`` damaged copies of ``src`` format back to ``src``."""
    from nvidia_terraform_modules_b200.fmt.command import format_source
    out = {"seeds": seeds, "lines_recovered": 0, "tokens_recovered": 0, "damaged_differs": 0}
    want = token_texts(src)
    for s in range(seeds):
        for kind, fn in (("lines", damage_lines), ("tokens", damage_tokens)):
            rng = random.Random(f"{name}/{kind}/{base_seed + s}")
            bad = fn(src, rng)
            assert token_texts(bad) == want, f"{name}: the damage changed a token ({kind}, seed {s})"
            out["damaged_differs"] += bad != src
            text, diags = format_source(bad, name)
            if not diags and text == src:
                out[f"{kind}_recovered"] += 1
    return out
