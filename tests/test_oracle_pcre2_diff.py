"""Third engine of SURVEY §8c: the oracle's java.util.regex restatement (and the product's host compiler)
against PCRE2 10.42 through ctypes, on generated patterns that CPython cannot parse (possessive and
atomic constructs inside look-behind, lazy counted loops around look-around, ...).

PCRE2 is not Java. Three classes of patterns are known to differ BY DESIGN and are kept out of the strict
comparison; each is pinned separately by tests/golden/regex_dialect_cases.json (the answer the JDK's
Pattern node semantics give, which the oracle and the product follow, next to PCRE2's):

  backref    the pattern holds a back-reference. java.util.regex leaves a group's capture in place when the
             match that set it is abandoned inside an independent sub-match (atomic group, look-around,
             possessive Curly: GroupTail only restores when its continuation fails, and there the
             continuation is `accept`), and Matcher clears groups once per find(), not per start position.
             PCRE2 restores captures on every backtrack.
  poss-group `(X){n,m}+`: Java's possessive Curly matches X atomically per iteration; PCRE2 treats the
             whole loop as one atomic group and may backtrack inside an earlier iteration.
  empty-iter a quantified group that can match the empty string: Java's Loop.match stops iterating as soon
             as an iteration consumed nothing, even below the minimum count; PCRE2 keeps iterating.
"""
import json
import os
import random
import re

import pytest

import pcre2
from rxgen import Gen, gen_text

pytestmark = pytest.mark.skipif(not pcre2.available, reason="libpcre2-8 not present")
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "regex_dialect_cases.json")


def quantified_groups(pat: str):
    """(body, quantifier) of every group directly followed by a quantifier other than a bare `?`."""
    out, stack, i, in_cls = [], [], 0, 0
    while i < len(pat):
        ch = pat[i]
        if ch == "\\":
            i += 2
            continue
        if in_cls:
            if ch == "[":
                in_cls += 1
            elif ch == "]":
                in_cls -= 1
        elif ch == "[":
            in_cls = 1
            if pat[i + 1:i + 2] == "^":
                i += 1
            if pat[i + 1:i + 2] == "]":
                i += 1
        elif ch == "(":
            stack.append(i)
        elif ch == ")":
            start = stack.pop()
            m = re.match(r"(\*|\+|\{\d+(,\d*)?\})([?+]?)", pat[i + 1:])
            if m:
                out.append((pat[start:i + 1], m.group(0)))
        i += 1
    return out


def dialect_class(pat: str):
    if re.search(r"\\[1-9]|\\k<", pat):
        return "backref"
    for body, q in quantified_groups(pat):
        if q.endswith("+") and len(q) > 1:
            return "poss-group"
        try:
            if pcre2.Regex(body).can_match_empty():
                return "empty-iter"
        except pcre2.Pcre2Error:
            return "empty-iter"
    return None


def test_quantified_group_scanner():
    assert quantified_groups(r"(a|b)+x(?:c)(d){2,3}?[(]+\(e\)*") == [("(a|b)", "+"), ("(d)", "{2,3}?")]
    assert dialect_class(r"(\W{1,3}){2}+") == "poss-group"
    assert dialect_class(r"(\.{2,}|\D|^){2,}") == "empty-iter"
    assert dialect_class(r"(a)\1") == "backref"
    assert dialect_class(r"(?>a+)b(?<=ab)c*+") is None


def test_random_differential_vs_pcre2(oracle):
    """Existence and span of the first find() agree on every generated pattern outside the three classes."""
    from log_parser_b200 import binding, build

    build.build()
    rng = random.Random(20261107)
    n_checked = n_pat = n_nonregular = 0
    for it in range(6000):
        j, _p = Gen(rng, max_depth=3, no_backrefs=True, no_group_possessive=True).regex()
        if dialect_class(j) is not None:
            continue
        try:
            p = pcre2.Regex(j)
        except pcre2.Pcre2Error:
            continue        # e.g. look-behind that PCRE2 10.42 cannot bound
        o = oracle.Regex(j)
        try:
            d = binding.DebugRegex(j)
        except binding.UnsupportedRegexError:
            d = None
        n_pat += 1
        n_nonregular += bool(re.search(r"\(\?[>=!<]|[*+?}]\+", j))
        for _ in range(8):
            t = gen_text(rng)
            try:
                exp = p.search(t)
            except pcre2.Pcre2Error:
                continue    # match limit
            got = o.span(t)
            assert (None if got is None else tuple(got)) == exp, f"java={j!r} text={t!r} oracle={got} pcre2={exp}"
            if d is not None:
                assert bool(d.match(t)) == (exp is not None), f"java={j!r} text={t!r} product={d.match(t)} pcre2={exp}"
            n_checked += 1
    assert n_checked > 25000 and n_nonregular > 800, (n_checked, n_pat, n_nonregular)


@pytest.mark.parametrize("pat,text,exp", [
    # Java-only syntax that CPython re/regex reject and PCRE2 parses: both engines must give the same answer
    (r"\Qa.b\E+", "a.bbb", (0, 5)), (r"(?<w>a+)-\k<w>", "aa-aa", (0, 5)), (r"\x{41}\x{1F600}", "A\U0001F600", (0, 2)),
    (r"\e\cA", "\x1b\x01", (0, 2)), (r"(?<=a|bc)d", "bcd", (2, 3)), (r"(?<!a|bc)d", "bcd", None), (r"a++b", "aaab", (0, 4)),
    (r"a*+a", "aaa", None), (r"(?>a|ab)c", "abc", None), (r"(?:a|ab)c", "abc", (0, 3)), (r"x*+\b", "xx y", (0, 2)),
    (r"\h+\v", "a \t\rb", (1, 4)), (r"a\R", "a\r\nb", (0, 3)), (r"a$", "a\r\n", (0, 1)), (r"a\Z", "a\u0085", (0, 1)),
    (r"a\z", "a\n", None), (r".+", "ab\rcd", (0, 2)), (r"(?s).+", "ab\rcd", (0, 5)), (r"(?i)ERROR", "an error", (3, 8)),
    (r"(?x) a b # comment", "ab", (0, 2)), (r"a{2,3}+a", "aaaa", (0, 4)), (r"a{2,3}+a", "aaa", None), (r"(?=(a+))a*b", "aaab", (0, 4)),
])
def test_java_syntax_cpython_cannot_parse(oracle, pat, text, exp):
    got = oracle.Regex(pat).span(text)
    assert (None if got is None else tuple(got)) == exp
    assert pcre2.Regex(pat).search(text) == exp


def test_dialect_difference_fixtures(oracle):
    """The committed cases where java.util.regex (per the JDK's node semantics; the oracle and the product follow
    it) and PCRE2 part ways: both recorded answers are reproduced, so a change on either side is noticed."""
    from log_parser_b200 import binding, build

    build.build()
    cases = json.load(open(GOLDEN))["cases"]
    assert {c["class"] for c in cases} == {"backref", "poss-group", "empty-iter"}
    for c in cases:
        exp_java = None if c["java"] is None else tuple(c["java"])
        exp_pcre = None if c["pcre2"] is None else tuple(c["pcre2"])
        got = oracle.Regex(c["pattern"]).span(c["text"])
        assert (None if got is None else tuple(got)) == exp_java, c
        assert pcre2.Regex(c["pattern"]).search(c["text"]) == exp_pcre, c
        assert bool(binding.DebugRegex(c["pattern"]).match(c["text"])) == (exp_java is not None), c
        assert dialect_class(c["pattern"]) == c["class"], c
