"""Pins the oracle against the known-answer vectors of SURVEY.md §8c.

The reference holds no tests, fixtures or golden vectors (src/test does not exist), so these are
hand-derived from ScoringService.java:123-151,177-178, ContextAnalysisService.java:62-106,
FrequencyTrackingService.java:69-72 and the one correct worked number of
docs/SCORING_ALGORITHM.md:107-114.
"""
import math
import struct

import pytest

from log_parser_b200.model import ScoringConfig


def hx(x: float) -> str:
    return struct.pack(">d", x).hex()


CFG = ScoringConfig()


@pytest.mark.parametrize("i,n,exp,exphex", [
    (0, 10, 2.5, "4004000000000000"),
    (1, 10, 2.0, "4000000000000000"),
    (2, 10, 1.5, "3ff8000000000000"),
    (3, 10, 1.3333333333333335, "3ff5555555555556"),
    (5, 10, 1.0, "3ff0000000000000"),
    (9, 10, 0.6, "3fe3333333333333"),
    (1499, 10000, 1.7505000000000002, "3ffc020c49ba5e36"),
    (2001, 10000, 1.4998333333333334, "3ff7ff513cc1e099"),
    (999999, 1000000, 0.500001, "3fe0000218def417"),
])
def test_chrono_vectors(oracle, i, n, exp, exphex):
    got = oracle.chrono(CFG, i, n)
    assert got == exp
    assert hx(got) == exphex


def test_proximity_vectors(oracle):
    # DOC:107-114: 1 + 0.8*e^(-0.5)
    assert 1.0 + oracle.proximity_term(CFG, 0.8, 5) == 1.4852245277701068
    assert hx(1.0 + oracle.proximity_term(CFG, 0.8, 5)) == "3ff7c37acb5fd59e"
    assert 1.0 + oracle.proximity_term(CFG, 0.6, 3) == 1.4444909324090307


@pytest.mark.parametrize("log,lines", [
    (b"", [b""]),
    (b"\n\n", []),
    (b"\na", [b"", b"a"]),
    (b"a\r\nb\n\n", [b"a", b"b"]),
    (b"a\rb", [b"a\rb"]),
    (b"a\n\r", [b"a", b"\r"]),
    (b"a\n\r\n\n", [b"a"]),
    (b"a\n \n", [b"a", b" "]),
    (b"\r\r\n", [b"\r"]),
    (b"x", [b"x"]),
    (b"x\n", [b"x"]),
    (b"\n", []),
    (b"\r\n", []),
    (b"a\n\nb", [b"a", b"", b"b"]),
])
def test_split_vectors(oracle, log, lines):
    assert [log[a:b] for a, b in oracle.split(log)] == lines


def test_context_vectors(oracle):
    # 3 lines each ERR+EXC, then 1 STACK line: score 2.3000000000000003 -> factor capped to 2.5
    lines = [b"ERROR FooException"] * 3 + [b"\tat a.b.C.m(File.java:1)"]
    assert oracle.context_factor(CFG, lines) == 2.5
    big = ScoringConfig(max_context_factor=100.0)
    assert oracle.context_factor(big, lines) == 3.3000000000000003
    # ERR wins over WARN on the same line (else-if, CAS:64-70)
    assert oracle.context_factor(big, [b"WARN and ERROR"]) == 1.4
    assert oracle.context_factor(big, [b"WARN only"]) == 1.2
    assert oracle.context_factor(big, [b"plain"]) == 1.0
    assert oracle.context_factor(big, []) == 1.0
    # case-insensitive ERR/WARN, case-sensitive EXC
    assert oracle.context_factor(big, [b"fatal"]) == 1.4
    assert oracle.context_factor(big, [b"warning"]) == 1.2
    assert oracle.context_factor(big, [b"fooexception"]) == 1.0
    assert oracle.context_factor(big, [b"OutOfMemoryError"]) == 1.3


def test_context_density_rule(oracle):
    big = ScoringConfig(max_context_factor=1000.0)
    # total=11, stack+err=8: 8 > 7.699999999999999 -> x0.8
    lines = [b"ERROR"] * 8 + [b"x"] * 3
    score = 0.0
    for _ in range(8):
        score += 0.4
    assert oracle.context_factor(big, lines) == 1.0 + score * 0.8
    # total=20, 14: 14 > 14.0 false
    lines = [b"ERROR"] * 14 + [b"x"] * 6
    score = 0.0
    for _ in range(14):
        score += 0.4
    assert oracle.context_factor(big, lines) == 1.0 + score
    # total=10: rule off
    lines = [b"ERROR"] * 10
    score = 0.0
    for _ in range(10):
        score += 0.4
    assert oracle.context_factor(big, lines) == 1.0 + score


def test_stack_regex(oracle):
    big = ScoringConfig(max_context_factor=1000.0)
    # STACK line: +0.1 per line, then +min(n*0.1, 0.5)
    assert oracle.context_factor(big, [b"   at com.foo.Bar$1.run(Bar.java:12)  "]) == 1.0 + (0.1 + 0.1)
    assert oracle.context_factor(big, [b"at x()"]) == 1.0 + (0.1 + 0.1)
    assert oracle.context_factor(big, [b"at x() trailing"]) == 1.0
    assert oracle.context_factor(big, [b"xat x()"]) == 1.0


def test_frequency_vectors(oracle):
    assert oracle.frequency_penalty(CFG, False, 1000.0) == 0.0          # FTS:69-72
    assert oracle.frequency_penalty(CFG, True, 5.0) == 0.0              # DOC:202
    assert oracle.frequency_penalty(CFG, True, 10.0) == 0.0             # <= threshold
    assert oracle.frequency_penalty(CFG, True, 15.0) == 0.5
    assert oracle.frequency_penalty(CFG, True, 100.0) == 0.8            # capped


def test_score_product_order(oracle):
    """SS:102-109: left-to-right product; one event, all factors visible."""
    from log_parser_b200.model import PatternSpec, SecondaryPattern, SequencePattern

    log = b"\n".join([b"start", b"connect begin", b"WARN slow", b"boom OutOfMemoryError", b"retry", b"x", b"y", b"z", b"w", b"v"])
    p = PatternSpec(id="p1", primary_regex="OutOfMemoryError", confidence=0.8, severity="high",
                    lines_before=2, lines_after=1,
                    secondaries=[SecondaryPattern("retry", 0.8, 10), SecondaryPattern("nomatch", 0.5, 10)],
                    sequences=[SequencePattern(["start", "connect", "boom"], 0.5), SequencePattern(["zzz", "boom"], 0.9)])
    r = oracle.analyze(CFG, [p], log)
    assert r.n_lines == 10 and list(r.line) == [4] and list(r.pattern) == [0]
    base, sev, ch, px, tp, cf, pen = r.factors[0]
    assert (base, sev) == (0.8, 3.0)
    assert ch == oracle.chrono(CFG, 3, 10)
    assert px == 1.0 + 0.8 * math.exp(-1 / 10.0)
    assert tp == 1.5
    # context: lines 2..5 -> "connect begin"(none) "WARN slow"(+0.2) "boom OutOfMemoryError"(+0.3) "retry"
    s = 0.0
    s += 0.2
    s += 0.3
    assert cf == 1.0 + s
    assert pen == 0.0
    assert r.score[0] == base * sev * ch * px * tp * cf * (1.0 - pen)
    assert list(r.freq_delta) == [1]


def test_frequency_sequence(oracle):
    """Read-before-increment (SS:84-88): k-th match of a pattern sees prior + k."""
    from log_parser_b200.model import PatternSpec

    log = b"\n".join([b"hit"] * 30)
    p = PatternSpec(id="p", primary_regex="hit", confidence=1.0, severity="INFO")
    blank = PatternSpec(id="  ", primary_regex="hit", confidence=1.0, severity="INFO")
    r = oracle.analyze(CFG, [p, blank], log)
    pens = r.factors[r.pattern == 0][:, 6]
    for k, pen in enumerate(pens):
        rate = float(k)
        exp = 0.0 if (k == 0 or rate <= 10.0) else min(0.8, (rate - 10.0) / 10.0)
        assert pen == exp
    assert (r.factors[r.pattern == 1][:, 6] == 0.0).all()   # blank id: never penalised (FTS:65-67)
    r2 = oracle.analyze(CFG, [p, blank], log, prior_count=[12, 0])
    assert r2.factors[r2.pattern == 0][0, 6] == (12.0 - 10.0) / 10.0
    assert list(r2.freq_delta) == [30, 30]


def test_discovery_order(oracle):
    from log_parser_b200.model import PatternSpec

    log = b"ab\nb\na\nab"
    ps = [PatternSpec(id="a", primary_regex="a", confidence=1.0), PatternSpec(id="b", primary_regex="b", confidence=1.0)]
    r = oracle.analyze(CFG, ps, log)
    assert list(zip(r.line, r.pattern)) == [(1, 0), (1, 1), (2, 1), (3, 0), (4, 0), (4, 1)]


def test_temporal_equivalence_bruteforce(oracle):
    """backward greedy walk (SS:237-261) == forward threshold T < i."""
    import random

    from log_parser_b200.model import PatternSpec, SequencePattern

    rng = random.Random(7)
    for _ in range(300):
        n = rng.randint(1, 14)
        k = rng.randint(1, 4)
        ev = ["e%d" % j for j in range(k)]
        lines = []
        for _l in range(n):
            toks = [e for e in ev if rng.random() < 0.3]
            if rng.random() < 0.5:
                toks.append("P")
            lines.append(" ".join(toks))
        log = "\n".join(lines).encode()
        p = PatternSpec(id="p", primary_regex="P", confidence=1.0, sequences=[SequencePattern(ev, 1.0)])
        r = oracle.analyze(CFG, [p], log)
        nl = r.n_lines
        has = [[ev[j] in lines[l].split() for l in range(nl)] for j in range(k)]
        # forward threshold
        T = -1
        ok_chain = True
        for j in range(k - 1):
            nxt = [l for l in range(nl) if l > T and has[j][l]]
            if not nxt:
                ok_chain = False
                break
            T = nxt[0]
        for idx, ln in enumerate(r.line):
            i = int(ln) - 1
            near = any(has[k - 1][l] for l in range(max(0, i - 5), min(nl, i + 6)))
            exp = near and (k == 1 or (ok_chain and T < i))
            assert r.factors[idx, 4] == (2.0 if exp else 1.0)


def _pin_corpus():
    import base64
    import gzip
    import os

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "regex_pin_corpus.tsv.gz")
    with gzip.open(path, "rt", encoding="ascii") as f:
        for row in f:
            if row.strip():
                p, flags, t, found, s, e = row.rstrip("\n").split("\t")
                yield base64.b64decode(p).decode("utf-8"), int(flags), base64.b64decode(t).decode("utf-8"), found == "1", int(s), int(e)


def test_pin_corpus_is_what_the_oracle_and_the_library_answer(oracle):
    """tests/golden/regex_pin_corpus.tsv.gz is the file oracle/RegexPin.java replays on a real JDK (none in this image):
    ~5 000 (pattern, flags, text) cases with the first find() and its span. The restatement must keep giving the recorded
    answers (otherwise the corpus is regenerated and the change explained), and the product's host compiler — automata or
    bytecode engine, the tables the device walks — gives the same ones."""
    from log_parser_b200 import binding

    n = n_found = n_lib = 0
    compiled = {}
    for pat, flags, text, found, s, e in _pin_corpus():
        key = (pat, flags)
        if key not in compiled:
            try:
                d = binding.DebugRegex(pat, flags)
            except binding.UnsupportedRegexError:
                d = None                               # beyond the library's limits (e.g. look-behind spanning > 64 lengths)
            compiled[key] = (oracle.Regex(pat, flags), d)
        o, d = compiled[key]
        sp = o.span(text)
        assert (sp is not None) == found and (sp or (-1, -1)) == (s, e), (pat, flags, text)
        n += 1
        n_found += found
        if d is None:
            continue
        got = d.match(text)
        if got < 0:
            continue                                   # backtracking budget
        assert bool(got) == found, (pat, flags, text)
        n_lib += 1
        if found:
            try:
                assert tuple(d.span(text)) == (len(text[:s].encode()), len(text[:e].encode())), (pat, flags, text)
            except binding.LpError:
                pass
    assert n > 4500 and n_found > 1200 and n_lib > 4400
