"""CPU-only tier: the C-ABI library loads and exports every symbol of include/logparser_b200.h,
fails loudly without a GPU, and its host-side regex compiler (the tables the device walks) agrees
with the oracle. No compute entry point is exercised here."""
import os
import random
import re

import pytest

from log_parser_b200 import binding
from log_parser_b200.model import PatternSpec, ScoringConfig
from rxgen import Gen, gen_text

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from log_parser_b200 import build

    build.build()
    return binding.load_library()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "logparser_b200.h")).read()
    declared = set(re.findall(r"\b(lp_[a-z_0-9]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(binding.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in the header but not exported"


def test_debug_hooks_are_not_in_the_product_library(lib):
    assert not hasattr(lib, "lp_debug_rx_compile") and not hasattr(lib, "lp_debug_plan_dump")
    assert hasattr(binding.load_debug_library(), "lp_debug_rx_compile")


def test_build_reuse_is_decided_by_content_not_by_file_times():
    """build.stale(): the prebuilt library is reused exactly when the hash stored beside it equals the hash of the sources,
    headers and compiler flags (a copy of the tree on a GPU box has meaningless file times)."""
    from log_parser_b200 import build

    build.build()
    assert not build.stale()
    assert build.source_hash() == open(build.SO + ".srchash").read().strip()
    assert build.source_hash(flags=build.NVCC_FLAGS + ["-DX"]) != build.source_hash()
    assert build.source_hash(sources=build.SOURCES[:-1]) != build.source_hash()
    os.utime(os.path.join(build.CSRC, "engine.cu"))          # touching a source does not make the library stale
    assert not build.stale()


def test_version_and_error_string(lib):
    assert b"sm_100a" in lib.lp_version()
    assert isinstance(lib.lp_last_error(), bytes)


def test_no_cpu_fallback_without_gpu(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(binding.LpError) as ei:
        binding.LogParser(ScoringConfig())
    assert ei.value.code == binding.LP_ERR_CUDA


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "log_parser_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import oracle" not in src and "from oracle" not in src and "oracle/" not in src, f


def test_dfa_tables_vs_oracle_random(lib, oracle):
    rng = random.Random(4242)
    checked = n_dfa = 0
    for _ in range(1500):
        j, _p = Gen(rng, regular_only=True).regex()
        try:
            o = oracle.Regex(j)
        except (oracle.OracleSyntaxError, oracle.OracleUnsupported):
            continue
        d = binding.DebugRegex(j)
        n_dfa += d.is_dfa      # a few land on the bytecode path (assertion-guarded empty loops)
        for _k in range(10):
            t = gen_text(rng)
            assert bool(d.match(t)) == o.find(t), (j, t)
            checked += 1
    assert checked > 10000 and n_dfa > 1200


def test_anchored_verification_equals_line_walk(lib, oracle):
    """Regexes whose literals cover the start (or the end) of every match are decided in k_confirm at
    the literal occurrences by an anchored automaton instead of a walk of the whole line. The host
    mirrors of both walks must agree, and agree with the oracle. Literal length is lowered to 1-2 bytes
    here so that the short random patterns exercise the cover test and both directions."""
    rng = random.Random(777)
    n_fwd = n_rev = checked = 0
    for it in range(2500):
        j, _p = Gen(rng, regular_only=True).regex()
        try:
            o = oracle.Regex(j)
        except (oracle.OracleSyntaxError, oracle.OracleUnsupported):
            continue
        d = binding.DebugRegex(j, min_literal=1 + it % 3)
        if not d.is_dfa or d.anchor == 0:
            continue
        n_fwd += d.anchor == 1
        n_rev += d.anchor == 2
        for _k in range(12):
            t = gen_text(rng)
            exp = o.find(t)
            assert bool(d.match(t)) == exp, (j, t)
            assert bool(d.match_anchored(t)) == exp, (j, t, d.anchor, d.literals())
            checked += 1
    assert n_fwd > 150 and n_rev > 40 and checked > 3000, (n_fwd, n_rev, checked)
    # the shapes of the synthetic pattern set and of CAS:27-34
    cases = [(r"\b\d{8,}ms\b", 2), (r"\b[45]\d\d \d+ms\b", 1), (r"retrying in \d+ms", 1), (r"\.\.\. \d+ more$", 0),
             (r"\.\.\. \d+ more", 2), (r"(?i)timeout after \d+s$", 1), (r"caused by: \S+Error\b", 1), (r"^ab+cdef$", 0)]
    texts = ["took 123456789ms", "x12345678ms", "12345678ms", "1234567ms!", "HTTP 503 12ms", "a503 12ms", "503 12msx", "retrying in 5ms",
             "retrying in ms", "\t... 12 more", "... more", "Timeout After 30s", "timeout after 30s\r", "timeout after 30sx",
             "caused by: java.lang.OutOfMemoryError", "caused by: Errors", "abbbcdef", "xabcdef", ""]
    for pat, mode in cases:
        d = binding.DebugRegex(pat)
        assert d.anchor == mode, (pat, d.anchor)
        o = oracle.Regex(pat)
        for t in texts:
            assert bool(d.match(t)) == o.find(t), (pat, t)
            if mode:
                assert bool(d.match_anchored(t)) == o.find(t), (pat, t)


def test_non_ascii_texts_vs_oracle_random(lib, oracle):
    """The automata work on UTF-8 bytes, java.util.regex on UTF-16 units: on texts with two- and three-byte
    characters and Java's extra line terminators the DFA, the anchored automata, the bytecode engine and the
    spans must still agree with the oracle. (A pattern that can match the empty string next to `\\B` once
    matched between the two bytes of a character: the search loop now advances by whole code points.)"""
    j = "(?i)ERRxx\\-*|((1{0,2}a)x[_]|\\S|(a*?\\.?:).)\\-{2}?|((a{2}R{2,})?|(:{2}?)){2,}\\B"
    assert binding.DebugRegex(j).match("x\u00e9E E") == 0 and not oracle.Regex(j).find("x\u00e9E E")
    alphabet = "abcERx_1 .-:\r\u00e9\u20ac\u0085\u2028"
    rng = random.Random(99)
    checked = 0
    for it in range(1200):
        rx, _p = Gen(rng, regular_only=(it % 3 != 0)).regex()
        try:
            o = oracle.Regex(rx)
            d = binding.DebugRegex(rx)
        except Exception:
            continue
        for _k in range(8):
            t = "".join(rng.choice(alphabet) for _ in range(rng.randint(0, 16)))
            m = d.match(t)
            if m < 0:
                continue
            e = o.find(t)
            assert bool(m) == e, (rx, t)
            if d.anchor:
                assert bool(d.match_anchored(t)) == e, (rx, t, d.anchor)
            try:
                sp = d.span(t)
            except binding.LpError:
                continue
            so = o.span(t)
            if so is not None:
                so = (len(t[:so[0]].encode("utf-8")), len(t[:so[1]].encode("utf-8")))
            assert sp == so, (rx, t)
            checked += 1
    assert checked > 6000


def test_boundary_form_decision_vs_oracle_random(lib, oracle):
    """Regexes of the shape [\\b][\\w*]LITERAL[\\b] are decided on the device by looking at the bytes next to the
    literal occurrence. The host mirror of that decision against the oracle, on patterns built to look
    like that shape while hiding assertions, groups and alternations inside (`word(?: \\A)` once slipped
    through: nested zero-width nodes must disqualify the form)."""
    assert binding.DebugRegex("1234ms(?i: \\A)").match_local("x1234ms ") == -2
    assert binding.DebugRegex("\\b(WARN|WARNING)\\b", 2).match_local("a warn b") == 1
    rng = random.Random(5)
    words = ["ERRx", "abc_1", "Exception", "x-1:E"]
    checked = 0
    for _ in range(2500):
        a, _p = Gen(rng, regular_only=True, max_depth=2).regex()
        w = rng.choice(words)
        wj = "".join("\\" + c if c in ".-:" else c for c in w)
        j = [wj + a, a + wj, "\\b" + wj + "\\b", "\\b\\w*" + wj + "\\b", "(?:" + a + "|" + wj + ")", wj + "(?:" + a + ")"][rng.randrange(6)]
        try:
            o = oracle.Regex(j)
            d = binding.DebugRegex(j)
        except Exception:
            continue
        for _k in range(6):
            t = "".join(rng.choice([rng.choice(words), gen_text(rng, 5)]) for _ in range(rng.randint(0, 3)))
            r = d.match_local(t)
            if r == -2:
                break
            assert bool(r) == o.find(t), (j, t)
            checked += 1
    assert checked > 3000


def test_match_spans_vs_oracle_random(lib, oracle):
    """Matcher.start()/end() of the first find() from the bytecode engine (what k_spans runs) against the
    oracle: leftmost start, end by java.util.regex's priorities (ordered alternation, greedy/lazy/possessive)."""
    rng = random.Random(31337)
    checked = with_match = 0
    for it in range(900):
        j, _p = Gen(rng, regular_only=(it % 2 == 0)).regex()
        try:
            o = oracle.Regex(j)
        except (oracle.OracleSyntaxError, oracle.OracleUnsupported):
            continue
        try:
            d = binding.DebugRegex(j)
        except binding.UnsupportedRegexError:
            continue
        for _k in range(8):
            t = gen_text(rng)
            exp = o.span(t)
            if exp is not None:      # oracle offsets are code points
                exp = (len(t[:exp[0]].encode("utf-8")), len(t[:exp[1]].encode("utf-8")))
                with_match += 1
            try:
                got = d.span(t)
            except binding.LpError:
                continue
            assert got == exp, (j, t)
            checked += 1
    assert checked > 5000 and with_match > 1500


def test_backtracking_bytecode_vs_oracle_random(lib, oracle):
    """The interpreter of csrc/vm_exec.h (same code on host and device) against the oracle, on
    patterns with look-around, back-references, possessive/atomic constructs — and on regular ones
    forced through the bytecode path."""
    rng = random.Random(77)
    checked = budget = 0
    for _ in range(900):
        j, _p = Gen(rng, max_depth=2).regex()
        try:
            o = oracle.Regex(j)
        except (oracle.OracleSyntaxError, oracle.OracleUnsupported):
            continue
        try:
            d = binding.DebugRegex(j, force_vm=True)
        except binding.UnsupportedRegexError:
            continue
        assert not d.is_dfa
        for _k in range(8):
            t = gen_text(rng, maxlen=16)
            got = d.match(t)
            if got < 0:
                budget += 1      # bounded backtracking: reported as LP_ERR_BACKTRACK by the device path
                continue
            assert bool(got) == o.find(t), (j, t)
            checked += 1
    assert checked > 5000 and budget < checked // 100


def test_deferred_leading_lookaround_vs_oracle_random(lib, oracle):
    """build_vm tests a leading look-ahead with an unbounded walk inside after the rest of the pattern has matched
    (vm.cpp); the oracle evaluates it where java.util.regex does, first. Same find() result, same span; also with
    captures and back-references in the rest, and with the look-ahead's own failure (VM_ABORT_MARK)."""
    rng = random.Random(4)
    checked = matched = 0
    while checked < 6000:
        def sub(regular):
            return Gen(rng, regular_only=regular, max_depth=2).regex()[0]
        cond = re.sub(r"\((?!\?)", "(?:", sub(True))
        walk = rng.choice([".*", r"\w+", ".+?", "[^E]*", "(?:ab|c)*"])
        look = rng.choice(["(?=", "(?!"]) + rng.choice([walk + cond, cond + walk, walk + cond + walk]) + ")"
        if rng.random() < 0.3:
            look += rng.choice(["(?=", "(?!"]) + walk + re.sub(r"\((?!\?)", "(?:", sub(True)) + ")"
        j = look + sub(rng.random() < 0.5)
        try:
            o = oracle.Regex(j)
            d = binding.DebugRegex(j)
        except (oracle.OracleSyntaxError, oracle.OracleUnsupported, binding.UnsupportedRegexError, binding.PatternSyntaxError):
            continue
        for _k in range(8):
            t = gen_text(rng, maxlen=24)
            got = d.match(t)
            if got < 0:
                continue
            assert bool(got) == o.find(t), (j, t)
            checked += 1
            if got:
                matched += 1
                cs, ce = o.span(t)
                assert tuple(d.span(t)) == (len(t[:cs].encode()), len(t[:ce].encode())), (j, t)
    assert matched > 300


@pytest.mark.parametrize("pat,text", [
    (r"(?=.*\[tag1\])(\w+) lost \1", "ab lost ab [tag1]"), (r"(?=.*\[tag1\])(\w+) lost \1", "ab lost ab [tag2]"),
    (r"(?!.*skip)(?=.*x)a(b|c)d", "acd x"), (r"(?!.*skip)(?=.*x)a(b|c)d", "acd x skip"), (r"(?=.*z)(?:a|ab)(?:c|bcd)", "abcd z"),
    (r"(?<![\w.])FooError(?!\w)", "x FooError: y"), (r"(?<![\w.])FooError(?!\w)", "x.FooError"),
    (r"Foo: (\w+) .*\b\1\b", "Foo: abc failed for abc again"), (r"Foo: (\w+) .*\b\1\b", "Foo: abc failed for abd again"),
    (r"\bFoo(?>: \w++)", "a Foo: bar"), (r"(?m)^a", "b\ra"), (r"(?m)a$", "a\rb"), (r"(?m)^$", ""),
    (r"x.{0,40}y", "x" + "a" * 30 + "y"), (r"x.{0,40}y", "x" + "a" * 41 + "y"), (r"(?=.*\[edge1\])lost \d+", "lost 12 [edge1]"),
    (r"(?i)(a)\1", "aA"), (r"(?<=é)x", "éx"), (r"(?<=\bab|c)x", "abx"), (r"^(?:a|ab)*+c", "abac"), (r"(a*)*b", "aaaa"),
    (r"(?:^_??[^x]*){2}a", "  a:.Ec_11E"), (r"(\w+)\s\1$", "hello hello\r"), (r"(?<=\d{2,4})ms", "x1234ms"),
    # a class that leaves out a whole block of UTF-8 lead bytes: the search for the next start position must not stop on the
    # continuation byte of a character whose lead byte no match begins with (first_set in vm.cpp)
    (r"(?!\u0434)\P{L}", "\u0434\u0434"), (r"(?!\u0434)\P{L}", "\u0434\u0434-"), (r"(?=.)[^\u0400-\u04ff]", "\u0436\u0436"), (r"(?>\P{IsCyrillic})x", "\u0436x"),
    # (?U)\b / \B: CharPredicates.WORD() on both sides, a combining mark is a word character
    (r"(?U)\b\u00e9", "a\u00e9 \u00e9"), (r"(?U)\b\u00e9", "a\u00e9"), (r"(?U)a\b", "a\u0301"), (r"(?U)a\B", "a\u0301"), (r"(?U)\B\W", "\u0434-"), (r"(?U)\B", ""),
    (r"(?U)\b\d+\b", "x \u0663\u0664 y"), (r"(?U)\bx\b", "\u0434x\u0434"), (r"\bx\b", "\u0434x\u0434"),
])
def test_backtracking_cases(lib, oracle, pat, text):
    d = binding.DebugRegex(pat)
    assert bool(d.match(text)) == oracle.Regex(pat).find(text)


def test_multi_script_text_vs_oracle_random(lib, oracle):
    """Unicode property classes, negated classes that leave out whole UTF-8 lead-byte blocks, (?U) classes and word
    boundaries inside look-arounds, atomic groups, possessive repeats and back-references, on texts mixing 1- to 4-byte
    characters and combining marks; automata and bytecode engine (forced for every other pattern). The generator is
    scripts/fuzz_host_scripts.py, here with a fixed seed."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("fuzz_host_scripts", os.path.join(ROOT, "scripts", "fuzz_host_scripts.py"))
    fz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fz)
    problems = []
    checked, matched, bad = fz.check(random.Random(7), 600, report=lambda *a: problems.append(a))
    assert bad == 0, problems[:5]
    assert checked > 4000 and matched > 800


@pytest.mark.parametrize("pat,flags", [
    (r"\b(ERROR|FATAL|CRITICAL|SEVERE)\b", 2), (r"\b(WARN|WARNING)\b", 2),
    (r"^\s*at\s+[\w\.\$]+\(.*\)\s*$", 0), (r"\b\w*Exception\b|\b\w*Error\b", 0),
])
def test_context_regex_tables(lib, oracle, pat, flags):
    rng = random.Random(3)
    words = ["ERROR", "error", "Fatal", "CRITICAL", "severe", "WARN", "warning", "Warnings", "xERROR", "ERRORS", "at", "\tat",
             "  at", "com.foo.Bar$1.run(Bar.java:12)", "NullPointerException", "OutOfMemoryError", "Errors", "exception",
             "MyError:", "(x)", "()", " ", "INFO", "foo", "-", "\r", "é", " ", "\x85"]
    d = binding.DebugRegex(pat, flags)
    o = oracle.Regex(pat, flags)
    for _ in range(3000):
        t = " ".join(rng.choice(words) for _ in range(rng.randint(0, 5)))
        if rng.random() < 0.3:
            t = rng.choice(["\tat ", "   at ", "at "]) + t
        if rng.random() < 0.1:
            t += rng.choice(["\r", " ", "\x85", "\r\r"])
        assert bool(d.match(t)) == o.find(t), (pat, t)


@pytest.mark.parametrize("pat,text", [
    ("a$", "a\r"), ("a$", "a "), ("a$", "a\x85"), ("a$", "a\r\r"), ("a\\z", "a\r"), ("a$\\r", "a\r"), ("^.$", "é"),
    ("^..$", "\U0001F600"), ("[^a]", "é"), ("é+$", "ééé"), ("[à-ÿ]x", "éx"), ("\\bé", "é"), ("x\\B", "xé"), (".", "\r"),
    ("(?s).", "\r"), ("(?d).", "\r"), ("(?d)a$", "a\r"), ("\\R", "\r"), ("^$", ""), ("\\B", ""), ("\\b", ""), ("a{3,}", "aa"),
    ("(a|b)*?c$", "abcab"), ("x*", ""), ("\\p{Alpha}+\\d", "ab1"), ("\\Qa.b\\E", "axb"), ("[\\w&&[^\\d]]+$", "ab1"),
])
def test_dfa_java_specifics(lib, oracle, pat, text):
    assert bool(binding.DebugRegex(pat).match(text)) == oracle.Regex(pat).find(text)


def test_syntax_errors_surface_like_java(lib):
    for bad in ["a{,2}", "(a", "[a", "*a", "a{3,2}", "\\k<x>", "(?<=a*)b"]:
        with pytest.raises(binding.PatternSyntaxError):
            binding.DebugRegex(bad)
    for ok in ["\\G", "\\p{L}", "(?u)a", "(?U)\\w", "\\p{IsGreek}", "(?U)\\bx", "(?U)x\\B"]:        # legal Java the reference serves: accepted since round 2
        binding.DebugRegex(ok)
    for unsup in ["\\X", "\\b{g}", "\\N{DIGIT ONE}", "(?iu)(a)\\1"]:
        with pytest.raises(binding.UnsupportedRegexError):
            binding.DebugRegex(unsup)


def test_required_literals_are_sound(lib, oracle):
    """Every matching text contains one of the extracted literals."""
    from log_parser_b200 import synth

    rng = random.Random(9)
    wl = synth.make_patterns(200, 77)
    for p, render in zip(wl.patterns, wl.renderers):
        try:
            d = binding.DebugRegex(p.primary_regex)
        except binding.UnsupportedRegexError:
            continue
        lits = d.literals()
        o = oracle.Regex(p.primary_regex)
        for _ in range(5):
            t = render(rng)
            assert o.find(t), (p.primary_regex, t)
            if lits:
                tb = t.encode()
                assert any((l.lower() in tb.lower()) if ic else (l in tb) for ic, l in lits), (p.primary_regex, t, lits)


def test_scan_group_product_dfa_vs_oracle(lib, oracle):
    """The always-on scan walks product automata of several regexes (k_dfa_scan). The serialised
    tables, including the idle-state fast path, are checked here on the host against the oracle."""
    import ctypes as C

    lib = binding.load_debug_library()
    groups = [
        ([r"\b(ERROR|FATAL|CRITICAL|SEVERE)\b", r"\b(WARN|WARNING)\b", r"\b\w*Exception\b|\b\w*Error\b", r"\b\d{8,}ms\b"], [2, 2, 0, 0]),
        ([r"\b[45]\d\d \d+ \d+\.\d{3}$", r"^\s*at\s+[\w\.\$]+\(.*\)\s*$"], [0, 0]),
        ([r"^\s*at\s+[\w\.\$]+\(.*\)\s*$", r"^\s+\.\.\. \d+ more$", r"^$", r"x$"], [0, 0, 0, 0]),
        ([r"a.*b", r"\Bc\B", r"[^a-z]+$"], [0, 0, 0]),
    ]
    rng = random.Random(21)
    words = ["ERROR", "error", "Fatal", "WARN", "warning", "xERROR", "at", "\tat", "com.foo.Bar$1.run(Bar.java:12)",
             "NullPointerException", "OutOfMemoryError", "Errors", "12345678ms", "123456789ms", "x1234567ms", "503 12 0.123",
             "404 1 22.5", "... 12 more", "\t... 3 more", " ", "é", "\r", "abc", "a", "b", "c", "x", "ms", "_", "-"]
    for regs, flags in groups:
        arr = (C.c_char_p * len(regs))(*[r.encode() for r in regs])
        fl = (C.c_int * len(regs))(*flags)
        oracles = [oracle.Regex(r, f) for r, f in zip(regs, flags)]
        for _ in range(250):     # the hook rebuilds the group for every call
            t = " ".join(rng.choice(words) for _ in range(rng.randint(0, 6)))
            if rng.random() < 0.2:
                t = rng.choice(["\tat ", "   at ", "\t"]) + t
            if rng.random() < 0.1:
                t += rng.choice(["\r", " ", "\x85"])
            tb = t.encode()
            got = lib.lp_debug_group_match(arr, fl, len(regs), tb, len(tb))
            assert got < 0xFFFFFFF0, (regs, t, hex(got))
            exp = sum((1 << k) for k, o in enumerate(oracles) if o.find(t))
            assert got == exp, (regs, t, bin(got), bin(exp))


def test_scan_group_first_byte_skip_set(oracle):
    """k_dfa_scan does not walk a non-empty line whose first byte decides every regex of its groups without a hit
    (the stack-trace regex of CAS:31 dies on anything but white space or 'a'): for every byte of the skip set, no
    line that starts with it may match any regex of the group."""
    import ctypes as C
    import random

    lib = binding.load_debug_library()
    groups = [[r"^\s*at\s+[\w\.\$]+\(.*\)\s*$"], [r"^\s*at\s+[\w\.\$]+\(.*\)\s*$", r"^\d+ errors?$", r"^Caused by"],
              [r"^\s*at\s+[\w\.\$]+\(.*\)\s*$", r"x\d"], [r"^$|^a"], [r"\Aabc\z", r"^ab*$"]]
    rng = random.Random(8)
    for regs in groups:
        arr = (C.c_char_p * len(regs))(*[r.encode() for r in regs])
        fl = (C.c_int * len(regs))(*([0] * len(regs)))
        skip = (C.c_uint32 * 8)()
        assert lib.lp_debug_group_skip(arr, fl, len(regs), skip) == 0
        in_set = [b for b in range(256) if (skip[b >> 5] >> (b & 31)) & 1]
        assert 13 not in in_set and 0xC2 not in in_set and 0xE2 not in in_set
        if regs == groups[0]:
            assert ord("2") in in_set and ord("C") in in_set and ord("\t") not in in_set and ord("a") not in in_set and ord(" ") not in in_set
        if regs == groups[2]:
            assert not in_set                      # an unanchored member keeps every line alive
        ors = [oracle.Regex(r) for r in regs]
        for b in in_set:
            if b >= 0x80:
                continue                            # keep the texts valid UTF-8
            for _ in range(6):
                t = chr(b) + "".join(rng.choice("at com.x.Y(Z.java:1) 123 errors\tCaused by\rabc$") for _ in range(rng.randint(0, 30)))
                tb = t.encode()
                assert lib.lp_debug_group_match(arr, fl, len(regs), tb, len(tb)) == 0, (regs, t)
                assert not any(o.find(t) for o in ors), (regs, t)
