"""CPU-only differential fuzz of the host compiler against the oracle on MULTI-SCRIPT text: Unicode property classes,
negated classes that leave out whole UTF-8 lead-byte blocks, (?U) classes and word boundaries, inside look-arounds,
atomic groups, possessive repeats and back-references, through the automata and through the bytecode engine (forced for
every other pattern). Texts mix 1-, 2-, 3- and 4-byte characters and combining marks. Hunting tool; a fixed seed of the same
generator runs in tests/test_library_cpu.py.
Usage: python scripts/fuzz_host_scripts.py [first_seed] [last_seed] [patterns_per_seed]"""
import random
import sys

sys.path.insert(0, ".")
from log_parser_b200 import binding  # noqa: E402
from oracle import oracle  # noqa: E402

ALPHA = "aé́ΩдЖ٣_ -.x1‍€漢\U0001F600Z"
TERMINATORS = "\r\u0085\u2028"              # what a line of a "\n"-split log can still hold and `$`, `.`, (?m) care about
ATOMS = [r"\p{L}", r"\P{L}", r"\p{Lu}", r"\P{Lu}", r"\p{IsCyrillic}", r"\P{IsCyrillic}", r"\p{InGreek}", r"\P{InGreek}", r"[^Ѐ-ӿ]",
         r"\p{Nd}", r"\P{Nd}", r"\p{M}", r"\P{M}", r"\p{IsAlphabetic}", r"\P{IsAlphabetic}", r"\p{So}", r"\P{So}", r"\p{IsHan}", r"[^\p{IsHan}a]",
         r"\w", r"\W", r"\d", r"\D", r"\s", r"\S", ".", "a", "x", "é", "д", "Ж", "漢", "Жx1", "éa", "xд", "1漢", "\U0001F600", r"[д-я]", r"[^a-zд]",
         r"[\p{L}&&[^\p{IsLatin}]]", r"\b", r"\B", "^", "$"]
QUANT = ["", "", "", "+", "*", "?", "{2}", "+?", "*?", "*+", "++", "{1,2}"]


def piece(rng: random.Random, depth: int) -> str:
    k = rng.randrange(12)
    if depth <= 0 or k < 6:
        a = rng.choice(ATOMS)
        return a + (rng.choice(QUANT) if a not in ("^", "$", r"\b", r"\B") else "")
    inner = "".join(piece(rng, depth - 1) for _ in range(rng.randint(1, 2)))
    if k == 6:
        return "(?=" + inner + ")"
    if k == 7:
        return "(?!" + inner + ")"
    if k == 8:
        one = rng.choice([a for a in ATOMS if a not in ("^", "$", r"\b", r"\B")])
        return rng.choice(["(?<=", "(?<!"]) + one + rng.choice(["", "{2}", "{1,2}"]) + ")"
    if k == 9:
        return "(?>" + inner + ")" + rng.choice(["", "+"])
    if k == 10:
        return "(?:" + inner + "|" + piece(rng, depth - 1) + ")" + rng.choice(QUANT)
    return "(" + inner + ")" + rng.choice(["", r"\1", r".*\1"])


def pattern(rng: random.Random) -> str:
    parts = [piece(rng, 2) for _ in range(rng.randint(1, 3))]
    if rng.random() < 0.25:                   # a flag switched on or off part of the way, or for one group only
        k = rng.randrange(len(parts))
        parts[k] = rng.choice(["(?i)", "(?-i)", "(?u)", "(?m)", "(?s)", "(?d)", "(?U)", "(?-U)", "(?x) ", "(?iu)"]) + parts[k] if rng.random() < 0.5 \
            else "(?" + rng.choice(["i", "iu", "s", "m", "U", "-i", "d", "x"]) + ":" + parts[k] + ")"
    return rng.choice(["", "", "(?U)", "(?i)", "(?iu)", "(?U)(?i)", "(?m)", "(?s)", "(?d)", "(?dm)"]) + "".join(parts)


def text(rng: random.Random, maxlen: int = 12) -> str:
    t = "".join(rng.choice(ALPHA) for _ in range(rng.randint(0, maxlen)))
    k = rng.random()
    if k < 0.15:
        t += rng.choice(TERMINATORS)
    elif k < 0.25:
        cut = rng.randint(0, len(t))
        t = t[:cut] + rng.choice(TERMINATORS) + t[cut:]
    return t


def check(rng: random.Random, n_patterns: int, report=print):
    checked = bad = matched = 0
    done = 0
    while done < n_patterns:
        j = pattern(rng)
        force = rng.random() < 0.5
        try:
            o = oracle.Regex(j)
            d = binding.DebugRegex(j, force_vm=force)
        except (oracle.OracleSyntaxError, oracle.OracleUnsupported, binding.UnsupportedRegexError, binding.PatternSyntaxError):
            continue
        done += 1
        for _k in range(8):
            t = text(rng)
            got = d.match(t)
            if got < 0:
                continue
            checked += 1
            want = o.find(t)
            if bool(got) != want:
                bad += 1
                report("FIND MISMATCH", "dfa" if d.is_dfa else "vm", repr(j), repr(t), "library", got, "oracle", want)
                continue
            if d.anchor and bool(d.match_anchored(t)) != want:
                bad += 1
                report("ANCHORED MISMATCH", repr(j), repr(t), "oracle", want)
            local = d.match_local(t)                       # -2: not of the "literal with boundary conditions" form
            if local != -2 and bool(local) != want:
                bad += 1
                report("BOUNDARY-FORM MISMATCH", repr(j), repr(t), "library", local, "oracle", want)
            if want:
                matched += 1
                try:
                    sp = d.span(t)
                except binding.LpError:
                    continue
                cs, ce = o.span(t)
                if sp != (len(t[:cs].encode()), len(t[:ce].encode())):
                    bad += 1
                    report("SPAN MISMATCH", repr(j), repr(t), sp, (cs, ce))
    return checked, matched, bad


def main():
    lo = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    hi = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    per = int(sys.argv[3]) if len(sys.argv) > 3 else 1500
    n = bad = 0
    for seed in range(lo, hi):
        c, _m, b = check(random.Random(seed), per)
        n += c
        bad += b
    print("checked", n, "bad", bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
