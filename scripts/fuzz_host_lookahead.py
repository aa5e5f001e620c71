"""CPU-only differential fuzz of the deferred leading look-arounds (vm.cpp build_vm, VM_LOOK_BEGIN / VM_ABORT_MARK in
vm_exec.h) against the oracle: random look-aheads with an unbounded walk inside, glued in front of random rests with
and without captures, back-references, atomic groups and possessive repeats. Hunting tool; the fixed seed of the same
generator lives in tests/test_library_cpu.py.
Usage: python scripts/fuzz_host_lookahead.py [first_seed] [last_seed] [checks_per_seed]"""
import random
import re
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from log_parser_b200 import binding  # noqa: E402
from oracle import oracle  # noqa: E402
from rxgen import Gen  # noqa: E402

ALPHA = "abcERx_1 .-:\ré"
WALKS = [".*", r"\w+", ".+?", "[^E]*", "(?:ab|c)*", r".*?\b", r"\S*", ".{2,}"]


def text(rng: random.Random, maxlen: int) -> str:
    return "".join(rng.choice(ALPHA) for _ in range(rng.randint(0, maxlen)))


def one_seed(seed: int, want: int):
    rng = random.Random(seed)
    checked = bad = 0
    while checked < want:
        def sub(regular, depth=2):
            return Gen(rng, regular_only=regular, max_depth=depth).regex()[0]
        cond = re.sub(r"\((?!\?)", "(?:", sub(True))
        walk = rng.choice(WALKS)
        look = rng.choice(["(?=", "(?!"]) + rng.choice([walk + cond, cond + walk, walk + cond + walk]) + ")"
        for _ in range(rng.choice([0, 0, 1, 2])):
            look += rng.choice(["(?=", "(?!"]) + rng.choice(WALKS) + re.sub(r"\((?!\?)", "(?:", sub(True)) + ")"
        head = rng.choice(["", "", "^", r"\b", "(?i)"])
        j = head + look + sub(rng.random() < 0.5, rng.choice([1, 2, 3]))
        try:
            o = oracle.Regex(j)
            d = binding.DebugRegex(j)
        except (oracle.OracleSyntaxError, oracle.OracleUnsupported, binding.UnsupportedRegexError, binding.PatternSyntaxError):
            continue
        for _k in range(8):
            t = text(rng, rng.choice([8, 24, 60]))
            got = d.match(t)
            if got < 0:
                continue
            checked += 1
            want_found = o.find(t)
            if bool(got) != want_found:
                bad += 1
                print("FIND MISMATCH", repr(j), repr(t), "library", got, "oracle", want_found)
                continue
            if got:
                cs, ce = o.span(t)
                if tuple(d.span(t)) != (len(t[:cs].encode()), len(t[:ce].encode())):
                    bad += 1
                    print("SPAN MISMATCH", repr(j), repr(t), tuple(d.span(t)), (cs, ce))
    return checked, bad


def main():
    lo = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    hi = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    want = int(sys.argv[3]) if len(sys.argv) > 3 else 4000
    n = bad = 0
    for seed in range(lo, hi):
        c, b = one_seed(seed, want)
        n += c
        bad += b
    print("checked", n, "bad", bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
