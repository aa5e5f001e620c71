"""CPU-only differential fuzz of the host compiler (DFA, anchored automata, bytecode engine, spans) against the
oracle on texts with non-ASCII characters and Java's extra line terminators. Hunting tool; the fixed-seed
version lives in tests/test_library_cpu.py."""
import random
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from log_parser_b200 import binding  # noqa: E402
from oracle import oracle  # noqa: E402
from rxgen import Gen  # noqa: E402

ALPHA_U = "abcERx_1 .-:\r\u00e9\u20ac\u0085\u2028"


def gt(rng, maxlen=16):
    return "".join(rng.choice(ALPHA_U) for _ in range(rng.randint(0, maxlen)))


def main():
    rng = random.Random(int(sys.argv[1]) if len(sys.argv) > 1 else 99)
    bad = n = 0
    for it in range(int(sys.argv[2]) if len(sys.argv) > 2 else 4000):
        j, _ = Gen(rng, regular_only=(it % 3 != 0)).regex()
        try:
            o = oracle.Regex(j)
            d = binding.DebugRegex(j)
        except Exception:
            continue
        for _k in range(8):
            t = gt(rng)
            try:
                m = d.match(t)
            except Exception:
                continue
            if m < 0:
                continue
            n += 1
            e = o.find(t)
            if bool(m) != e:
                bad += 1
                if bad < 8:
                    print("BAD", "dfa" if d.is_dfa else "vm", repr(j), repr(t), m, e)
            if d.anchor and bool(d.match_anchored(t)) != e:
                bad += 1
                if bad < 8:
                    print("BAD-ANCH", repr(j), repr(t), e)
            try:
                sp = d.span(t)
            except Exception:
                continue
            so = o.span(t)
            if so is not None:
                so = (len(t[:so[0]].encode()), len(t[:so[1]].encode()))
            if sp != so:
                bad += 1
                if bad < 8:
                    print("BAD-SPAN", repr(j), repr(t), sp, so)
    print("checked", n, "bad", bad)


if __name__ == "__main__":
    main()
