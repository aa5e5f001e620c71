"""CPU-only differential fuzz with deeper random regexes glued to literal words: whole-line DFA, anchored
automata, boundary-form decision, bytecode engine and spans against the oracle. Hunting tool.
Usage: python scripts/fuzz_host_deep.py <seed> [rounds]"""
import random
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from log_parser_b200 import binding  # noqa: E402
from oracle import oracle  # noqa: E402
from rxgen import Gen  # noqa: E402

ALPHA = "abcERx_1 .-:\r\u00e9"
WORDS = ["ERRx", "abc_1", "Exception", "x-1:E", "timeout "]


def main():
    seed = int(sys.argv[1]) if len(sys.argv) > 1 else 21
    rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 2500
    rng = random.Random(seed)
    bad = n = 0
    for it in range(rounds):
        a, _ = Gen(rng, regular_only=(it % 4 != 0), max_depth=3).regex()
        w = rng.choice(WORDS)
        wj = "".join("\\" + c if c in ".-:" else c for c in w)
        j = [a, wj + a, a + wj, "(?:" + a + ")" + wj + "\\b", "\\b" + wj + "(?:" + a + ")?", "^" + wj + a, a + wj + "$"][rng.randrange(7)]
        try:
            o = oracle.Regex(j)
            d = binding.DebugRegex(j)
        except Exception:
            continue
        for _k in range(6):
            t = "".join(rng.choice([rng.choice(WORDS), "".join(rng.choice(ALPHA) for _ in range(rng.randint(0, 8)))])
                        for _ in range(rng.randint(0, 4)))
            try:
                m = d.match(t)
            except Exception:
                continue
            if m < 0:
                continue
            e = o.find(t)
            n += 1
            if bool(m) != e:
                bad += 1
                print("BAD", "dfa" if d.is_dfa else "vm", repr(j), repr(t), m, e)
            if d.anchor and bool(d.match_anchored(t)) != e:
                bad += 1
                print("BAD-ANCH", d.anchor, repr(j), repr(t), e)
            r = d.match_local(t)
            if r != -2 and bool(r) != e:
                bad += 1
                print("BAD-LOCAL", repr(j), repr(t), r, e)
            try:
                sp = d.span(t)
            except Exception:
                continue
            so = o.span(t)
            if so is not None:
                so = (len(t[:so[0]].encode()), len(t[:so[1]].encode()))
            if sp != so:
                bad += 1
                print("BAD-SPAN", repr(j), repr(t), sp, so)
        if bad > 6:
            break
    print("seed", seed, "checked", n, "bad", bad)


if __name__ == "__main__":
    main()
