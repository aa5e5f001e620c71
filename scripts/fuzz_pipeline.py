"""Wider differential fuzz of the device pipeline (events and spans) against the oracle: random regex fragments
glued around literal words. One-off hunting tool; the fixed-seed version lives in tests/test_gpu_parity.py."""
import sys, random
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from log_parser_b200 import binding
from log_parser_b200.model import PatternSpec, ScoringConfig
from oracle import oracle
from rxgen import Gen
parser = binding.LogParser(ScoringConfig())
words = ["ERRx", "abc_1", "Exception", "timeout", "x-1:E", "R.E.c", "WARN", "é_ab", "1234ms"]
alphabet = "abcERx_1 .-:\r\té"
bad = 0
for seed in range(int(sys.argv[1]) if len(sys.argv) > 1 else 100, int(sys.argv[2]) if len(sys.argv) > 2 else 112):
    rng = random.Random(seed)
    for rnd_i in range(20):
        pats, rxs = [], []
        tries = 0
        while len(pats) < 16 and tries < 400:
            tries += 1
            a, _ = Gen(rng, regular_only=(len(pats) % 3 != 0), max_depth=2).regex()
            b, _ = Gen(rng, regular_only=True, max_depth=1).regex()
            w = rng.choice(words)
            wj = "".join("\\" + ch if ch in ".-:" else ch for ch in w)
            shape = rng.randrange(7)
            j = [wj + a, a + wj, a + wj + b, "\\b" + wj + "\\b", "(?:" + a + "|" + wj + b + ")", "(?i)" + wj + a, "^" + a + wj + "$"][shape]
            try:
                o = oracle.Regex(j); binding.DebugRegex(j)
            except Exception:
                continue
            pats.append(PatternSpec(id=f"p{len(pats)}", primary_regex=j, confidence=0.5, severity="LOW")); rxs.append(o)
        lines = []
        for _ in range(300):
            parts = []
            for _k in range(rng.randint(0, 4)):
                parts.append(rng.choice(words) if rng.random() < 0.5 else "".join(rng.choice(alphabet) for _ in range(rng.randint(0, 6))))
            lines.append("".join(parts).replace("\n", ""))
        log = ("\n".join(lines) + "\n").encode()
        try:
            parser.load_patterns(pats)
        except binding.LpError as e:
            continue
        try:
            res = parser.analyze(log, options=binding.OPT_NO_RANK | binding.OPT_SPANS)
        except binding.LpError as e:
            print("analyze error", seed, rnd_i, e); continue
        got = set(zip(res.line.tolist(), res.pattern.tolist()))
        # java split of our text: lines may contain \r at end -> "\r\n" handling: recompute via oracle split semantics
        text_lines = log.decode().split("\n")[:-1]
        exp = set()
        for i, ln in enumerate(text_lines):
            if ln.endswith("\r"): ln = ln[:-1]      # \r\n separator
            if i >= res.n_lines: break
            for p, o in enumerate(rxs):
                if o.find(ln): exp.add((i + 1, p))
        if got != exp:
            bad += 1
            d = sorted(got ^ exp)[:3]
            print("MISMATCH seed", seed, rnd_i, d, [(pats[p].primary_regex, text_lines[l-1]) for l, p in d])
        else:
            # spans
            for k in range(0, res.n_events, 5):
                ln = text_lines[int(res.line[k]) - 1]
                if ln.endswith("\r"): ln = ln[:-1]
                s, e = rxs[int(res.pattern[k])].span(ln)
                ex = (len(ln[:s].encode()), len(ln[:e].encode()))
                if (int(res.span_start[k]), int(res.span_end[k])) != ex:
                    bad += 1; print("SPAN MISMATCH", seed, pats[int(res.pattern[k])].primary_regex, repr(ln), ex, int(res.span_start[k]), int(res.span_end[k])); break
print("done, bad =", bad)
