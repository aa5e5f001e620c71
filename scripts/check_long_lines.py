"""Lines of tens to hundreds of kilobytes with literals placed across the 32 KiB tile boundaries, against the oracle
(events, scores, spans). One-off check for the device path."""
import sys, random
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from log_parser_b200 import binding
from log_parser_b200.model import PatternSpec, ScoringConfig
from oracle import oracle
rng = random.Random(3)
pats = [PatternSpec(id=f"p{i}", primary_regex=r, confidence=0.5, severity="LOW", lines_before=1, lines_after=1)
        for i, r in enumerate(["ERRx", r"\bWARN\b", r"timeout \d+ms$", "^start", "(?<=a)Exception", r"\d{8,}ms", r"x.*y$", r"\s+at\s"])]
def filler(n): return "".join(rng.choice("abc de.fg-h1 2") for _ in range(n))
lines = []
for L in (70_000, 200_000, 10, 0, 32768 * 2, 40_000, 5):
    s = list(filler(L))
    for pos in (0, 32764, 32766, 32767, 32768, 65534, L - 6, L // 2):
        if 0 <= pos and pos + 12 < L:
            w = rng.choice(["ERRx", " WARN ", "timeout 12ms", "aException", " 123456789ms ", "start", " at "])
            s[pos:pos + len(w)] = list(w)
    lines.append("".join(s))
lines[1] = "start" + lines[1][5:-12] + "timeout 99ms"
log = ("\n".join(lines) + "\n").encode()
lp = binding.LogParser(ScoringConfig()); lp.load_patterns(pats)
for variant in ("host", "host2"):
    res = lp.analyze(log, options=binding.OPT_FACTORS | binding.OPT_SPANS)
    ora = oracle.analyze(ScoringConfig(), pats, log)
    ok = np.array_equal(res.line, ora.line) and np.array_equal(res.pattern, ora.pattern) and np.array_equal(res.score, ora.score)
    print(variant, "events", res.n_events, len(ora.line), "OK" if ok else "MISMATCH")
    if not ok:
        print(sorted(set(zip(res.line.tolist(), res.pattern.tolist())) ^ set(zip(ora.line.tolist(), ora.pattern.tolist())))[:10])
rx = [oracle.Regex(p.primary_regex) for p in pats]
badspan = 0
for k in range(res.n_events):
    t = lines[int(res.line[k]) - 1]
    s, e = rx[int(res.pattern[k])].span(t)
    if (int(res.span_start[k]), int(res.span_end[k])) != (s, e): badspan += 1
print("span mismatches", badspan)
