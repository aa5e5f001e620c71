"""Self-consistency fuzz on the GPU: a batch of small logs (many without a final newline, cut in the middle of
words so that literals straddle log boundaries) must give, log by log, what the same logs give one at a time.
No oracle involved, so it is fast. Usage: python scripts/fuzz_batch_consistency.py [first_seed] [end_seed]"""
import random
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import numpy as np  # noqa: E402
from log_parser_b200 import binding  # noqa: E402
from log_parser_b200.model import PatternSpec, ScoringConfig  # noqa: E402
from rxgen import Gen  # noqa: E402

WORDS = ["ERRx", "abc_1", "Exception", "timeout", "x-1:E", "R.E.c", "WARN", "1234ms"]
ALPHA = "abcERx_1 .-:\r"


def main():
    lo = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    hi = int(sys.argv[2]) if len(sys.argv) > 2 else 41
    parser = binding.LogParser(ScoringConfig())
    bad = rounds = 0
    for seed in range(lo, hi):
        rng = random.Random(seed)
        pats = []
        tries = 0
        while len(pats) < 16 and tries < 400:
            tries += 1
            a, _ = Gen(rng, regular_only=(len(pats) % 3 != 0), max_depth=2).regex()
            w = rng.choice(WORDS)
            wj = "".join("\\" + ch if ch in ".-:" else ch for ch in w)
            j = [wj + a, a + wj, "\\b" + wj + "\\b", "(?:" + a + "|" + wj + ")", "^" + wj, wj + "$", wj][rng.randrange(7)]
            try:
                binding.DebugRegex(j)
            except Exception:
                continue
            pats.append(PatternSpec(id=f"p{len(pats) % 5}", primary_regex=j, confidence=0.5, severity="LOW"))
        try:
            parser.load_patterns(pats)
        except binding.LpError:
            continue
        text = []
        for _ in range(400):
            parts = [rng.choice(WORDS) if rng.random() < 0.6 else "".join(rng.choice(ALPHA) for _ in range(rng.randint(0, 6)))
                     for _k in range(rng.randint(0, 4))]
            text.append("".join(parts))
        blob = ("\n".join(text) + "\n").encode()
        cuts = sorted(rng.sample(range(1, len(blob)), 25))          # anywhere: mid-word, mid-line
        logs = [blob[a:b] for a, b in zip([0] + cuts, cuts + [len(blob)])]
        try:
            whole = parser.analyze_batch(logs)
            prior = np.zeros(len(pats), np.uint64)
            for k, lg in enumerate(logs):
                one = parser.analyze(lg, prior_count=prior)
                b, e = int(whole.log_event_begin[k]), int(whole.log_event_begin[k + 1])
                ok = (int(whole.log_n_lines[k]) == one.n_lines and np.array_equal(whole.line[b:e], one.line) and
                      np.array_equal(whole.pattern[b:e], one.pattern) and np.array_equal(whole.score[b:e], one.score) and
                      np.array_equal(whole.rank[b:e], one.rank))
                if not ok:
                    bad += 1
                    print("MISMATCH seed", seed, "log", k, repr(lg[-12:]), repr(logs[k + 1][:12]) if k + 1 < len(logs) else "")
                    break
                # FrequencyTrackingService keys its counts by pattern id: every pattern of an id sees the id's total
                delta = np.asarray(one.freq_delta, np.uint64)
                for ident in set(q.id for q in pats):
                    idx = [i for i, q in enumerate(pats) if q.id == ident]
                    prior[idx] += delta[idx].sum()
        except binding.LpError as ex:
            print("error", seed, ex)
            continue
        rounds += 1
    print("rounds", rounds, "bad", bad)


if __name__ == "__main__":
    main()
