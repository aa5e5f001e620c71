"""Self-consistency fuzz on the GPU: one log cut by line range into 2-5 shards (halos, summaries, composition
in lp_range_score) must concatenate to what the whole log gives, with random regexes as primaries, secondaries
and sequence events. No oracle involved. Usage: python scripts/fuzz_range_consistency.py [first_seed] [end_seed]"""
import random
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import numpy as np  # noqa: E402
from log_parser_b200 import binding, parallel  # noqa: E402
from log_parser_b200.model import PatternSpec, ScoringConfig, SecondaryPattern, SequencePattern  # noqa: E402
from rxgen import Gen  # noqa: E402

WORDS = ["ERRx", "abc_1", "Exception", "timeout", "x-1:E", "R.E.c", "WARN", "1234ms"]
ALPHA = "abcERx_1 .-:\r"


def rx(rng):
    while True:
        a, _ = Gen(rng, regular_only=(rng.random() < 0.8), max_depth=2).regex()
        w = rng.choice(WORDS)
        wj = "".join("\\" + ch if ch in ".-:" else ch for ch in w)
        j = [wj + a, a + wj, "\\b" + wj + "\\b", "(?:" + a + "|" + wj + ")", wj][rng.randrange(5)]
        try:
            binding.DebugRegex(j)
            return j
        except Exception:
            continue


def main():
    lo = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    hi = int(sys.argv[2]) if len(sys.argv) > 2 else 31
    bad = rounds = 0
    for seed in range(lo, hi):
        rng = random.Random(seed)
        cfg = ScoringConfig(max_window=rng.choice([5, 40, 100]))
        pats = []
        for k in range(10):
            secs = [SecondaryPattern(regex=rx(rng), weight=rng.choice([0.3, 0.5, 0.8]), proximity_window=rng.choice([1, 7, 30, 200]))
                    for _ in range(rng.randint(0, 2))]
            seqs = [SequencePattern(events=[rx(rng) for _ in range(rng.randint(1, 4))], bonus_multiplier=rng.choice([0.2, 0.5]))
                    for _ in range(rng.randint(0, 2))]
            ctx = rng.random() < 0.6
            pats.append(PatternSpec(id=f"p{k % 4}", primary_regex=rx(rng), confidence=0.5 + 0.05 * k, severity=rng.choice(["LOW", "HIGH", "CRITICAL"]),
                                    lines_before=rng.randint(0, 30) if ctx else None, lines_after=rng.randint(0, 30) if ctx else None,
                                    secondaries=secs, sequences=seqs))
        lines = []
        for _ in range(rng.choice([300, 1500, 4000])):
            parts = [rng.choice(WORDS) if rng.random() < 0.4 else "".join(rng.choice(ALPHA) for _ in range(rng.randint(0, 6)))
                     for _k in range(rng.randint(0, 4))]
            lines.append("".join(parts))
        log = ("\n".join(lines) + rng.choice(["", "\n", "\n\n\r\n"])).encode()
        world = rng.choice([2, 3, 5])
        lps = [binding.LogParser(cfg) for _ in range(world + 1)]
        try:
            for lp in lps:
                lp.load_patterns(pats)
            prior = (np.arange(len(pats)) % 3).astype(np.uint64) * 4
            for i, q in enumerate(pats):          # one count per id
                prior[i] = prior[[j for j, r in enumerate(pats) if r.id == q.id][0]]
            whole = lps[world].analyze(log, prior_count=prior)
            end = parallel.strip_trailing_separators(log)
            if end == 0 and len(log) > 0:
                continue
            engines = [parallel.GpuRangeEngine(lp) for lp in lps[:world]]
            pieces = parallel.split_line_ranges(log, world, engines[0].halo_lines(), end)
            summaries = np.stack([engines[r].scan(bytes(log[tb:te]), hb, ha) for r, (tb, te, hb, ha) in enumerate(pieces)])
            results = [engines[r].score(summaries, r, world, prior) for r in range(world)]
            ok = (np.array_equal(np.concatenate([r.line for r in results]), whole.line) and
                  np.array_equal(np.concatenate([r.pattern for r in results]), whole.pattern) and
                  np.array_equal(np.concatenate([r.score for r in results]), whole.score))
            if not ok:
                bad += 1
                print("MISMATCH seed", seed, "world", world, "events", whole.n_events)
            rounds += 1
        except binding.LpError as ex:
            print("error", seed, ex)
        finally:
            for lp in lps:
                lp.close()
    print("rounds", rounds, "bad", bad)


if __name__ == "__main__":
    main()
