"""Differential fuzz of the whole path (events, all seven factors, scores, ranks) against the oracle with random
regexes as primaries, secondaries and sequence events and random scoring.* values.
Usage: python scripts/fuzz_scoring_vs_oracle.py [first_seed] [end_seed]"""
import random
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import numpy as np  # noqa: E402
from log_parser_b200 import binding  # noqa: E402
from log_parser_b200.model import PatternSpec, ScoringConfig, SecondaryPattern, SequencePattern  # noqa: E402
from oracle import oracle  # noqa: E402
from fuzz_range_consistency import rx, WORDS, ALPHA  # noqa: E402

NAMES = ["base", "sev", "chrono", "prox", "temporal", "ctx", "penalty"]


def gen_case(seed):
    rng = random.Random(seed)
    cfg = ScoringConfig(max_window=rng.choice([3, 40, 100]), decay_constant=rng.choice([2.0, 10.0]),
                        frequency_threshold=rng.choice([2.0, 10.0]), max_context_factor=rng.choice([1.5, 2.5]),
                        early_bonus_threshold=rng.choice([0.1, 0.2]), penalty_threshold=rng.choice([0.5, 0.7]))
    pats = []
    for k in range(8):
        secs = [SecondaryPattern(regex=rx(rng), weight=rng.choice([0.3, 0.5, 0.8]), proximity_window=rng.choice([0, 1, 7, 30, 200]))
                for _ in range(rng.randint(0, 3))]
        seqs = [SequencePattern(events=[rx(rng) for _ in range(rng.randint(0, 4))], bonus_multiplier=rng.choice([0.2, 0.5]))
                for _ in range(rng.randint(0, 2))]
        ctx = rng.random() < 0.6
        pats.append(PatternSpec(id=rng.choice([f"p{k % 3}", f"u{k}", "", None]), primary_regex=rx(rng), confidence=0.3 + 0.08 * k,
                                severity=rng.choice(["LOW", "HIGH", "CRITICAL", "weird"]),
                                lines_before=rng.randint(0, 12) if ctx else None, lines_after=rng.randint(0, 25) if ctx else None,
                                secondaries=secs, sequences=seqs))
    lines = []
    for _ in range(rng.choice([1, 40, 400, 900])):
        parts = [rng.choice(WORDS + ["ERROR", "WARN", "\tat a.B(C.java:1)", "FooException"]) if rng.random() < 0.4
                 else "".join(rng.choice(ALPHA) for _ in range(rng.randint(0, 6))) for _k in range(rng.randint(0, 4))]
        lines.append("".join(parts))
    log = ("\n".join(lines) + rng.choice(["", "\n", "\n\n\r\n"])).encode()
    prior = np.array([rng.choice([0, 0, 3, 11]) for _ in pats], np.uint64)
    ids = {}
    for i, q in enumerate(pats):                  # one count per id
        if q.id and q.id.strip():
            prior[i] = prior[ids.setdefault(q.id, i)]
    return cfg, pats, log, prior


def main():
    lo = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    hi = int(sys.argv[2]) if len(sys.argv) > 2 else 31
    bad = rounds = 0
    for seed in range(lo, hi):
        cfg, pats, log, prior = gen_case(seed)
        lp = binding.LogParser(cfg)
        try:
            lp.load_patterns(pats)
            res = lp.analyze(log, prior_count=prior, options=binding.OPT_FACTORS)
            ora = oracle.analyze(cfg, pats, log, prior_count=prior)
            ok = (res.n_lines == ora.n_lines and np.array_equal(res.line, ora.line) and np.array_equal(res.pattern, ora.pattern))
            why = "events"
            if ok and len(ora.line):
                for c, nm in enumerate(NAMES):
                    if not np.array_equal(res.factors[:, c], ora.factors[:, c]):
                        ok, why = False, nm
                        break
                if ok and not np.array_equal(res.score, ora.score):
                    ok, why = False, "score"
                if ok and not np.array_equal(res.freq_delta, ora.freq_delta):
                    ok, why = False, "freq_delta"
            if not ok:
                bad += 1
                print("MISMATCH seed", seed, why, "events", res.n_events, len(ora.line))
                if why in NAMES and "-v" in sys.argv:
                    c = NAMES.index(why)
                    k = int(np.nonzero(res.factors[:, c] != ora.factors[:, c])[0][0])
                    i, p = int(res.line[k]) - 1, int(res.pattern[k])
                    sp = pats[p]
                    print("  event", k, "line", i + 1, "pattern", p, repr(sp.primary_regex), "before/after", sp.lines_before, sp.lines_after,
                          "got", res.factors[k, c], "expected", ora.factors[k, c], "n_lines", res.n_lines)
                    text_lines = log.decode().split("\n")
                    b0 = max(0, i - (sp.lines_before or 0)) if sp.lines_before is not None else i
                    a1 = min(res.n_lines, i + 1 + (sp.lines_after or 0)) if sp.lines_before is not None else i + 1
                    for l in range(b0, a1):
                        print("   ", l + 1, repr(text_lines[l]))
            rounds += 1
        except binding.LpError as ex:
            print("error", seed, ex)
        finally:
            lp.close()
    print("rounds", rounds, "bad", bad)


if __name__ == "__main__":
    main()
