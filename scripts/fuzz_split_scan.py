"""Differential fuzz of the one pass over the log (k_fscan / k_scan_tiles / k_expand / k_log_meta) against the oracle on
texts built to stress its seams: lengths around multiples of the 1 KiB span and the 16 KiB tile, literals and "\r\n" pairs
straddling span and tile boundaries, runs of newlines dense enough to overflow a span's list (the re-read path of
k_expand), lines longer than a tile, texts that end inside a line / in "\r" / in many separators, batches whose logs
begin and end at such places. Hunting tool (GPU); the fixed cases live in tests/test_gpu_parity.py.
Usage: python scripts/fuzz_split_scan.py [first_seed] [last_seed]"""
import random
import sys

sys.path.insert(0, ".")
import numpy as np  # noqa: E402

from log_parser_b200 import binding  # noqa: E402
from log_parser_b200.model import PatternSpec, ScoringConfig  # noqa: E402
from oracle import oracle  # noqa: E402

WORDS = [b"needle-1", b"ERRx", b"abcd", b"Exception", b"timeout now"]
PATS = [PatternSpec(id="a", primary_regex=r"needle-1", confidence=0.6), PatternSpec(id="b", primary_regex=r"\bERRx\b", confidence=0.5),
        PatternSpec(id="c", primary_regex=r"abcd$", confidence=0.5), PatternSpec(id="d", primary_regex=r"^$", confidence=0.4),
        PatternSpec(id="e", primary_regex=r"Exception.*timeout now", confidence=0.7), PatternSpec(id="f", primary_regex=r"^\s+at \w+", confidence=0.3)]


def piece(rng: random.Random) -> bytes:
    k = rng.randrange(10)
    if k == 0:
        return b"\n" * rng.randint(1, 300)                       # dense newlines: more than 64 per span
    if k == 1:
        return b"\r\n" * rng.randint(1, 80)
    if k == 2:
        return bytes(rng.choice(b"xyz .:\t") for _ in range(rng.randint(0, 40)))
    if k == 3:
        return rng.choice(WORDS)
    if k == 4:
        return b"q" * rng.randint(900, 1100)                     # pushes what follows across a span boundary
    if k == 5:
        return b"w" * rng.randint(16000, 17000)                  # ... across a tile boundary; a line longer than a tile
    if k == 6:
        return b"\tat com.Foo.bar(Foo.java:1)\n"
    if k == 7:
        return rng.choice(WORDS) + rng.choice([b"\n", b"\r\n", b"\r", b""])
    if k == 8:
        return b"Exception in x " + b"." * rng.randint(0, 60) + b" timeout now" + rng.choice([b"\n", b"\r\n"])
    return rng.choice([b"\n", b"\r\n", b"\r", b" "])


def make_text(rng: random.Random) -> bytes:
    target = rng.choice([0, 1, 15, 16, 17, 1023, 1024, 1025, 1040, 2048, 16383, 16384, 16385, 16400, 32768, 40000, 70000])
    out = bytearray()
    while len(out) < target:
        out += piece(rng)
    if rng.random() < 0.5:
        del out[target:]                                        # cut exactly at a seam length
    if rng.random() < 0.3:
        out += rng.choice([b"\n", b"\r\n", b"\r", b"\n\n\n", b"needle-1"])
    return bytes(out).replace(b"\0", b" ")


def same(res, ora, k=0) -> bool:
    b, e = int(res.log_event_begin[k]), int(res.log_event_begin[k + 1])
    return (int(res.log_n_lines[k]) == ora.n_lines and e - b == len(ora.line) and np.array_equal(res.line[b:e], ora.line)
            and np.array_equal(res.pattern[b:e], ora.pattern) and np.array_equal(res.score[b:e], ora.score))


def main():
    lo = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    hi = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    cfg = ScoringConfig()
    lp = binding.LogParser(cfg)
    lp.load_patterns(PATS)
    table = oracle.PatternTable(PATS)
    bad = n = 0
    for seed in range(lo, hi):
        rng = random.Random(seed)
        for _ in range(12):
            text = make_text(rng)
            res = lp.analyze(text, options=binding.OPT_LINE_OFFSETS)
            ora = oracle.analyze(cfg, PATS, text, table=table)
            ok = same(res, ora)
            if ok:
                for i, (s, e) in enumerate(oracle.split(text)):
                    if (int(res.line_start[i]), int(res.line_end[i])) != (s, e):
                        ok = False
                        break
            n += 1
            if not ok:
                bad += 1
                print("MISMATCH seed", seed, "len", len(text), "lines", ora.n_lines, int(res.log_n_lines[0]), "events", len(ora.line), res.n_events)
        logs = [make_text(rng) for _ in range(rng.randint(2, 6))]
        res = lp.analyze_batch(logs)
        prior = np.zeros(len(PATS), np.uint64)
        for k, one in enumerate(logs):
            ora = oracle.analyze(cfg, PATS, one, prior_count=prior, table=table)
            n += 1
            if not same(res, ora, k):
                bad += 1
                print("BATCH MISMATCH seed", seed, "log", k, "len", len(one))
            prior = prior + ora.freq_delta.astype(np.uint64)
    print("checked", n, "bad", bad)
    lp.close()
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
