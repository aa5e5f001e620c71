"""GPU parity tests proper: the CUDA path, called through the C ABI, against the oracle on the same
seeded inputs. Bar (BASELINE.json north_star): pattern ids and line numbers bit-exact, discovery
order identical, scores <= 1e-9 relative (they are in fact compared for bit equality first), final
ranking identical."""
import os
import random
import sys

import numpy as np
import pytest

from log_parser_b200 import binding, synth
from log_parser_b200.model import PatternSpec, ScoringConfig, SecondaryPattern, SequencePattern

pytestmark = pytest.mark.gpu

REL_TOL = 1e-9   # north_star: scores accumulated in fp64 must agree to <= 1e-9 relative


def _check(res, ora, log_index=0, check_factors=True):
    b, e = int(res.log_event_begin[log_index]), int(res.log_event_begin[log_index + 1])
    assert int(res.log_n_lines[log_index]) == ora.n_lines
    assert e - b == len(ora.line), f"event count {e - b} != oracle {len(ora.line)}"
    np.testing.assert_array_equal(res.line[b:e], ora.line)
    np.testing.assert_array_equal(res.pattern[b:e], ora.pattern)
    if check_factors and res.factors is not None and len(ora.line):
        names = ["base", "sev", "chrono", "prox", "temporal", "ctx", "penalty"]
        for k, nm in enumerate(names):
            bad = np.nonzero(res.factors[b:e, k] != ora.factors[:, k])[0]
            assert bad.size == 0, f"factor {nm} differs at event {bad[0]} (line {ora.line[bad[0]]}, pattern {ora.pattern[bad[0]]}): " \
                                  f"{res.factors[b + bad[0], k]!r} vs {ora.factors[bad[0], k]!r}"
    if len(ora.line):
        rel = np.abs(res.score[b:e] - ora.score) / np.maximum(np.abs(ora.score), 1e-300)
        assert rel.max() <= REL_TOL
        # ranking: score descending, ties by discovery order (stable)
        exp_order = np.argsort(-ora.score, kind="stable")
        exp_rank = np.empty_like(exp_order)
        exp_rank[exp_order] = np.arange(len(exp_order))
        np.testing.assert_array_equal(res.rank[b:e], exp_rank.astype(np.uint32))


@pytest.fixture(scope="module")
def parser():
    p = binding.LogParser(ScoringConfig())
    yield p
    p.close()


def test_c1_full_parity(parser, oracle):
    """configs[0]: one 10k-line Java stack-trace pod log vs ~50 patterns."""
    wl, log = synth.config_c1()
    parser.load_patterns(wl.patterns)
    res = parser.analyze(log, options=binding.OPT_FACTORS)
    ora = oracle.analyze(ScoringConfig(), wl.patterns, log)
    assert len(ora.line) > 20
    _check(res, ora)
    np.testing.assert_array_equal(res.freq_delta, ora.freq_delta)
    # bit-exact scores, not just within tolerance
    np.testing.assert_array_equal(res.score, ora.score)


@pytest.mark.parametrize("log,n", [
    (b"", 1), (b"\n\n", 0), (b"\na", 2), (b"a\r\nb\n\n", 2), (b"a\rb", 1), (b"a\n\r", 2), (b"a\n\r\n\n", 1),
    (b"a\n \n", 2), (b"\r\r\n", 1), (b"x", 1), (b"x\n", 1), (b"\n", 0), (b"\r\n", 0), (b"a\n\nb", 3),
    (b"hit\n" * 40, 40), (b"\n" * 100 + b"hit", 101), (b"hit" + b"\n" * 100, 1),
])
def test_split_edge_cases(parser, oracle, log, n):
    pats = [PatternSpec(id="p", primary_regex="hit|^$|a", confidence=0.7, severity="HIGH", lines_before=2, lines_after=2)]
    parser.load_patterns(pats)
    res = parser.analyze(log, options=binding.OPT_FACTORS | binding.OPT_LINE_OFFSETS)
    ora = oracle.analyze(ScoringConfig(), pats, log)
    assert ora.n_lines == n
    _check(res, ora)
    spans = oracle.split(log)
    for i, (s, e) in enumerate(spans):
        assert (int(res.line_start[i]), int(res.line_end[i])) == (s, e)


def test_long_tail_of_empty_lines_is_trimmed(parser, oracle):
    """A log (and logs of a batch) ending in hundreds of thousands of blank lines: String.split drops them all
    (AS:53); k_log_meta finds the last non-empty line 256 lines at a time, one block per log."""
    pats = [PatternSpec(id="p", primary_regex="hit", confidence=0.7, severity="HIGH")]
    parser.load_patterns(pats)
    log = b"a\nhit 1\n\nhit 2" + b"\n" * 300_000
    res = parser.analyze(log)
    ora = oracle.analyze(ScoringConfig(), pats, log)
    assert ora.n_lines == 4
    _check(res, ora)
    logs = [b"hit\r\n" + b"\r\n" * 70_000, b"x\n\n\nhit", b"\n" * 1000, b"hit" + b"\n" * 513]
    batch = parser.analyze_batch(logs)
    for k, one in enumerate(logs):
        _check(batch, oracle.analyze(ScoringConfig(), pats, one), log_index=k)


def test_scan_seams_against_oracle(oracle):
    """Fixed-seed slice of scripts/fuzz_split_scan.py: texts whose lengths, literals, "\\r\\n" pairs and newline runs sit on
    the seams of the scan pass (1 KiB spans, 16 KiB tiles, list overflow, lines longer than a tile), alone and in batches."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("fuzz_split_scan", os.path.join(os.path.dirname(__file__), "..", "scripts", "fuzz_split_scan.py"))
    fz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fz)
    argv = sys.argv
    try:
        sys.argv = ["fuzz_split_scan.py", "1", "13"]
        assert fz.main() == 0
    finally:
        sys.argv = argv


def test_all_factors_small(parser, oracle):
    log = b"\n".join([b"start", b"connect begin", b"WARN slow", b"boom OutOfMemoryError", b"retry", b"x", b"y", b"z", b"w", b"v"])
    pats = [PatternSpec(id="p1", primary_regex="OutOfMemoryError", confidence=0.8, severity="high", lines_before=2, lines_after=1,
                        secondaries=[SecondaryPattern("retry", 0.8, 10), SecondaryPattern("nomatch", 0.5, 10)],
                        sequences=[SequencePattern(["start", "connect", "boom"], 0.5), SequencePattern(["zzz", "boom"], 0.9)]),
            PatternSpec(id=None, primary_regex="^x$|retry", confidence=0.5, severity="weird"),
            PatternSpec(id="p1", primary_regex="boom", confidence=0.9, severity="LOW", lines_before=0, lines_after=0)]
    parser.load_patterns(pats)
    res = parser.analyze(log, options=binding.OPT_FACTORS)
    ora = oracle.analyze(ScoringConfig(), pats, log)
    assert len(ora.line) == 4
    _check(res, ora)
    np.testing.assert_array_equal(res.score, ora.score)


def test_random_small_logs_many_configs(parser, oracle):
    """Random pattern sets over a small vocabulary: dense matches, windows clipped at both ends,
    sequences of 1-4 events, shared regexes between roles, shared frequency ids."""
    rng = random.Random(1234)
    toks = ["ERROR", "warn", "Exception", "FooError", "alpha", "beta", "gamma", "delta", "at x.y(Z.java:1)", "retry", "timeout", "\tat a.b(C.java:2)"]
    for it in range(40):
        n = rng.randint(1, 120)
        lines = []
        for _ in range(n):
            k = rng.randint(0, 4)
            lines.append(" ".join(rng.choice(toks) for _ in range(k)))
        sep = rng.choice(["\n", "\r\n"])
        log = sep.join(lines).encode() + (b"\n" if rng.random() < 0.5 else b"")
        pats = []
        for p in range(rng.randint(1, 6)):
            voc = ["alpha", "beta", "gamma", "delta", "retry", "timeout", "Foo\\w+", "^$", "\\bat\\b", "a.*a", "[bg]\\w+a\\b"]
            secs = [SecondaryPattern(rng.choice(voc), round(rng.uniform(0.1, 0.9), 2), rng.randint(0, 30)) for _ in range(rng.randint(0, 3))]
            seqs = [SequencePattern([rng.choice(voc) for _ in range(rng.randint(0, 4))], round(rng.uniform(0.1, 0.9), 2)) for _ in range(rng.randint(0, 2))]
            ctx = rng.random() < 0.8
            pats.append(PatternSpec(id=rng.choice(["a", "b", "", None, "c"]), primary_regex=rng.choice(voc), confidence=round(rng.uniform(0.1, 1), 3),
                                    severity=rng.choice(["CRITICAL", "high", "Medium", "LOW", "INFO", "x"]),
                                    lines_before=rng.randint(0, 15) if ctx else None, lines_after=rng.randint(0, 15) if ctx else None,
                                    secondaries=secs, sequences=seqs))
        cfg = ScoringConfig(max_window=rng.choice([0, 3, 100]), frequency_threshold=rng.choice([2.0, 10.0]),
                            max_context_factor=rng.choice([2.5, 50.0]))
        lp = binding.LogParser(cfg)
        lp.load_patterns(pats)
        prior = [rng.randint(0, 15) for _ in pats] if rng.random() < 0.5 else None
        # patterns sharing an id share one frequency slot: the prior of the first one counts
        res = lp.analyze(log, prior_count=prior, options=binding.OPT_FACTORS)
        ora = oracle.analyze(cfg, pats, log, prior_count=prior)
        _check(res, ora)
        np.testing.assert_array_equal(res.score, ora.score)
        np.testing.assert_array_equal(res.freq_delta, ora.freq_delta)
        lp.close()


def test_non_ascii_lines(parser, oracle):
    log = "café ERROR\néERROR x\nnaïve FooException  \nplain\n\U0001F600 boom\n".encode("utf-8")
    pats = [PatternSpec(id="a", primary_regex=r"\bERROR\b", confidence=1.0, lines_before=1, lines_after=1),
            PatternSpec(id="b", primary_regex=r"[^\x00-\x7F]+ boom$", confidence=1.0),
            PatternSpec(id="c", primary_regex=r"plain$", confidence=1.0),
            PatternSpec(id="d", primary_regex=r"^.{5} ", confidence=1.0)]
    parser.load_patterns(pats)
    res = parser.analyze(log, options=binding.OPT_FACTORS)
    ora = oracle.analyze(ScoringConfig(), pats, log)
    _check(res, ora)


def test_batch_matches_sequential_requests(parser, oracle):
    """lp_analyze_batch == the reference handling the logs one request after the other
    (frequency state carried from log k to log k+1)."""
    wl = synth.make_patterns(40, 99, noliteral=1)
    logs = [synth.make_log(wl, n, 1000 + i, pool_blocks=500, hit_rate=0.05) for i, n in enumerate([300, 1, 77, 1200])]
    logs.insert(2, b"")
    logs.append(b"no newline at end")
    parser.load_patterns(wl.patterns)
    res = parser.analyze_batch(logs, options=binding.OPT_FACTORS)
    prior = np.zeros(len(wl.patterns), np.uint64)
    for k, lg in enumerate(logs):
        ora = oracle.analyze(ScoringConfig(), wl.patterns, lg, prior_count=prior)
        _check(res, ora, log_index=k)
        prior = prior + ora.freq_delta
    np.testing.assert_array_equal(res.freq_delta, prior)


def test_two_phase_equals_one_call(parser):
    import torch

    wl = synth.make_patterns(30, 5, noliteral=0)
    logs = [synth.make_log(wl, 400, 50 + i, pool_blocks=400, hit_rate=0.05) for i in range(3)]
    parser.load_patterns(wl.patterns)
    one = parser.analyze_batch(logs)
    blob = np.frombuffer(b"".join(logs), np.uint8)
    d = torch.from_numpy(blob.copy()).cuda()
    off = np.concatenate([[0], np.cumsum([len(x) for x in logs])]).astype(np.uint64)
    counts = parser.batch_scan_device(d.data_ptr(), off)
    two = parser.batch_score(None)
    np.testing.assert_array_equal(counts, one.freq_delta)
    np.testing.assert_array_equal(one.score, two.score)
    np.testing.assert_array_equal(one.line, two.line)
    np.testing.assert_array_equal(one.rank, two.rank)


def test_device_resident_equals_host_path(parser):
    import torch

    wl, log = synth.config_c1()
    parser.load_patterns(wl.patterns)
    a = parser.analyze(log)
    d = torch.from_numpy(np.frombuffer(log, np.uint8).copy()).cuda()
    b = parser.analyze_device(d.data_ptr(), d.numel())
    np.testing.assert_array_equal(a.score, b.score)
    np.testing.assert_array_equal(a.line, b.line)
    t = parser.timings()
    assert t["kernel_launches"] > 5 and t["scan_ms"] > 0


def test_backtracking_budget_degrades_to_a_count(parser):
    """A (line, regex) evaluation that exhausts the verifier's step budget (catastrophic backtracking; java.util.regex
    would keep going) no longer fails the request: the pair counts as no match and the result says how many evaluations
    were abandoned. LP_OPT_STRICT_BUDGET brings the old behaviour (LP_ERR_BACKTRACK) back."""
    pats = [PatternSpec(id="boom", primary_regex=r"(a+)+b\1", confidence=0.5), PatternSpec(id="plain", primary_regex="needle-here", confidence=0.5)]
    log = b"aba\n" + b"a" * 48 + b"c\nxx needle-here\n"
    parser.load_patterns(pats)
    res = parser.analyze(log)
    assert res.budget_exceeded >= 1
    assert [(int(l), int(p)) for l, p in zip(res.line, res.pattern)] == [(1, 0), (3, 1)]
    with pytest.raises(binding.LpError) as ei:
        parser.analyze(log, options=binding.OPT_STRICT_BUDGET)
    assert ei.value.code == binding.LP_ERR_BACKTRACK
    assert parser.analyze(b"aba\nneedle-here\n").budget_exceeded == 0


def test_deep_backtracking_stack_comes_from_the_pool(parser, oracle):
    """An evaluation whose backtracking stack outgrows the thread's local part (an alternation loop pushes one choice
    per iteration) continues on a deep stack taken from the pool: same events as the oracle, nothing abandoned.
    Many such lines at once, so that several threads hold a deep stack at the same time."""
    pats = [PatternSpec(id="deep", primary_regex=r"(k)(?:a|b)*c\1", confidence=0.5), PatternSpec(id="plain", primary_regex="needle-here", confidence=0.5)]
    rng = random.Random(5)
    lines = []
    for i in range(300):
        body = "".join(rng.choice("ab") for _ in range(rng.randint(150, 700)))
        lines.append("k" + body + ("ck" if i % 3 else "cz"))
        if i % 50 == 0:
            lines.append("xx needle-here")
    log = ("\n".join(lines) + "\n").encode()
    parser.load_patterns(pats)
    res = parser.analyze(log)
    ora = oracle.analyze(ScoringConfig(), pats, log)
    assert len(ora.line) >= 200
    assert res.budget_exceeded == 0
    _check(res, ora)


def test_c3_nonregular_patterns(parser, oracle):
    """configs[2] shape: look-around / back-reference / possessive+atomic primaries take the
    backtracking verifier on prefiltered candidate lines; results must still equal the oracle's."""
    wl = synth.make_patterns(80, 31337, nonregular_frac=0.5, noliteral=2)
    extra = [PatternSpec(id="nl1", primary_regex=r"(?<=\[)\w+-(\d+)\] \w+\.\w+", confidence=0.5, lines_before=3, lines_after=3),
             PatternSpec(id="nl2", primary_regex=r"x.{0,60}yz", confidence=0.5),
             PatternSpec(id="nl3", primary_regex=r"(?m)^\tat .*\)$", confidence=0.4, secondaries=[SecondaryPattern(r"(?i)caused by(?=:)", 0.5, 20)])]
    pats = wl.patterns + extra
    log = synth.make_log(wl, 4000, 4242, pool_blocks=1500, hit_rate=0.05)
    parser.load_patterns(pats)
    res = parser.analyze(log, options=binding.OPT_FACTORS)
    ora = oracle.analyze(ScoringConfig(), pats, log)
    assert len(ora.line) > 100
    nonreg = sum(1 for i in set(ora.pattern.tolist()) if any(t in pats[i].primary_regex for t in ("(?<", "(?=", "\\1", "(?>", "(?m)")))
    assert nonreg >= 5, "the log must exercise the backtracking path"
    _check(res, ora)
    np.testing.assert_array_equal(res.score, ora.score)


def test_bad_pattern_reports_index(parser):
    pats = [PatternSpec(id="ok", primary_regex="fine", confidence=1.0), PatternSpec(id="bad", primary_regex="a{,2}", confidence=1.0)]
    with pytest.raises(binding.PatternSyntaxError) as ei:
        parser.load_patterns(pats)
    assert ei.value.bad_pattern == 1


def test_c2_sample_parity_and_properties(parser, oracle):
    """configs[1] shape at a size the oracle finishes in seconds, plus size-independent properties
    on a larger log: events sorted in discovery order, rank is a permutation ordered by score,
    idempotence, and line count equal to the newline count."""
    wl, log = synth.config_c2(n_lines=20_000, n_patterns=500)
    parser.load_patterns(wl.patterns)
    res = parser.analyze(log, options=binding.OPT_FACTORS)
    ora = oracle.analyze(ScoringConfig(), wl.patterns, log)
    assert len(ora.line) > 50
    _check(res, ora)
    np.testing.assert_array_equal(res.score, ora.score)

    wl, big = synth.config_c2(n_lines=300_000, n_patterns=500)
    parser.load_patterns(wl.patterns)
    r1 = parser.analyze(big)
    r2 = parser.analyze(big)
    assert r1.n_lines == 300_000 == big.count(b"\n")
    key = r1.line.astype(np.uint64) * 1000 + r1.pattern
    assert (np.diff(key.astype(np.int64)) > 0).all()
    order = np.argsort(r1.rank)
    assert (np.sort(r1.rank) == np.arange(r1.n_events)).all()
    assert (np.diff(r1.score[order]) <= 0).all()
    np.testing.assert_array_equal(r1.score, r2.score)
    np.testing.assert_array_equal(r1.rank, r2.rank)
    # linearity of the event set: the first half of the log yields exactly the events of those lines
    # whose scoring windows do not reach the cut (checked on ids only)
    half = big[: big.index(b"\n", len(big) // 2) + 1]
    rh = parser.analyze(half)
    nh = rh.n_lines
    sel = r1.line <= nh
    np.testing.assert_array_equal(r1.line[sel], rh.line)
    np.testing.assert_array_equal(r1.pattern[sel], rh.pattern)


def test_literal_with_boundaries_patterns(parser, oracle):
    """Patterns of the form [\\b][\\w*]LITERAL[\\b] are decided at the literal occurrence (no automaton
    walk); mixed-case and case-insensitive variants must keep java.util.regex semantics."""
    lines = ["ERROR x", "xERROR y", "ERRORS", "an error here", "Failed to bind", "FAILED TO BIND", "failed to bind", "warn", "WARNING!",
             "prewarn", "FooException: x", "FooExceptions", "my_Error", "Error", "a SocketTimeoutException b", "aSocketTimeoutException",
             "shutdown hook invoked", "Shutdown hook invoked twice", "é ERROR é", "ERROR", "x", ""]
    log = ("\n".join(lines) + "\n").encode()
    regs = [r"\b(ERROR|FATAL)\b", r"(?i)\b(warn|warning)\b", r"[Ff]ailed to bind", r"(?i)failed TO bind", r"\b\w*Exception\b|\b\w*Error\b",
            r"\bSocketTimeoutException\b", r"Shutdown hook invoked", r"\w*Exception", r"ERROR\b", r"\bERROR", r"(?i:shut)down hook"]
    pats = [PatternSpec(id="p%d" % i, primary_regex=r, confidence=0.5, lines_before=2, lines_after=2,
                        secondaries=[SecondaryPattern(regs[(i + 1) % len(regs)], 0.5, 5)]) for i, r in enumerate(regs)]
    parser.load_patterns(pats)
    res = parser.analyze(log, options=binding.OPT_FACTORS)
    ora = oracle.analyze(ScoringConfig(), pats, log)
    assert len(ora.line) > 25
    _check(res, ora)
    np.testing.assert_array_equal(res.score, ora.score)


def test_sharded_blocks_equal_one_batch(parser):
    """The multi-GPU decomposition on one GPU: two contiguous blocks of logs, scanned and scored
    with the exclusive prefix of the per-pattern counts, concatenate to the single-batch result."""
    from log_parser_b200 import parallel

    wl = synth.make_patterns(40, 777, noliteral=1)
    logs = [synth.make_log(wl, n, 300 + i, pool_blocks=400, hit_rate=0.06) for i, n in enumerate([500, 200, 900, 50, 400, 300])]
    parser.load_patterns(wl.patterns)
    whole = parser.analyze_batch(logs)
    eng = parallel.GpuBatchEngine(parser)
    bounds = parallel.shard_bounds([len(x) for x in logs], 2)
    assert bounds[0][1] == bounds[1][0] and 0 < bounds[0][1] < len(logs)
    prior = np.zeros(len(wl.patterns), np.uint64)
    lines, pats, scores = [], [], []
    for lo, hi in bounds:
        counts = eng.scan(logs[lo:hi])
        r = eng.score(prior)
        np.testing.assert_array_equal(counts, r.freq_delta)
        lines.append(r.line); pats.append(r.pattern); scores.append(r.score)
        prior = prior + counts
    np.testing.assert_array_equal(np.concatenate(lines), whole.line)
    np.testing.assert_array_equal(np.concatenate(pats), whole.pattern)
    np.testing.assert_array_equal(np.concatenate(scores), whole.score)
    np.testing.assert_array_equal(prior, whole.freq_delta)


def test_c4_shape_large_windows(oracle):
    """configs[3] shape at a size the oracle finishes: dense secondary matches, proximity windows up
    to scoring.proximity.max-window = 1000, context +-500 lines, 3 secondaries + 2 sequences per pattern."""
    cfg = ScoringConfig(max_window=1000)
    wl = synth.make_patterns(30, 4040, scoring_heavy=True, noliteral=0)
    log = synth.make_log(wl, 6000, 4041, hit_rate=0.05, symptom_rate=0.4, phase_rate=0.01, pool_blocks=1500)
    lp = binding.LogParser(cfg)
    lp.load_patterns(wl.patterns)
    res = lp.analyze(log, options=binding.OPT_FACTORS)
    ora = oracle.analyze(cfg, wl.patterns, log)
    assert len(ora.line) > 100
    assert (ora.factors[:, 3] > 1.0).mean() > 0.5      # proximity bonus present on most events
    _check(res, ora)
    np.testing.assert_array_equal(res.score, ora.score)
    lp.close()


def _range_sharded(cfg, patterns, log, world, prior=None):
    """The line-range decomposition run on one GPU: `world` contexts play the ranks, the summaries
    are stacked instead of all-gathered."""
    from log_parser_b200 import parallel

    lps = [binding.LogParser(cfg) for _ in range(world)]
    try:
        for lp in lps:
            lp.load_patterns(patterns)
        engines = [parallel.GpuRangeEngine(lp) for lp in lps]
        end = parallel.strip_trailing_separators(log)
        assert not (end == 0 and len(log) > 0)
        pieces = parallel.split_line_ranges(log, world, engines[0].halo_lines(), end)
        summaries = np.stack([engines[r].scan(bytes(log[tb:te]), hb, ha) for r, (tb, te, hb, ha) in enumerate(pieces)])
        results = [engines[r].score(summaries, r, world, prior) for r in range(world)]
        return results, summaries
    finally:
        for lp in lps:
            lp.close()


@pytest.mark.parametrize("world", [1, 2, 3, 5])
def test_line_range_shards_equal_whole_log(parser, world):
    """SURVEY §8e second mode: one log cut by line range with halos, scored per shard after one
    exchange of shard summaries, concatenates (in shard order) to the whole-log result bit for bit:
    global line numbers, chronological factor on the whole log's N, windows through the halos, the
    unbounded backward sequence walk (SS:296-305) across shards, frequency counts in discovery order."""
    from log_parser_b200 import parallel

    cfg = ScoringConfig()
    wl = synth.make_patterns(40, 5150 + world, scoring_heavy=True, noliteral=1)
    log = synth.make_log(wl, 7000, 5151, hit_rate=0.05, symptom_rate=0.3, phase_rate=0.01, pool_blocks=1500) + b"\r\n\n\n"
    prior = (np.arange(len(wl.patterns)) % 7).astype(np.uint64)
    parser.load_patterns(wl.patterns)
    whole = parser.analyze(log, prior_count=prior, options=binding.OPT_FACTORS)
    assert whole.n_events > 200
    assert (whole.factors[:, 4] > 1.0).any() and (whole.factors[:, 6] > 0.0).any()     # sequences and penalties in play
    results, summaries = _range_sharded(cfg, wl.patterns, log, world, prior)
    assert int(summaries[:, 0].sum()) == int(whole.log_n_lines[0])
    assert all(int(r.log_n_lines[0]) == int(whole.log_n_lines[0]) for r in results)
    np.testing.assert_array_equal(np.concatenate([r.line for r in results]), whole.line)
    np.testing.assert_array_equal(np.concatenate([r.pattern for r in results]), whole.pattern)
    np.testing.assert_array_equal(np.concatenate([r.score for r in results]), whole.score)
    np.testing.assert_array_equal(sum(r.freq_delta for r in results), whole.freq_delta)
    merged = parallel.merge_ranks([r.score for r in results])
    np.testing.assert_array_equal(np.concatenate(merged), whole.rank)


def test_line_range_shards_c4_windows():
    """configs[3]: proximity windows up to 1000 lines and +-500 context lines => 1000-line halos."""
    cfg = ScoringConfig(max_window=1000)
    wl = synth.make_patterns(30, 4040, scoring_heavy=True, noliteral=0)
    log = synth.make_log(wl, 9000, 4042, hit_rate=0.05, symptom_rate=0.4, phase_rate=0.01, pool_blocks=1500)
    lp = binding.LogParser(cfg)
    lp.load_patterns(wl.patterns)
    assert lp.range_halo_lines() >= 500
    whole = lp.analyze(log)
    lp.close()
    results, _ = _range_sharded(cfg, wl.patterns, log, 3)
    np.testing.assert_array_equal(np.concatenate([r.line for r in results]), whole.line)
    np.testing.assert_array_equal(np.concatenate([r.pattern for r in results]), whole.pattern)
    np.testing.assert_array_equal(np.concatenate([r.score for r in results]), whole.score)


@pytest.mark.parametrize("log", [b"", b"x", b"ERROR a\nb", b"a\n\n\nb\n", b"\n\nz", b"a\r\nb\r\n\r\n"])
def test_line_range_shards_tiny_logs(parser, log):
    """More ranks than lines: shards without own lines still take part in the exchange."""
    cfg = ScoringConfig()
    pats = [PatternSpec(id="p", primary_regex="ERROR|b|z|^$", confidence=0.9, severity="HIGH",
                        secondaries=[SecondaryPattern(regex="a", weight=0.5, proximity_window=3)],
                        sequences=[SequencePattern(events=["a", "b|z"], bonus_multiplier=0.3)])]
    parser.load_patterns(pats)
    whole = parser.analyze(log)
    results, summaries = _range_sharded(cfg, pats, log, 4)
    assert int(summaries[:, 0].sum()) == int(whole.log_n_lines[0])
    np.testing.assert_array_equal(np.concatenate([r.line for r in results]), whole.line)
    np.testing.assert_array_equal(np.concatenate([r.score for r in results]), whole.score)


def test_line_range_sequence_chain_spans_shards(parser):
    """The backward sequence walk is unbounded (SS:296-305): its events may sit in different shards,
    far outside any halo. 4 events: alpha@53, beta@32+214, gamma@102+333, then the last event next to
    the primary; with 5 shards of ~90 lines the chain enters shards in states 1 and 2."""
    cfg = ScoringConfig()
    lines = [b"INFO tick %d" % i for i in range(450)]
    lines[53] = b"alpha"; lines[32] = b"beta"; lines[214] = b"beta"; lines[102] = b"gamma"; lines[333] = b"gamma"
    for i in range(0, 441, 7):
        lines[i + 1] = b"ERR boom"
        lines[i + 2] = b"delta"
    log = b"\n".join(lines) + b"\n"
    pats = [PatternSpec(id="p", primary_regex="ERR", confidence=0.9, severity="HIGH",
                        sequences=[SequencePattern(events=["alpha", "beta", "gamma", "delta"], bonus_multiplier=0.25),
                                   SequencePattern(events=["gamma", "alpha", "delta"], bonus_multiplier=0.5),      # never completes
                                   SequencePattern(events=["delta"], bonus_multiplier=0.125)])]
    parser.load_patterns(pats)
    whole = parser.analyze(log, options=binding.OPT_FACTORS)
    t = whole.factors[:, 4]
    assert set(np.unique(t)) == {1.125, 1.375}
    assert int(whole.line[np.argmax(t > 1.2)]) == 338          # first primary line after gamma@333 (0-based 337)
    for world in (2, 5):
        results, _ = _range_sharded(cfg, pats, log, world)
        assert len(results) == world and min(len(r.line) for r in results) > 0
        np.testing.assert_array_equal(np.concatenate([r.line for r in results]), whole.line)
        np.testing.assert_array_equal(np.concatenate([r.score for r in results]), whole.score)


def test_device_side_exchange_equals_one_batch(parser):
    """lp_batch_scan_device_async / lp_batch_score_gathered: the by-log exchange without a host round
    trip. Two contexts play two ranks; each scan phase writes its [counts, status] straight into its
    slot of the gathered buffer (what the NCCL all-gather produces)."""
    import torch
    from log_parser_b200 import parallel

    wl = synth.make_patterns(40, 778, noliteral=1)
    logs = [synth.make_log(wl, n, 320 + i, pool_blocks=400, hit_rate=0.06) for i, n in enumerate([500, 200, 900, 50, 400, 300])]
    prior = (np.arange(len(wl.patterns)) % 5).astype(np.uint64)
    parser.load_patterns(wl.patterns)
    whole = parser.analyze_batch(logs, prior_count=prior)
    bounds = parallel.shard_bounds([len(x) for x in logs], 2)
    lps = [binding.LogParser(ScoringConfig()) for _ in range(2)]
    try:
        for lp in lps:
            lp.load_patterns(wl.patterns)
        n = len(wl.patterns) + 1
        gathered = torch.zeros(2 * n, dtype=torch.int64, device="cuda")
        texts, offs = [], []
        for lo, hi in bounds:
            blob = b"".join(logs[lo:hi])
            off = np.zeros(hi - lo + 1, np.uint64)
            np.cumsum([len(x) for x in logs[lo:hi]], out=off[1:])
            texts.append(torch.from_numpy(np.frombuffer(blob, dtype=np.uint8).copy()).cuda())
            offs.append(off)
        rounds = 0
        while True:
            rounds += 1
            assert rounds < 10
            for r in range(2):
                lps[r].batch_scan_device_async(texts[r].data_ptr(), offs[r], gathered[r * n:].data_ptr())
            torch.cuda.synchronize()
            results = [lps[r].batch_score_gathered(gathered.data_ptr(), r, 2, prior) for r in range(2)]
            assert (results[0] is None) == (results[1] is None)       # every rank sees the same status words
            if results[0] is not None:
                break
        counts = gathered.cpu().numpy().reshape(2, n)
        assert (counts[:, -1] == 0).all()
        np.testing.assert_array_equal(np.concatenate([r.line for r in results]), whole.line)
        np.testing.assert_array_equal(np.concatenate([r.pattern for r in results]), whole.pattern)
        np.testing.assert_array_equal(np.concatenate([r.score for r in results]), whole.score)
        np.testing.assert_array_equal(counts[:, :-1].sum(axis=0).astype(np.uint64), whole.freq_delta)
        # world_size 1 through the helper
        ex = parallel.DeviceExchange(lps[0], 0, 1)
        blob = b"".join(logs)
        off = np.zeros(len(logs) + 1, np.uint64)
        np.cumsum([len(x) for x in logs], out=off[1:])
        t = torch.from_numpy(np.frombuffer(blob, dtype=np.uint8).copy()).cuda()
        one = ex.step(t.data_ptr(), off, prior)
        ex.close()
        np.testing.assert_array_equal(one.score, whole.score)
    finally:
        for lp in lps:
            lp.close()


def test_piecewise_host_copy_equals_device_resident(parser, oracle):
    """lp_analyze on a log of more than 8 MB copies it in pieces on a second stream and scans each piece
    as it lands (after the prefilter has been calibrated by a first call): same result as the
    device-resident path, including candidates that straddle a cut, and both equal to the oracle
    (Parse.java:44 -> AnalysisService.java:53 is one host string there)."""
    import torch

    wl = synth.make_patterns(60, 909, noliteral=1)
    log = synth.make_log(wl, 90_000, 910, hit_rate=0.03)
    assert len(log) > (8 << 20)
    parser.load_patterns(wl.patterns)
    first = parser.analyze(log)                       # calibrates; plain copy
    second = parser.analyze(log)                      # piecewise copy overlapped with the scan
    dev = torch.from_numpy(np.frombuffer(log, dtype=np.uint8).copy()).cuda()
    resident = parser.analyze_device(dev.data_ptr(), len(log))
    assert resident.n_events > 1000
    for r in (first, second):
        np.testing.assert_array_equal(r.line, resident.line)
        np.testing.assert_array_equal(r.pattern, resident.pattern)
        np.testing.assert_array_equal(r.score, resident.score)
        np.testing.assert_array_equal(r.rank, resident.rank)
    assert parser.timings()["kernel_launches"] > 0
    ora = oracle.analyze(ScoringConfig(), wl.patterns, log)
    _check(second, ora, check_factors=False)
    np.testing.assert_array_equal(second.score, ora.score)


def test_service_analyze_on_device_equals_oracle_fed_assembly(tmp_path, oracle):
    """SURVEY §8f N1+N2 through the GPU: a YAML pattern directory and a PodFailureData-shaped dict go through
    service.AnalysisService.analyze (lp_analyze underneath); the whole AnalysisResult (events with their
    context strings, metadata.totalLines / patternsUsed, summary) equals the assembly fed with the oracle's
    columns (AnalysisService.java:53-121, 132-156, 166-215). A second request sees the first one's
    frequency state (FrequencyTrackingService.java:41-93)."""
    import yaml

    from log_parser_b200 import patterns, service

    set_a = {"metadata": {"library_id": "lib-a"}, "patterns": [
        {"id": "oom", "name": "OOM", "severity": "critical", "primary_pattern": {"regex": "OutOfMemoryError", "confidence": 0.9},
         "secondary_patterns": [{"regex": "GC overhead", "weight": 0.6, "proximity_window": 20}],
         "sequence_patterns": [{"description": "heap grows", "bonus_multiplier": 0.3, "events": [{"regex": "heap"}, {"regex": "OutOfMemory"}]}],
         "context_extraction": {"lines_before": 1, "lines_after": 2, "include_stack_trace": True}},
        {"id": "warn", "name": "Warn", "severity": "LOW", "primary_pattern": {"regex": "\\bWARN\\b", "confidence": 0.4}},
        {"id": "uni", "name": "Unicode word", "severity": "medium", "primary_pattern": {"regex": "\\p{Lu}\\p{L}+Exception", "confidence": 0.7},
         "context_extraction": {"lines_before": 2, "lines_after": 0}}]}
    set_b = {"metadata": {"library_id": "lib-b"}, "patterns": [
        {"id": "odd", "name": "Odd", "severity": "weird", "primary_pattern": {"regex": "^x$", "confidence": 0.5},
         "context_extraction": {"lines_before": 0, "lines_after": 0}}]}
    (tmp_path / "a.yml").write_text(yaml.safe_dump(set_a))
    (tmp_path / "b.yaml").write_text(yaml.safe_dump(set_b))
    sets = patterns.load_pattern_directory(str(tmp_path))
    assert len(sets) == 2
    rng = random.Random(77)
    filler = ["INFO request served", "heap usage 71%", "WARN slow call", "x", "java.lang.OutOfMemoryError: GC overhead limit exceeded",
              "\tat com.example.Svc.run(Svc.java:42)", "Ünexpected ÉtatException raised", "", "WARNING not a word match", "GC overhead high"]
    body = [rng.choice(filler) for _ in range(400)]
    logs = "\r\n".join(body[:200]) + "\n" + "\n".join(body[200:]) + "\n\n"
    now = [1_000_000.0]
    svc = service.AnalysisService(sets, clock=lambda: now[0])
    specs, raw = patterns.flatten(sets)
    prior = np.zeros(len(specs), np.uint64)
    for request in range(2):
        got = svc.analyze({"podName": "p", "logs": logs})
        ora = oracle.analyze(ScoringConfig(), specs, logs.encode("utf-8"), prior_count=prior)
        assert len(ora.line) > 50
        exp = service.assemble_result(ora.line, ora.pattern, ora.score, logs, raw, sets, 0.0, ora.n_lines)
        assert got["metadata"]["totalLines"] == exp["metadata"]["totalLines"]
        assert got["metadata"]["patternsUsed"] == exp["metadata"]["patternsUsed"]
        assert got["summary"] == exp["summary"]
        assert len(got["events"]) == len(exp["events"])
        for g, e in zip(got["events"], exp["events"]):
            assert g["lineNumber"] == e["lineNumber"] and g["matchedPattern"] == e["matchedPattern"] and g["context"] == e["context"]
            assert abs(g["score"] - e["score"]) <= REL_TOL * abs(e["score"])
        assert {k: v for k, v in svc.freq.get_frequency_stats().items()} == \
               {s.id: int(prior[i] + ora.freq_delta[i]) for i, s in enumerate(specs) if prior[i] + ora.freq_delta[i] > 0}
        prior = prior + ora.freq_delta.astype(np.uint64)
        now[0] += 60.0
    svc.parser.close()


def test_match_spans_equal_oracle(parser, oracle):
    """LP_OPT_SPANS (SURVEY §8-R extension; north_star: spans bit-exact): Matcher.start()/end() of the
    first find() of each event's primary regex, from the device backtracking engine, against the
    oracle's java.util.regex restatement. Oracle offsets are code points, the library's are UTF-8 bytes."""
    pats = [PatternSpec(id="a", primary_regex=r"\b\w*Exception\b", confidence=0.9, severity="HIGH"),
            PatternSpec(id="b", primary_regex=r"(?:GET|POST) (/\S*?)(\?\S+)? HTTP", confidence=0.5, severity="LOW"),
            PatternSpec(id="c", primary_regex=r"retry(?:ing)? in \d+?ms|timeout", confidence=0.5, severity="LOW"),
            PatternSpec(id="d", primary_regex=r"(\w+)=\1", confidence=0.5, severity="LOW"),
            PatternSpec(id="e", primary_regex=r"(?<=user=)\w++(?!\w)", confidence=0.5, severity="LOW"),
            PatternSpec(id="f", primary_regex=r"é+x|a.c$", confidence=0.5, severity="LOW"),
            PatternSpec(id="g", primary_regex=r"^$", confidence=0.5, severity="LOW")]
    lines = ["java.lang.NullPointerException at x", "GET /a/b?q=1 HTTP/1.1", "POST / HTTP", "retrying in 250ms then timeout",
             "timeout; retry in 5ms", "k=k v=w w=w", "login user=alice ok", "ééx then abc", "abc\r", "", "nothing here",
             "xException yException", "préfix NullPointerException é"]
    log = "\n".join(lines).encode("utf-8") + b"\n"
    parser.load_patterns(pats)
    res = parser.analyze(log, options=binding.OPT_SPANS)
    ora = oracle.analyze(ScoringConfig(), pats, log)
    np.testing.assert_array_equal(res.line, ora.line)
    np.testing.assert_array_equal(res.pattern, ora.pattern)
    assert res.n_events >= 12 and res.span_start is not None
    rx = [oracle.Regex(p.primary_regex) for p in pats]
    for k in range(res.n_events):
        text = lines[int(res.line[k]) - 1]
        s, e = rx[int(res.pattern[k])].span(text)
        exp = (len(text[:s].encode("utf-8")), len(text[:e].encode("utf-8")))
        assert (int(res.span_start[k]), int(res.span_end[k])) == exp, (pats[int(res.pattern[k])].primary_regex, text)
    # a larger log: every span lies inside its line and the spanned text matches on its own
    wl = synth.make_patterns(40, 1212, noliteral=1)
    log = synth.make_log(wl, 4000, 1213, hit_rate=0.05, pool_blocks=600)
    parser.load_patterns(wl.patterns)
    res = parser.analyze(log, options=binding.OPT_SPANS | binding.OPT_LINE_OFFSETS)
    assert res.n_events > 100
    rx = {}
    for k in range(0, res.n_events, 7):
        g = int(res.line[k]) - 1
        text = log[int(res.line_start[g]):int(res.line_end[g])].decode("utf-8")
        p = int(res.pattern[k])
        o = rx.setdefault(p, oracle.Regex(wl.patterns[p].primary_regex))
        s, e = o.span(text)
        assert (int(res.span_start[k]), int(res.span_end[k])) == (len(text[:s].encode()), len(text[:e].encode()))


def test_c5_shape_many_logs_in_one_batch(parser, oracle):
    """configs[4] shape at a size the oracle finishes: a batch of 1500 pod logs of very different sizes
    (empty, separators only, no final newline, \\r\\n) processed as the reference would process the
    requests one after the other, frequency counts carried across them; per-log ranks."""
    rnd = random.Random(5005)
    wl = synth.make_patterns(30, 5006, noliteral=1)
    pool = synth.make_log(wl, 6000, 5007, hit_rate=0.06, pool_blocks=900).split(b"\n")
    logs = []
    for k in range(1500):
        r = rnd.random()
        if r < 0.03:
            logs.append(b"")
        elif r < 0.05:
            logs.append(rnd.choice([b"\n", b"\r\n\n", b"\n\n\n"]))
        else:
            n = rnd.choice([1, 2, 5, 20, 60, 150])
            at = rnd.randrange(0, len(pool) - n)
            sep = b"\r\n" if rnd.random() < 0.1 else b"\n"
            body = sep.join(pool[at:at + n])
            logs.append(body + (b"" if rnd.random() < 0.3 else sep))
    parser.load_patterns(wl.patterns)
    res = parser.analyze_batch(logs)
    assert res.n_logs == len(logs) and res.n_events > 2000
    prior = np.zeros(len(wl.patterns), np.uint64)
    for k, lg in enumerate(logs):
        ora = oracle.analyze(ScoringConfig(), wl.patterns, lg, prior_count=prior)
        _check(res, ora, log_index=k, check_factors=False)
        b, e = int(res.log_event_begin[k]), int(res.log_event_begin[k + 1])
        np.testing.assert_array_equal(res.score[b:e], ora.score)
        prior = prior + ora.freq_delta
    np.testing.assert_array_equal(res.freq_delta, prior)
    assert (prior > 10).any()          # the frequency penalty is in play across requests


def _dist_worker(rank, world, port, path):
    import os
    import torch
    import torch.distributed as dist
    from log_parser_b200 import parallel

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)      # only carries the ncclUniqueId
    wl = synth.make_patterns(40, 778, noliteral=1)
    logs = [synth.make_log(wl, n, 320 + i, pool_blocks=400, hit_rate=0.06) for i, n in enumerate([500, 200, 900, 50, 400, 300])]
    prior = (np.arange(len(wl.patterns)) % 5).astype(np.uint64)
    lo, hi = parallel.shard_bounds([len(x) for x in logs], world)[rank]
    lp = binding.LogParser(ScoringConfig(), device=rank)
    lp.load_patterns(wl.patterns)
    parallel.init_library_comm(lp, rank, world)
    blob = b"".join(logs[lo:hi])
    off = np.zeros(hi - lo + 1, np.uint64)
    np.cumsum([len(x) for x in logs[lo:hi]], out=off[1:])
    text = torch.from_numpy(np.frombuffer(blob, dtype=np.uint8).copy()).cuda()
    for _ in range(3):       # the first steps repeat internally while the scratch estimates settle
        r = lp.analyze_batch_dist_device(text.data_ptr(), off, prior)
    np.savez(os.path.join(path, f"d{rank}.npz"), line=np.array(r.line), pattern=np.array(r.pattern), score=np.array(r.score))
    dist.barrier()
    dist.destroy_process_group()
    lp.close()


def test_library_nccl_batch_two_gpus(parser, tmp_path):
    """lp_comm_init + lp_analyze_batch_dist_device on two GPUs (skipped on a one-GPU box): two ranks, one
    collective call per step, concatenate to the single-batch result."""
    import socket
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    wl = synth.make_patterns(40, 778, noliteral=1)
    logs = [synth.make_log(wl, n, 320 + i, pool_blocks=400, hit_rate=0.06) for i, n in enumerate([500, 200, 900, 50, 400, 300])]
    prior = (np.arange(len(wl.patterns)) % 5).astype(np.uint64)
    parser.load_patterns(wl.patterns)
    whole = parser.analyze_batch(logs, prior_count=prior)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_dist_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    parts = [np.load(os.path.join(tmp_path, f"d{r}.npz")) for r in range(2)]
    np.testing.assert_array_equal(np.concatenate([p["line"] for p in parts]), whole.line)
    np.testing.assert_array_equal(np.concatenate([p["pattern"] for p in parts]), whole.pattern)
    np.testing.assert_array_equal(np.concatenate([p["score"] for p in parts]), whole.score)


def _range_dist_worker(rank, world, port, path):
    import os
    import torch
    import torch.distributed as dist
    from log_parser_b200 import parallel

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)      # only carries the ncclUniqueId
    wl = synth.make_patterns(40, 5152, scoring_heavy=True, noliteral=1)
    log = synth.make_log(wl, 7000, 5151, hit_rate=0.05, symptom_rate=0.3, phase_rate=0.01, pool_blocks=1500) + b"\r\n\n\n"
    prior = (np.arange(len(wl.patterns)) % 7).astype(np.uint64)
    lp = binding.LogParser(ScoringConfig(), device=rank)
    lp.load_patterns(wl.patterns)
    parallel.init_library_comm(lp, rank, world)
    for _ in range(3):       # the first steps repeat internally while the scratch estimates settle
        r = parallel.analyze_range_dist(lp, log, rank, world, prior)
    np.savez(os.path.join(path, f"r{rank}.npz"), line=np.array(r.line), pattern=np.array(r.pattern), score=np.array(r.score),
             delta=np.array(r.freq_delta), n=np.array(r.log_n_lines))
    dist.barrier()
    dist.destroy_process_group()
    lp.close()


def test_line_range_dist_two_gpus(parser, tmp_path):
    """lp_analyze_range_dist_device on two GPUs (skipped on a one-GPU box): one log cut by line range, summaries
    all-gathered by NCCL inside the library and composed on the device; the shards concatenate to the whole-log result."""
    import socket
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    wl = synth.make_patterns(40, 5152, scoring_heavy=True, noliteral=1)
    log = synth.make_log(wl, 7000, 5151, hit_rate=0.05, symptom_rate=0.3, phase_rate=0.01, pool_blocks=1500) + b"\r\n\n\n"
    prior = (np.arange(len(wl.patterns)) % 7).astype(np.uint64)
    parser.load_patterns(wl.patterns)
    whole = parser.analyze(log, prior_count=prior)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_range_dist_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    parts = [np.load(os.path.join(tmp_path, f"r{r}.npz")) for r in range(2)]
    np.testing.assert_array_equal(np.concatenate([p["line"] for p in parts]), whole.line)
    np.testing.assert_array_equal(np.concatenate([p["pattern"] for p in parts]), whole.pattern)
    np.testing.assert_array_equal(np.concatenate([p["score"] for p in parts]), whole.score)
    np.testing.assert_array_equal(sum(p["delta"] for p in parts), whole.freq_delta)
    assert all(int(p["n"][0]) == int(whole.log_n_lines[0]) for p in parts)


def test_line_range_dist_single_rank_equals_whole_log(parser):
    """lp_analyze_range_dist_device with a communicator of one rank (runs on a one-GPU box): summary built on the
    device, all-gather, k_range_compose, k_score reading the shard's place and the sequence thresholds from device
    memory; must give exactly lp_analyze's result, also from a host buffer and on a fresh context (repeat loop)."""
    from log_parser_b200 import parallel

    wl = synth.make_patterns(40, 5152, scoring_heavy=True, noliteral=1)
    log = synth.make_log(wl, 7000, 5151, hit_rate=0.05, symptom_rate=0.3, phase_rate=0.01, pool_blocks=1500) + b"\n\n"
    prior = (np.arange(len(wl.patterns)) % 6).astype(np.uint64)
    parser.load_patterns(wl.patterns)
    whole = parser.analyze(log, prior_count=prior, options=binding.OPT_FACTORS)
    assert whole.n_events > 200 and (whole.factors[:, 4] > 1.0).any()
    lp = binding.LogParser(ScoringConfig())
    try:
        lp.load_patterns(wl.patterns)
        lp.comm_init(lp.comm_unique_id(), 0, 1)
        for _ in range(3):
            r = parallel.analyze_range_dist(lp, log, 0, 1, prior, options=binding.OPT_FACTORS)
            np.testing.assert_array_equal(r.line, whole.line)
            np.testing.assert_array_equal(r.pattern, whole.pattern)
            np.testing.assert_array_equal(r.score, whole.score)
            np.testing.assert_array_equal(r.rank, whole.rank)
            np.testing.assert_array_equal(r.freq_delta, whole.freq_delta)
            assert int(r.log_n_lines[0]) == int(whole.log_n_lines[0])
        end = parallel.strip_trailing_separators(log)
        host = np.frombuffer(log[:end], dtype=np.uint8).copy()
        r = lp.analyze_range_dist_host_ptr(host.ctypes.data, end, 0, 0, prior)
        np.testing.assert_array_equal(r.line, whole.line)
        np.testing.assert_array_equal(r.score, whole.score)
    finally:
        lp.close()


@pytest.mark.parametrize("layout", ["same-device", "all-devices"])
def test_one_process_several_devices_equals_one_device(parser, layout):
    """lp_create_multi / lp_create(cfg, -1): ONE process, a batch cut into one block of logs per device context, counts
    exchanged on the host, one result in batch order; must equal lp_analyze_batch on a single device, repeatedly (the
    estimates of every device settle) and with fewer logs than devices. "same-device" lists device 0 three times, so the
    cut, the exchange and the merge run on a one-GPU box; "all-devices" takes what the box has."""
    wl = synth.make_patterns(40, 4243, noliteral=1)
    logs = [synth.make_log(wl, n, 800 + i, pool_blocks=400, hit_rate=0.06) for i, n in enumerate([700, 40, 1300, 5, 400, 900, 60])]
    prior = (np.arange(len(wl.patterns)) % 9).astype(np.uint64)
    parser.load_patterns(wl.patterns)
    whole = parser.analyze_batch(logs, prior_count=prior, options=binding.OPT_FACTORS)
    if layout == "same-device":
        lp = binding.LogParser(ScoringConfig(), devices=[0, 0, 0])
    else:
        lp = binding.LogParser(ScoringConfig(), device=-1)       # on a one-GPU box: one device context, the plain path
    try:
        lp.load_patterns(wl.patterns)
        for _ in range(3):
            r = lp.analyze_batch(logs, prior_count=prior, options=binding.OPT_FACTORS)
            np.testing.assert_array_equal(r.log_n_lines, whole.log_n_lines)
            np.testing.assert_array_equal(r.log_event_begin, whole.log_event_begin)
            np.testing.assert_array_equal(r.line, whole.line)
            np.testing.assert_array_equal(r.pattern, whole.pattern)
            np.testing.assert_array_equal(r.score, whole.score)
            np.testing.assert_array_equal(r.factors, whole.factors)
            np.testing.assert_array_equal(r.rank, whole.rank)
            np.testing.assert_array_equal(r.freq_delta, whole.freq_delta)
        two = lp.analyze_batch(logs[:2], prior_count=prior)           # fewer logs than devices
        ref2 = parser.analyze_batch(logs[:2], prior_count=prior)
        np.testing.assert_array_equal(two.line, ref2.line)
        np.testing.assert_array_equal(two.score, ref2.score)
        one = lp.analyze(logs[2], prior_count=prior)                   # a single log runs on the first device
        np.testing.assert_array_equal(one.score, parser.analyze(logs[2], prior_count=prior).score)
    finally:
        lp.close()


def test_library_nccl_single_rank_equals_batch(parser):
    """lp_analyze_batch_dist_device with a communicator of one rank: the exchange path (count after the
    unique pass, all-gather and prefix on the second stream, k_score waiting for it, the repeat loop on
    a fresh context) must give exactly lp_analyze_batch's result."""
    import torch

    wl = synth.make_patterns(40, 4242, noliteral=1)
    logs = [synth.make_log(wl, n, 700 + i, pool_blocks=400, hit_rate=0.06) for i, n in enumerate([800, 30, 1500, 400])]
    prior = (np.arange(len(wl.patterns)) % 9).astype(np.uint64)
    parser.load_patterns(wl.patterns)
    whole = parser.analyze_batch(logs, prior_count=prior, options=binding.OPT_FACTORS)
    lp = binding.LogParser(ScoringConfig())
    try:
        lp.load_patterns(wl.patterns)
        lp.comm_init(lp.comm_unique_id(), 0, 1)
        blob = b"".join(logs)
        off = np.zeros(len(logs) + 1, np.uint64)
        np.cumsum([len(x) for x in logs], out=off[1:])
        text = torch.from_numpy(np.frombuffer(blob, dtype=np.uint8).copy()).cuda()
        for _ in range(3):
            r = lp.analyze_batch_dist_device(text.data_ptr(), off, prior, options=binding.OPT_FACTORS)
            np.testing.assert_array_equal(r.line, whole.line)
            np.testing.assert_array_equal(r.pattern, whole.pattern)
            np.testing.assert_array_equal(r.score, whole.score)
            np.testing.assert_array_equal(r.rank, whole.rank)
            np.testing.assert_array_equal(r.freq_delta, whole.freq_delta)
    finally:
        lp.close()


@pytest.mark.parametrize("seed", [11, 12, 13, 14])
def test_medium_logs_all_paths_vs_oracle(parser, oracle, seed):
    """20 000-line logs with a mixed pattern set (literal, class-expanded, anchored forwards/backwards,
    look-around / back-reference, literal-free) and enough events for more than one sort tile: every event,
    score and rank against the oracle."""
    wl = synth.make_patterns(60, 9000 + seed, nonregular_frac=0.12 if seed % 2 else 0.0, scoring_heavy=(seed % 3 == 0), noliteral=2)
    log = synth.make_log(wl, 20_000, 9100 + seed, hit_rate=0.25, symptom_rate=0.2, phase_rate=0.02, pool_blocks=3000)
    if seed == 12:
        log = log.replace(b"\n", b"\r\n", 4000)
    parser.load_patterns(wl.patterns)
    prior = (np.arange(len(wl.patterns)) % 4).astype(np.uint64) * 3
    res = parser.analyze(log, prior_count=prior, options=binding.OPT_FACTORS)
    ora = oracle.analyze(ScoringConfig(), wl.patterns, log, prior_count=prior)
    assert len(ora.line) > 2100          # more than one 2048-key sort tile
    _check(res, ora)
    np.testing.assert_array_equal(res.score, ora.score)
    np.testing.assert_array_equal(res.freq_delta, ora.freq_delta)


def test_many_events_device_radix_path_equals_tile_sort_path(parser):
    """More than 131 072 events in one log: the event list and the ranking go through the device-wide radix
    sorts instead of the tile sort. The same log cut into line-range shards keeps every shard on the tile-sort
    path; both must agree bit for bit (events, scores, merged ranking)."""
    from log_parser_b200 import parallel

    cfg = ScoringConfig()
    wl = synth.make_patterns(30, 6006, noliteral=1)
    log = synth.make_log(wl, 260_000, 6007, hit_rate=0.9, symptom_rate=0.3, pool_blocks=4000)
    parser.load_patterns(wl.patterns)
    whole = parser.analyze(log)
    if whole.n_events <= 131072:
        pytest.skip(f"only {whole.n_events} events")
    results, _ = _range_sharded(cfg, wl.patterns, log, 4)
    assert max(len(r.line) for r in results) < 131072
    np.testing.assert_array_equal(np.concatenate([r.line for r in results]), whole.line)
    np.testing.assert_array_equal(np.concatenate([r.pattern for r in results]), whole.pattern)
    np.testing.assert_array_equal(np.concatenate([r.score for r in results]), whole.score)
    np.testing.assert_array_equal(np.concatenate(parallel.merge_ranks([r.score for r in results])), whole.rank)


def test_random_regexes_with_literals_through_the_pipeline(parser, oracle):
    """Adversarial patterns end to end: random regex fragments (classes, anchors, \\b, lazy / possessive
    quantifiers, look-around, back-references) glued around fixed words, so that the literal prefilter,
    the cover test for anchored verification, the boundary test, the line walkers and the backtracking
    engine all see shapes the synthetic pattern set never produces. Every (line, pattern) event against
    the oracle's find() on that line."""
    from rxgen import Gen

    rng = random.Random(20261107)
    words = ["ERRx", "abc_1", "Exception", "timeout", "x-1:E", "R.E.c"]
    alphabet = "abcERx_1 .-:"
    checked = n_events = 0
    for rnd_i in range(25):
        pats, rxs = [], []
        while len(pats) < 14:
            g = Gen(rng, regular_only=(len(pats) % 3 != 0), max_depth=2)
            a, _pa = g.regex()
            b, _pb = Gen(rng, regular_only=True, max_depth=1).regex()
            w = rng.choice(words)
            wj = "".join("\\" + ch if ch in ".-:" else ch for ch in w)
            shape = rng.randrange(5)
            j = [wj + a, a + wj, a + wj + b, "\\b" + wj + "\\b", "(?:" + a + "|" + wj + b + ")"][shape]
            try:
                o = oracle.Regex(j)
                binding.DebugRegex(j)
            except Exception:
                continue
            pats.append(PatternSpec(id=f"p{len(pats)}", primary_regex=j, confidence=0.5, severity="LOW"))
            rxs.append(o)
        lines = []
        for _ in range(400):
            parts = []
            for _k in range(rng.randint(0, 4)):
                parts.append(rng.choice(words) if rng.random() < 0.5 else "".join(rng.choice(alphabet) for _ in range(rng.randint(0, 6))))
            lines.append("".join(parts))
        log = ("\n".join(lines) + "\n").encode()
        try:
            parser.load_patterns(pats)
        except binding.UnsupportedRegexError:
            continue
        res = parser.analyze(log, options=binding.OPT_NO_RANK)
        got = set(zip(res.line.tolist(), res.pattern.tolist()))
        exp = set()
        for i, ln in enumerate(lines):
            if i >= res.n_lines:
                break
            for p, o in enumerate(rxs):
                if o.find(ln):
                    exp.add((i + 1, p))
        assert got == exp, (rnd_i, sorted(got ^ exp)[:5], [pats[p].primary_regex for _l, p in sorted(got ^ exp)[:3]])
        checked += 1
        n_events += len(exp)
    assert checked >= 20 and n_events > 2000


def test_multi_script_patterns_through_the_pipeline(parser, oracle):
    """Unicode property classes, (?U) classes and word boundaries, (?iu) folding, negated classes that leave out whole
    UTF-8 lead-byte blocks, inside look-arounds, atomic groups and back-references, on lines mixing 1- to 4-byte
    characters, combining marks and Java's extra line terminators: every (line, pattern) event of the device pipeline
    against the oracle's find() on that line. The generator is scripts/fuzz_host_scripts.py (its host-side run is
    tests/test_library_cpu.py::test_multi_script_text_vs_oracle_random)."""
    import importlib.util

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("fuzz_host_scripts", os.path.join(root, "scripts", "fuzz_host_scripts.py"))
    fz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fz)
    rng = random.Random(31)
    words = ["ERRx", "\u0434\u0430\u0442\u0430", "\u0416x1", "\u00e9a", "timeout"]
    checked = n_events = 0
    for rnd_i in range(8):
        pats, rxs = [], []
        while len(pats) < 12:
            a = fz.pattern(rng)
            w = rng.choice(words)
            j = [a, w + a, a + w, "(?U)\\b" + w + "\\b", "(?:" + a + "|" + w + ")", "(?iu)" + w + a][rng.randrange(6)]
            try:
                o = oracle.Regex(j)
                binding.DebugRegex(j)
            except Exception:
                continue
            pats.append(PatternSpec(id=f"p{len(pats)}", primary_regex=j, confidence=0.5, severity="LOW"))
            rxs.append(o)
        made = []
        for _ in range(300):
            made.append("".join(rng.choice(words) if rng.random() < 0.4 else fz.text(rng, 6) for _k in range(rng.randint(0, 3))))
        log = ("\n".join(made) + "\n").encode()
        lines = [log[s:e].decode() for s, e in oracle.split(log)]       # "\r" in front of a "\n" belongs to the separator
        parser.load_patterns(pats)
        res = parser.analyze(log, options=binding.OPT_NO_RANK)
        assert res.n_lines == len(lines)
        got = set(zip(res.line.tolist(), res.pattern.tolist()))
        exp = set()
        for i, ln in enumerate(lines):
            for p, o in enumerate(rxs):
                if o.find(ln):
                    exp.add((i + 1, p))
        assert got == exp, (rnd_i, sorted(got ^ exp)[:5], [(pats[p].primary_regex, lines[l - 1]) for l, p in sorted(got ^ exp)[:3]])
        checked += 1
        n_events += len(exp)
    assert checked == 8 and n_events > 2000


def test_concurrent_requests_from_threads(parser):
    """AnalysisService.analyze is called from several Quarkus worker threads at once (AS:35,50): calls on one
    shared context run on its workers (two by default, each with its own stream and scratch) and overlap on the
    GPU, as do calls on separate contexts; every result must equal the single-threaded one, and results stay
    valid after later calls reused the buffers."""
    import threading

    wl = synth.make_patterns(40, 3131, noliteral=1)
    logs = [synth.make_log(wl, 3000 + 500 * i, 3200 + i, hit_rate=0.05, pool_blocks=500) for i in range(4)]
    parser.load_patterns(wl.patterns)
    expect = [parser.analyze(lg) for lg in logs]
    kept = [(np.array(r.line), np.array(r.pattern), np.array(r.score), np.array(r.rank)) for r in expect]
    second = binding.LogParser(ScoringConfig())
    second.load_patterns(wl.patterns)
    out, errors = {}, []

    def worker(tid, lp):
        try:
            for rep in range(6):
                k = (tid + rep) % len(logs)
                r = lp.analyze(logs[k])
                out[(tid, rep)] = (k, r)
        except Exception as e:      # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(t, parser if t < 3 else second)) for t in range(5)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    second.close()
    assert not errors, errors
    assert len(out) == 30
    for (_tid, _rep), (k, r) in out.items():
        np.testing.assert_array_equal(r.line, kept[k][0])
        np.testing.assert_array_equal(r.pattern, kept[k][1])
        np.testing.assert_array_equal(r.score, kept[k][2])
        np.testing.assert_array_equal(r.rank, kept[k][3])
    for r, kp in zip(expect, kept):        # the first results are still intact (each owns its page-locked buffer)
        np.testing.assert_array_equal(r.score, kp[2])


def test_argument_errors_are_reported_not_computed():
    """Error behaviour at the boundary: bad calls come back as LP_ERR_INVALID with a message, never as a result."""
    import torch

    lp = binding.LogParser(ScoringConfig())
    try:
        with pytest.raises(binding.LpError) as e:
            lp.analyze(b"x\n")                                      # no pattern set loaded
        assert e.value.code == binding.LP_ERR_INVALID and "lp_load_patterns" in str(e.value)
        lp.load_patterns([PatternSpec(id="a", primary_regex="x", confidence=0.5, severity="LOW")])
        dev = torch.zeros(64, dtype=torch.uint8, device="cuda")
        with pytest.raises(binding.LpError) as e:
            lp.analyze_device(dev.data_ptr() + 1, 8)                # device pointer not 16-byte aligned
        assert e.value.code == binding.LP_ERR_INVALID
        with pytest.raises(binding.LpError) as e:
            lp.analyze_batch_device(dev.data_ptr(), np.array([0, 8, 4], np.uint64))     # offsets go backwards
        assert e.value.code == binding.LP_ERR_INVALID
        with pytest.raises(binding.LpError) as e:
            lp.batch_score()                                        # no scan pending
        assert e.value.code == binding.LP_ERR_INVALID
        with pytest.raises(binding.LpError) as e:
            lp.range_score(np.zeros((1, lp.range_summary_words()), np.uint64), 0, 1)    # no range scan pending
        assert e.value.code == binding.LP_ERR_INVALID
        with pytest.raises(binding.LpError) as e:
            lp.analyze_batch_dist_device(dev.data_ptr(), np.array([0, 8], np.uint64))   # lp_comm_init not called
        assert e.value.code == binding.LP_ERR_INVALID
        with pytest.raises(binding.LpError) as e:
            lp.load_patterns([PatternSpec(id="a", primary_regex="x", confidence=0.5, severity="LOW", lines_before=-1, lines_after=2)])
        assert e.value.code == binding.LP_ERR_INVALID
        # the context is still usable after every one of them
        lp.load_patterns([PatternSpec(id="a", primary_regex="x", confidence=0.5, severity="LOW")])
        assert lp.analyze(b"x\ny\n").n_events == 1
    finally:
        lp.close()


def test_shared_pattern_ids_across_shards(parser, oracle):
    """FrequencyTrackingService keys its counts by pattern id (FTS:46-50), so several patterns can feed one
    count. Sharded over ranks, what the lower ranks add to a pattern's count is the sum over every pattern
    with its id -- checked for line-range shards and for the by-log exchange, against the whole-input
    result and the oracle, with a low threshold so that the penalty is in play everywhere."""
    import dataclasses
    import torch
    from log_parser_b200 import parallel

    cfg = ScoringConfig(frequency_threshold=3.0)
    wl = synth.make_patterns(24, 8181, noliteral=1)
    pats = [dataclasses.replace(p, id=f"shared-{k % 4}") for k, p in enumerate(wl.patterns)]
    log = synth.make_log(wl, 6000, 8182, hit_rate=0.08, pool_blocks=900)
    lp = binding.LogParser(cfg)
    lp.load_patterns(pats)
    whole = lp.analyze(log, options=binding.OPT_FACTORS)
    ora = oracle.analyze(cfg, pats, log)
    _check(whole, ora)
    assert (whole.factors[:, 6] > 0).mean() > 0.5
    # one log by line range
    results, _ = _range_sharded(cfg, pats, log, 3)
    np.testing.assert_array_equal(np.concatenate([r.line for r in results]), whole.line)
    np.testing.assert_array_equal(np.concatenate([r.score for r in results]), whole.score)
    # a batch by log: two contexts play two ranks of the device-side exchange
    logs = [synth.make_log(wl, n, 8190 + i, hit_rate=0.08, pool_blocks=400) for i, n in enumerate([700, 300, 900, 500])]
    batch = lp.analyze_batch(logs)
    lps = [binding.LogParser(cfg) for _ in range(2)]
    try:
        for q in lps:
            q.load_patterns(pats)
        n = len(pats) + 1
        gathered = torch.zeros(2 * n, dtype=torch.int64, device="cuda")
        bounds = parallel.shard_bounds([len(x) for x in logs], 2)
        texts, offs = [], []
        for lo, hi in bounds:
            off = np.zeros(hi - lo + 1, np.uint64)
            np.cumsum([len(x) for x in logs[lo:hi]], out=off[1:])
            texts.append(torch.from_numpy(np.frombuffer(b"".join(logs[lo:hi]), dtype=np.uint8).copy()).cuda())
            offs.append(off)
        for _round in range(8):
            for r in range(2):
                lps[r].batch_scan_device_async(texts[r].data_ptr(), offs[r], gathered[r * n:].data_ptr())
            torch.cuda.synchronize()
            res = [lps[r].batch_score_gathered(gathered.data_ptr(), r, 2) for r in range(2)]
            if res[0] is not None and res[1] is not None:
                break
        np.testing.assert_array_equal(np.concatenate([r.score for r in res]), batch.score)
        # and with the counts passing through the host
        eng = parallel.GpuBatchEngine(lps[0])
        prior = np.zeros(len(pats), np.uint64)
        scores = []
        for lo, hi in bounds:
            counts = eng.scan(logs[lo:hi])
            scores.append(eng.score(prior).score)
            prior = prior + eng.by_id(counts)
        np.testing.assert_array_equal(np.concatenate(scores), batch.score)
    finally:
        for q in lps:
            q.close()
        lp.close()


def test_context_flags_with_many_scan_groups(oracle):
    """More literal-free regexes than one k_dfa_scan launch holds (4 groups): the stack-trace regex of
    CAS:27-34 then sits in a later launch, whose flag bits must be added to the bytes the first launch
    stored (they were once dropped). The case is seed 64 of scripts/fuzz_scoring_vs_oracle.py, which found
    it: 5 scan groups, the stack-trace regex in the fifth; two more seeds ride along."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
    from fuzz_scoring_vs_oracle import gen_case

    for seed in (64, 18, 5):
        cfg, pats, log, prior = gen_case(seed)
        lp = binding.LogParser(cfg)
        try:
            lp.load_patterns(pats)
            res = lp.analyze(log, prior_count=prior, options=binding.OPT_FACTORS)
        finally:
            lp.close()
        ora = oracle.analyze(cfg, pats, log, prior_count=prior)
        _check(res, ora)
        np.testing.assert_array_equal(res.score, ora.score)
