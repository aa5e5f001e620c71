"""The Java FFM binding (java/.../LogParserB200.java) cannot be compiled here (no JDK): at least its hand-written
struct offsets are checked against the ctypes mirror of include/logparser_b200.h that the test-suite binds with."""
import ctypes as C
import os
import re

from log_parser_b200 import binding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
JAVA = os.path.join(ROOT, "java", "com", "redhat", "podmortem", "parser", "gpu", "LogParserB200.java")


def _offsets(struct):
    return {name: getattr(struct, name).offset for name, _ in struct._fields_}


def test_pattern_desc_offsets_match():
    src = open(JAVA).read()
    off = _offsets(binding.PatternDesc)
    assert C.sizeof(binding.PatternDesc) == 72 and C.sizeof(binding.SecondaryDesc) == 24 and C.sizeof(binding.SequenceDesc) == 24
    expect = {"id": r"d\.set\(ADDRESS, %d, p\.getId\(\)", "primary_regex": r"d\.set\(ADDRESS, %d, a\.allocateUtf8String\(p\.getPrimaryPattern",
              "confidence": r"d\.set\(JAVA_DOUBLE, %d, p\.getPrimaryPattern\(\)\.getConfidence", "severity_multiplier": r"d\.set\(JAVA_DOUBLE, %d, SEVERITY_MULTIPLIERS",
              "has_context": r"d\.set\(JAVA_INT, %d, ctxRules \? 1 : 0\)", "lines_before": r"d\.set\(JAVA_INT, %d, ctxRules \? p\.getContextExtraction\(\)\.getLinesBefore",
              "lines_after": r"d\.set\(JAVA_INT, %d, ctxRules \? p\.getContextExtraction\(\)\.getLinesAfter", "n_secondaries": r"d\.set\(JAVA_INT, %d, secs\.size\(\)\)",
              "secondaries": r"d\.set\(ADDRESS, %d, sa\)", "n_sequences": r"d\.set\(JAVA_INT, %d, seqs\.size\(\)\)", "sequences": r"d\.set\(ADDRESS, %d, qa\)"}
    for field, pat in expect.items():
        assert re.search(pat % off[field], src), (field, off[field])
    so, qo = _offsets(binding.SecondaryDesc), _offsets(binding.SequenceDesc)
    assert (so["regex"], so["weight"], so["proximity_window"]) == (0, 8, 16) and "sa.set(JAVA_INT, o + 16" in src
    assert (qo["event_regex"], qo["n_events"], qo["bonus_multiplier"]) == (0, 8, 16) and "qa.set(JAVA_DOUBLE, o + 16" in src


def test_result_view_and_cfg_offsets_match():
    src = open(JAVA).read()
    v = _offsets(binding.ResultView)
    assert C.sizeof(binding.ResultView) == 136 and v["budget_exceeded"] == 128 and "v.get(JAVA_LONG, 128)" in src
    for field, pat in {"n_events": r"v\.get\(JAVA_LONG, %d\)", "log_n_lines": r"v\.get\(ADDRESS, %d\)\.reinterpret\(8\)", "line": r"int\[\] line = v\.get\(ADDRESS, %d\)",
                       "pattern": r"int\[\] pat = v\.get\(ADDRESS, %d\)", "score": r"double\[\] score = v\.get\(ADDRESS, %d\)", "rank": r"int\[\] rank = v\.get\(ADDRESS, %d\)",
                       "freq_delta": r"long\[\] delta = v\.get\(ADDRESS, %d\)"}.items():
        assert re.search(pat % v[field], src), (field, v[field])
    c = _offsets(binding.ScoringCfg)
    for field, kind in [("decay_constant", "JAVA_DOUBLE"), ("max_window", "JAVA_INT"), ("time_window_hours", "JAVA_INT"), ("early_bonus_threshold", "JAVA_DOUBLE"),
                        ("max_early_bonus", "JAVA_DOUBLE"), ("penalty_threshold", "JAVA_DOUBLE"), ("max_context_factor", "JAVA_DOUBLE"),
                        ("frequency_threshold", "JAVA_DOUBLE"), ("max_penalty", "JAVA_DOUBLE"), ("decay_lut", "ADDRESS")]:
        assert "cfg.set(%s, %d," % (kind, c[field]) in src, (field, c[field])


def test_every_bound_symbol_is_exported():
    src = open(JAVA).read()
    for name in re.findall(r'h\("(lp_[a-z_]+)"', src):
        assert name in binding.EXPORTS, name
