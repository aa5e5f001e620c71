"""N1/N2 host pieces on CPU: YAML pattern loading in (set, pattern) order and AnalysisResult
assembly (context slicing AS:132-156, metadata AS:166-180, summary AS:188-215). The event columns
come from the oracle here; on a GPU box AnalysisService.analyze feeds the same assembly."""
import os

import numpy as np
import yaml

from log_parser_b200 import patterns, service
from log_parser_b200.model import ScoringConfig

SET_A = {"metadata": {"library_id": "lib-a"}, "patterns": [
    {"id": "oom", "name": "OOM", "severity": "critical", "primary_pattern": {"regex": "OutOfMemoryError", "confidence": 0.9},
     "secondary_patterns": [{"regex": "GC overhead", "weight": 0.6, "proximity_window": 20}],
     "sequence_patterns": [{"description": "heap grows", "bonus_multiplier": 0.3, "events": [{"regex": "heap"}, {"regex": "OutOfMemory"}]}],
     "context_extraction": {"lines_before": 1, "lines_after": 2, "include_stack_trace": True}},
    {"id": "warn", "name": "Warn", "severity": "LOW", "primary_pattern": {"regex": "\\bWARN\\b", "confidence": 0.4}},
]}
SET_B = {"metadata": {"library_id": "lib-b"}, "patterns": None}
SET_C = {"metadata": {"library_id": "lib-c"}, "patterns": [
    {"id": "odd", "name": "Odd", "severity": "weird", "primary_pattern": {"regex": "^x$", "confidence": 0.5},
     "context_extraction": {"lines_before": 0, "lines_after": 0}}]}


def _write(tmp_path):
    for name, doc in [("a.yml", SET_A), ("b.yaml", SET_B), ("sub/c.yml", SET_C), ("notes.txt", {"x": 1})]:
        p = tmp_path / name
        p.parent.mkdir(parents=True, exist_ok=True)
        p.write_text(yaml.safe_dump(doc) if name != "notes.txt" else "ignored")
    (tmp_path / "broken.yml").write_text("patterns: [unclosed")


def test_loader_and_flatten(tmp_path):
    _write(tmp_path)
    sets = patterns.load_pattern_directory(str(tmp_path))
    assert sorted(s.library_id for s in sets) == ["lib-a", "lib-b", "lib-c"]     # broken.yml skipped, notes.txt ignored
    specs, raw = patterns.flatten(sets)
    assert len(specs) == 3 and len(raw) == 3
    by_id = {s.id: s for s in specs}
    assert by_id["oom"].lines_before == 1 and by_id["oom"].lines_after == 2 and by_id["oom"].has_context
    assert not by_id["warn"].has_context
    assert by_id["oom"].secondaries[0].proximity_window == 20 and by_id["oom"].sequences[0].events == ["heap", "OutOfMemory"]
    # order inside a set is list order; sets in directory-walk order
    order = [s.library_id for s in specs]
    assert order == sorted(order, key=[x.library_id for x in sets].index)
    assert patterns.load_pattern_directory(str(tmp_path / "missing")) == []


def test_result_assembly_matches_reference_shape(tmp_path, oracle):
    _write(tmp_path)
    sets = [patterns.load_pattern_file(str(tmp_path / "a.yml")), patterns.load_pattern_file(str(tmp_path / "b.yaml")),
            patterns.load_pattern_file(str(tmp_path / "sub/c.yml"))]
    specs, raw = patterns.flatten(sets)
    logs = "start heap 1\r\nWARN slow\nx\nboom java.lang.OutOfMemoryError: GC overhead\nafter1\nafter2\nafter3\n\n\n"
    ora = oracle.analyze(ScoringConfig(), specs, logs.encode())
    res = service.assemble_result(ora.line, ora.pattern, ora.score, logs, raw, sets, 0.0, ora.n_lines)
    assert res["metadata"]["totalLines"] == 7 and res["metadata"]["patternsUsed"] == ["lib-a", "lib-b", "lib-c"]
    ev = res["events"]
    assert [(e["lineNumber"], e["matchedPattern"]["id"]) for e in ev] == [(2, "warn"), (3, "odd"), (4, "oom")]
    assert "linesBefore" not in ev[0]["context"] and ev[0]["context"]["matchedLine"] == "WARN slow"
    assert ev[1]["context"] == {"matchedLine": "x", "linesBefore": [], "linesAfter": []}
    assert ev[2]["context"]["linesBefore"] == ["x"] and ev[2]["context"]["linesAfter"] == ["after1", "after2"]
    assert res["summary"] == {"significantEvents": 3, "severityDistribution": {"LOW": 1, "WEIRD": 1, "CRITICAL": 1}, "highestSeverity": "CRITICAL"}
    empty = service.assemble_result(np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0), "\n\n", raw, sets, 0.0, 0)
    assert empty["summary"] == {"significantEvents": 0, "highestSeverity": "NONE", "severityDistribution": {}}


def test_highest_severity_keeps_first_of_equal_maxima():
    """AS:204-209: Stream.max(comparingInt(order::indexOf)); unknown severities all compare as -1 and the first one wins."""
    raw = [{"id": "a", "severity": "strange"}, {"id": "b", "severity": "odd"}, {"id": "c", "severity": "info"}]
    sets = []
    res = service.assemble_result(np.array([1, 2], np.uint32), np.array([0, 1], np.uint32), np.array([1.0, 1.0]), "x\ny\n", raw, sets, 0.0, 2)
    assert res["summary"]["highestSeverity"] == "STRANGE"
    res = service.assemble_result(np.array([1, 2, 2], np.uint32), np.array([1, 0, 2], np.uint32), np.array([1.0, 1.0, 1.0]), "x\ny\n", raw, sets, 0.0, 2)
    assert res["summary"]["highestSeverity"] == "INFO"          # INFO (index 0) beats the unknowns (-1)
    res = service.assemble_result(np.array([1, 2], np.uint32), np.array([1, 0], np.uint32), np.array([1.0, 1.0]), "x\ny\n", raw, sets, 0.0, 2)
    assert res["summary"]["highestSeverity"] == "ODD"


def test_frequency_tracker_window():
    now = [1000.0]
    ft = service.FrequencyTracker(ScoringConfig(time_window_hours=1), clock=lambda: now[0])
    ft.record("p", 5)
    ft.record("  ", 3)
    ft.record(None, 3)
    assert ft.count_in_window("p") == 5 and ft.count_in_window("  ") == 0 and ft.count_in_window(None) == 0
    now[0] += 1800
    ft.record("p", 2)
    assert ft.count_in_window("p") == 7
    now[0] += 1801
    assert ft.count_in_window("p") == 2          # the first five left the one-hour window
    assert ft.get_frequency_stats() == {"p": 2}
    ft.reset("p")
    assert ft.count_in_window("p") == 0
