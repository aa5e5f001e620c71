"""Host-side mirror of the reference's service seam, for tests and examples:
`AnalysisService.analyze(PodFailureData) -> AnalysisResult` with the hot loop
(AnalysisService.java:53-113) replaced by one lp_analyze call, the frequency state kept on the host
(FrequencyTrackingService.java:25-56, 101-161), and the response assembled as the reference does
(AnalysisService.java:115-121, 132-156, 166-215). JSON field names follow the Java bean getters.
"""
from __future__ import annotations

import time
import uuid
from collections import deque
from datetime import datetime, timezone
from typing import Callable, Dict, List, Optional

import numpy as np

from .model import ScoringConfig
from .patterns import PatternSet, flatten

SEVERITY_ORDER = ["INFO", "LOW", "MEDIUM", "HIGH", "CRITICAL"]   # AS:207


class FrequencyTracker:
    """FrequencyTrackingService: per pattern id, the matches inside a sliding window of
    `time_window_hours` (PatternFrequency of the absent common library; SURVEY §8a-F model)."""

    def __init__(self, cfg: ScoringConfig, clock: Callable[[], float] = time.time):
        self.cfg, self.clock = cfg, clock
        self._hits: Dict[str, deque] = {}

    def _expire(self, q: deque, now: float):
        horizon = now - self.cfg.time_window_hours * 3600.0
        while q and q[0] <= horizon:
            q.popleft()

    def count_in_window(self, pattern_id: Optional[str]) -> int:
        if pattern_id is None or not pattern_id.strip():
            return 0
        q = self._hits.get(pattern_id)
        if not q:
            return 0
        self._expire(q, self.clock())
        return len(q)

    def record(self, pattern_id: Optional[str], n: int = 1):                      # FTS:41-56
        if pattern_id is None or not pattern_id.strip() or n <= 0:
            return
        now = self.clock()
        q = self._hits.setdefault(pattern_id, deque())
        q.extend([now] * n)

    def get_frequency_stats(self) -> Dict[str, int]:                              # FTS:101-131
        return {k: self.count_in_window(k) for k in list(self._hits)}

    def reset(self, pattern_id: Optional[str] = None):                             # FTS:137-161
        if pattern_id is None:
            self._hits.clear()
        else:
            self._hits.pop(pattern_id, None)


def extract_context(lines: List[str], index: int, raw_pattern: dict) -> dict:
    """AnalysisService.java:132-156."""
    ctx = {"matchedLine": lines[index]}
    rules = raw_pattern.get("context_extraction")
    if rules is None:
        return ctx
    before, after = int(rules.get("lines_before", 0)), int(rules.get("lines_after", 0))
    ctx["linesBefore"] = lines[max(0, index - before):index]
    ctx["linesAfter"] = lines[index + 1:min(len(lines), index + 1 + after)]
    return ctx


def java_split_lines(logs: str) -> List[str]:
    """String.split("\\\\r?\\\\n") (AS:53): trailing empty strings removed, no separator -> [input]."""
    import re

    parts = re.split(r"\r?\n", logs)
    if len(parts) == 1:
        return parts
    while parts and parts[-1] == "":
        parts.pop()
    return parts


def assemble_result(line: np.ndarray, pattern: np.ndarray, score: np.ndarray, logs: str, raw_patterns: List[dict],
                    pattern_sets: List[PatternSet], start_time: float, total_lines: int) -> dict:
    """AnalysisResult: events in discovery order (never sorted, AS:109,116), metadata (AS:166-180),
    summary (AS:188-215)."""
    lines = java_split_lines(logs)
    assert len(lines) == total_lines, (len(lines), total_lines)
    events = []
    for ln, p, sc in zip(line.tolist(), pattern.tolist(), score.tolist()):
        events.append({"lineNumber": int(ln), "matchedPattern": raw_patterns[p],
                       "context": extract_context(lines, int(ln) - 1, raw_patterns[p]), "score": float(sc)})
    result = {"events": events, "analysisId": str(uuid.uuid4())}
    result["metadata"] = {
        "processingTimeMs": int((time.time() - start_time) * 1000),
        "totalLines": total_lines,
        "analyzedAt": datetime.now(timezone.utc).isoformat().replace("+00:00", "Z"),
        "patternsUsed": [ps.library_id for ps in pattern_sets],
    }
    if not events:
        result["summary"] = {"significantEvents": 0, "highestSeverity": "NONE", "severityDistribution": {}}
        return result
    sev = [str(e["matchedPattern"].get("severity")).upper() for e in events]
    dist: Dict[str, int] = {}
    for s in sev:
        dist[s] = dist.get(s, 0) + 1
    rank = lambda s: SEVERITY_ORDER.index(s) if s in SEVERITY_ORDER else -1   # List.indexOf -> -1 for unknown severities
    best = sev[0]
    for s in sev[1:]:
        if rank(s) > rank(best):    # Stream.max reduces with BinaryOperator.maxBy: (a, b) -> cmp(a, b) >= 0 ? a : b, the FIRST of equal maxima stays
            best = s
    result["summary"] = {"significantEvents": len(events), "severityDistribution": dist, "highestSeverity": best}
    return result


class AnalysisService:
    """`analyze(PodFailureData) -> AnalysisResult` backed by liblogparser_b200.so."""

    def __init__(self, pattern_sets: List[PatternSet], cfg: Optional[ScoringConfig] = None, device: int = 0,
                 clock: Callable[[], float] = time.time):
        from .binding import LogParser

        self.cfg = cfg or ScoringConfig()
        self.pattern_sets = pattern_sets
        self.specs, self.raw = flatten(pattern_sets)
        self.freq = FrequencyTracker(self.cfg, clock)
        self.parser = LogParser(self.cfg, device=device)
        self.parser.load_patterns(self.specs)          # PatternSyntaxError here == HTTP 500 on every request in the reference

    def analyze(self, data: dict) -> dict:
        start = time.time()
        logs = data["logs"]                             # NPE -> 500 in the reference when null (AS:53)
        prior = np.array([self.freq.count_in_window(s.id) for s in self.specs], dtype=np.uint64)
        res = self.parser.analyze(logs.encode("utf-8"), prior_count=prior)
        for s, n in zip(self.specs, res.freq_delta.tolist()):
            self.freq.record(s.id, int(n))              # SS:88, once per event
        return assemble_result(res.line, res.pattern, res.score, logs, self.raw, self.pattern_sets, start, res.n_lines)
