"""Pattern-set loading: YAML files of a pattern directory -> flat PatternSpec list.

Mirrors PatternService.loadPatterns (PatternService.java:45-85): every regular file ending in
.yml/.yaml under `pattern.directory`, in directory-walk order, parsed into one PatternSet; files
that fail to parse are logged and skipped; a missing directory yields zero sets. The YAML key
names follow the common library's snake_case model (`primary_pattern.regex/confidence` are pinned
by docs/SCORING_ALGORITHM.md:29-33; the rest are inferred from the getters used in
AnalysisService.java:57-80,104,175 and ScoringService.java:65-319 — the model source is absent).
"""
from __future__ import annotations

import logging
import os
from dataclasses import dataclass, field
from typing import List, Optional

import yaml

from .model import PatternSpec, SecondaryPattern, SequencePattern

log = logging.getLogger(__name__)


@dataclass
class PatternSet:
    library_id: Optional[str]
    patterns: Optional[List[dict]]          # raw pattern dicts, kept for the response (matchedPattern)
    source: str = ""
    metadata: dict = field(default_factory=dict)


def load_pattern_file(path: str) -> PatternSet:
    with open(path, "r", encoding="utf-8") as f:
        doc = yaml.safe_load(f) or {}
    if not isinstance(doc, dict):
        raise ValueError("pattern set must be a mapping")
    meta = doc.get("metadata") or {}
    return PatternSet(library_id=meta.get("library_id"), patterns=doc.get("patterns"), source=path, metadata=meta)


def load_pattern_directory(directory: str) -> List[PatternSet]:
    """PatternService.java:45-69. Order = walk order (Files.walk is unsorted directory order)."""
    if not os.path.isdir(directory):
        log.error("Pattern directory does not exist or is not a directory: %s", directory)   # PS:50-55
        return []
    sets: List[PatternSet] = []
    for root, _dirs, files in os.walk(directory):
        for name in files:
            if not (name.endswith(".yml") or name.endswith(".yaml")):
                continue
            path = os.path.join(root, name)
            if not os.path.isfile(path):
                continue
            try:
                sets.append(load_pattern_file(path))
            except Exception as e:   # noqa: BLE001  (PS:82-84: log and skip)
                log.error("Failed to parse pattern file: %s (%s)", path, e)
    log.info("Successfully loaded %d pattern sets.", len(sets))
    return sets


def pattern_from_dict(d: dict, set_index: int, library_id: Optional[str]) -> PatternSpec:
    prim = d.get("primary_pattern") or {}
    ctx = d.get("context_extraction")
    secs = [SecondaryPattern(s["regex"], float(s.get("weight", 0.0)), int(s.get("proximity_window", 0)))
            for s in (d.get("secondary_patterns") or [])]
    seqs = []
    for q in (d.get("sequence_patterns") or []):
        events = q.get("events")
        seqs.append(SequencePattern([e["regex"] for e in events] if events else [], float(q.get("bonus_multiplier", 0.0)),
                                    q.get("description", "") or ""))
    return PatternSpec(
        id=d.get("id"), name=d.get("name", "") or "", primary_regex=prim["regex"], confidence=float(prim.get("confidence", 0.0)),
        severity=d.get("severity"), set_index=set_index, library_id=library_id or "",
        lines_before=int(ctx.get("lines_before", 0)) if ctx is not None else None,
        lines_after=int(ctx.get("lines_after", 0)) if ctx is not None else None,
        secondaries=secs, sequences=seqs)


def flatten(pattern_sets: List[PatternSet]):
    """(set, pattern) order of AnalysisService.java:91-92; sets with `patterns: null` are skipped
    (AS:57-59). Returns (specs, raw pattern dicts) index-aligned with lp_result_view.pattern."""
    specs: List[PatternSpec] = []
    raw: List[dict] = []
    for si, ps in enumerate(pattern_sets):
        if ps.patterns is None:
            continue
        for d in ps.patterns:
            specs.append(pattern_from_dict(d, si, ps.library_id))
            raw.append(d)
    return specs, raw
