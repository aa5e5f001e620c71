"""Plain data carriers for the hot path: the pattern descriptor that crosses the C-ABI and the
`scoring.*` configuration.

These mirror the reference's POJOs as far as the hot path reads them
(com.redhat.podmortem.common.model.pattern.* — un-vendored; getter names taken from
AnalysisService.java:57-80,104 and ScoringService.java:65,69,163,178,201,210,214,232,319) and the
nine `scoring.*` keys of src/main/resources/application.properties:5-20.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

# ScoringService.java:30-36; lookup key is severity.toUpperCase(), default 1.0 (SS:68-69)
SEVERITY_MULTIPLIERS = {"CRITICAL": 5.0, "HIGH": 3.0, "MEDIUM": 2.0, "LOW": 1.5, "INFO": 1.0}


def severity_multiplier(severity: Optional[str]) -> float:
    if severity is None:
        raise ValueError("severity is null (the reference throws NullPointerException at SS:69)")
    return SEVERITY_MULTIPLIERS.get(severity.upper(), 1.0)


@dataclass
class ScoringConfig:
    """application.properties:5-20 with the @ConfigProperty defaults of SS:38-51, CAS:24, FTS:27-34."""

    decay_constant: float = 10.0
    max_window: int = 100
    early_bonus_threshold: float = 0.2
    max_early_bonus: float = 2.5
    penalty_threshold: float = 0.5
    max_context_factor: float = 2.5
    frequency_threshold: float = 10.0
    max_penalty: float = 0.8
    time_window_hours: int = 1


@dataclass
class SecondaryPattern:
    regex: str
    weight: float
    proximity_window: int


@dataclass
class SequencePattern:
    events: List[str]
    bonus_multiplier: float
    description: str = ""


@dataclass
class PatternSpec:
    id: Optional[str]
    primary_regex: str
    confidence: float
    severity: str = "INFO"
    name: str = ""
    set_index: int = 0
    library_id: str = ""
    # context_extraction == None mirrors `rules == null` at AS:137-139
    lines_before: Optional[int] = None
    lines_after: Optional[int] = None
    secondaries: List[SecondaryPattern] = field(default_factory=list)
    sequences: List[SequencePattern] = field(default_factory=list)

    @property
    def has_context(self) -> bool:
        return self.lines_before is not None and self.lines_after is not None
