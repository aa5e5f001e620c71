"""B200-native hot path of podmortem/log-parser: multi-pattern regex scan + confidence scoring.

The product is log_parser_b200/liblogparser_b200.so (hand-written CUDA for sm_100a behind the C
ABI of include/logparser_b200.h). This package is the host-side mirror of the reference's
AnalysisService interface used by the tests and the benchmark; it has no CPU fallback.
"""
from .model import PatternSpec, ScoringConfig, SecondaryPattern, SequencePattern, severity_multiplier  # noqa: F401

__all__ = ["PatternSpec", "ScoringConfig", "SecondaryPattern", "SequencePattern", "severity_multiplier"]
