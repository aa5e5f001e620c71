"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes loader for oracle/_build/liboracle.so (the CPU restatement of the reference hot path).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module; the product package (log_parser_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")

F_UNIX_LINES, F_CASE_INSENSITIVE, F_COMMENTS, F_MULTILINE, F_DOTALL = 0x01, 0x02, 0x04, 0x08, 0x20
OPT_MEMO, OPT_NO_FIRST_OPT = 1, 2


class OrcCfg(C.Structure):
    _fields_ = [
        ("decay_constant", C.c_double), ("max_window", C.c_int32), ("time_window_hours", C.c_int32),
        ("early_bonus_threshold", C.c_double), ("max_early_bonus", C.c_double),
        ("penalty_threshold", C.c_double), ("max_context_factor", C.c_double),
        ("frequency_threshold", C.c_double), ("max_penalty", C.c_double),
    ]


class OrcSecondary(C.Structure):
    _fields_ = [("regex", C.c_char_p), ("weight", C.c_double), ("proximity_window", C.c_int32)]


class OrcSequence(C.Structure):
    _fields_ = [("events", C.POINTER(C.c_char_p)), ("n_events", C.c_uint32), ("bonus_multiplier", C.c_double)]


class OrcPattern(C.Structure):
    _fields_ = [
        ("id", C.c_char_p), ("primary_regex", C.c_char_p), ("confidence", C.c_double),
        ("severity_multiplier", C.c_double), ("has_context", C.c_int32), ("lines_before", C.c_int32),
        ("lines_after", C.c_int32), ("n_secondaries", C.c_uint32), ("secondaries", C.POINTER(OrcSecondary)),
        ("n_sequences", C.c_uint32), ("sequences", C.POINTER(OrcSequence)),
    ]


class OrcResult(C.Structure):
    _fields_ = [
        ("n_lines", C.c_uint64), ("n_events", C.c_uint64), ("line", C.POINTER(C.c_uint32)),
        ("pattern", C.POINTER(C.c_uint32)), ("score", C.POINTER(C.c_double)),
        ("factors", C.POINTER(C.c_double)), ("freq_delta", C.POINTER(C.c_uint64)), ("n_patterns", C.c_uint64),
    ]


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (g++)."""
    if force or not os.path.exists(_SO) or any(
        os.path.getmtime(os.path.join(_HERE, f)) > os.path.getmtime(_SO)
        for f in ("jregex.cpp", "jregex.hpp", "analyze.cpp", "oracle_api.h", "Makefile", "unicode_tables.inc")
    ):
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.orc_regex_compile.restype = C.c_void_p
        L.orc_regex_compile.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_int)]
        L.orc_regex_free.argtypes = [C.c_void_p]
        L.orc_regex_find.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        L.orc_regex_find_span.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_regex_set_first_opt.argtypes = [C.c_void_p, C.c_int]
        L.orc_split.restype = C.c_uint64
        L.orc_split.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_uint64]
        L.orc_chrono.restype = C.c_double
        L.orc_chrono.argtypes = [C.POINTER(OrcCfg), C.c_uint64, C.c_uint64]
        L.orc_context_factor.restype = C.c_double
        L.orc_context_factor.argtypes = [C.POINTER(OrcCfg), C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.c_size_t]
        L.orc_frequency_penalty.restype = C.c_double
        L.orc_frequency_penalty.argtypes = [C.POINTER(OrcCfg), C.c_int, C.c_double]
        L.orc_proximity_term.restype = C.c_double
        L.orc_proximity_term.argtypes = [C.POINTER(OrcCfg), C.c_double, C.c_double]
        L.orc_analyze.restype = C.c_int
        L.orc_analyze.argtypes = [C.POINTER(OrcCfg), C.POINTER(OrcPattern), C.c_size_t, C.c_void_p, C.c_size_t,
                                  C.POINTER(C.c_uint64), C.c_int, C.POINTER(C.POINTER(OrcResult)), C.c_char_p, C.c_size_t]
        L.orc_result_free.argtypes = [C.POINTER(OrcResult)]
        L.orc_analyze_many.restype = C.c_int64
        L.orc_analyze_many.argtypes = [C.POINTER(OrcCfg), C.POINTER(OrcPattern), C.c_size_t, C.POINTER(C.c_void_p),
                                       C.POINTER(C.c_size_t), C.c_size_t, C.c_int, C.c_int, C.POINTER(C.c_uint64),
                                       C.POINTER(C.c_double), C.c_char_p, C.c_size_t]
        _lib = L
    return _lib


class OracleSyntaxError(ValueError):
    pass


class OracleUnsupported(ValueError):
    pass


class Regex:
    def __init__(self, pattern: str, flags: int = 0):
        err = C.create_string_buffer(512)
        kind = C.c_int(0)
        self._h = lib().orc_regex_compile(pattern.encode("utf-8"), flags, err, 512, C.byref(kind))
        if not self._h:
            msg = err.value.decode("utf-8", "replace")
            if kind.value == 2:
                raise OracleUnsupported(msg)
            raise OracleSyntaxError(msg)

    def find(self, text) -> bool:
        b = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        return bool(lib().orc_regex_find(self._h, b, len(b)))

    def span(self, text):
        b = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        s, e = C.c_int(-1), C.c_int(-1)
        ok = lib().orc_regex_find_span(self._h, b, len(b), C.byref(s), C.byref(e))
        return (s.value, e.value) if ok else None

    def set_first_opt(self, on: bool):
        lib().orc_regex_set_first_opt(self._h, 1 if on else 0)

    def __del__(self):
        try:
            if self._h:
                lib().orc_regex_free(self._h)
        except Exception:
            pass


def make_cfg(cfg) -> OrcCfg:
    return OrcCfg(cfg.decay_constant, cfg.max_window, cfg.time_window_hours, cfg.early_bonus_threshold,
                  cfg.max_early_bonus, cfg.penalty_threshold, cfg.max_context_factor,
                  cfg.frequency_threshold, cfg.max_penalty)


class PatternTable:
    """Flattens PatternSpec objects (log_parser_b200.model) into orc_pattern[] and keeps the buffers alive."""

    def __init__(self, patterns: Sequence):
        from log_parser_b200.model import severity_multiplier

        self.n = len(patterns)
        self._keep = []
        arr = (OrcPattern * max(1, self.n))()
        for i, p in enumerate(patterns):
            secs = (OrcSecondary * max(1, len(p.secondaries)))()
            for j, s in enumerate(p.secondaries):
                secs[j] = OrcSecondary(s.regex.encode("utf-8"), s.weight, s.proximity_window)
            seqs = (OrcSequence * max(1, len(p.sequences)))()
            for j, s in enumerate(p.sequences):
                ev = (C.c_char_p * max(1, len(s.events)))(*[e.encode("utf-8") for e in s.events])
                self._keep.append(ev)
                seqs[j] = OrcSequence(C.cast(ev, C.POINTER(C.c_char_p)), len(s.events), s.bonus_multiplier)
            self._keep += [secs, seqs]
            arr[i] = OrcPattern(
                None if p.id is None else p.id.encode("utf-8"), p.primary_regex.encode("utf-8"), p.confidence,
                severity_multiplier(p.severity), 1 if p.has_context else 0,
                p.lines_before if p.has_context else 0, p.lines_after if p.has_context else 0,
                len(p.secondaries), C.cast(secs, C.POINTER(OrcSecondary)),
                len(p.sequences), C.cast(seqs, C.POINTER(OrcSequence)))
        self.arr = arr


@dataclass
class OracleResult:
    n_lines: int
    line: np.ndarray      # uint32, 1-based
    pattern: np.ndarray   # uint32
    score: np.ndarray     # float64
    factors: np.ndarray   # [M,7]
    freq_delta: np.ndarray


def split(log: bytes):
    n = lib().orc_split(log, len(log), None, None, 0)
    st = (C.c_uint64 * max(1, n))()
    en = (C.c_uint64 * max(1, n))()
    lib().orc_split(log, len(log), st, en, n)
    return [(st[i], en[i]) for i in range(n)]


def chrono(cfg, idx: int, n: int) -> float:
    c = make_cfg(cfg)
    return lib().orc_chrono(C.byref(c), idx, n)


def context_factor(cfg, lines: List[bytes]) -> float:
    c = make_cfg(cfg)
    arr = (C.c_char_p * max(1, len(lines)))(*lines)
    lens = (C.c_size_t * max(1, len(lines)))(*[len(x) for x in lines])
    return lib().orc_context_factor(C.byref(c), arr, lens, len(lines))


def frequency_penalty(cfg, exists: bool, count: float) -> float:
    c = make_cfg(cfg)
    return lib().orc_frequency_penalty(C.byref(c), 1 if exists else 0, count)


def proximity_term(cfg, weight: float, distance: float) -> float:
    c = make_cfg(cfg)
    return lib().orc_proximity_term(C.byref(c), weight, distance)


def analyze(cfg, patterns: Sequence, log: bytes, prior_count: Optional[Sequence[int]] = None,
            memo: bool = True, table: Optional[PatternTable] = None) -> OracleResult:
    t = table or PatternTable(patterns)
    c = make_cfg(cfg)
    out = C.POINTER(OrcResult)()
    err = C.create_string_buffer(1024)
    prior = None
    if prior_count is not None:
        prior = (C.c_uint64 * max(1, t.n))(*[int(x) for x in prior_count])
    buf = np.frombuffer(log, dtype=np.uint8) if not isinstance(log, np.ndarray) else log
    rc = lib().orc_analyze(C.byref(c), t.arr, t.n, buf.ctypes.data if len(buf) else None, len(buf), prior,
                           OPT_MEMO if memo else 0, C.byref(out), err, 1024)
    if rc != 0:
        msg = err.value.decode("utf-8", "replace")
        if rc == -2:
            raise OracleSyntaxError(msg)
        if rc == -3:
            raise OracleUnsupported(msg)
        raise RuntimeError(msg)
    r = out.contents
    m = r.n_events
    res = OracleResult(
        n_lines=r.n_lines,
        line=np.ctypeslib.as_array(r.line, (m,)).copy() if m else np.zeros(0, np.uint32),
        pattern=np.ctypeslib.as_array(r.pattern, (m,)).copy() if m else np.zeros(0, np.uint32),
        score=np.ctypeslib.as_array(r.score, (m,)).copy() if m else np.zeros(0, np.float64),
        factors=np.ctypeslib.as_array(r.factors, (m * 7,)).copy().reshape(m, 7) if m else np.zeros((0, 7)),
        freq_delta=np.ctypeslib.as_array(r.freq_delta, (t.n,)).copy() if t.n else np.zeros(0, np.uint64),
    )
    lib().orc_result_free(out)
    return res


def analyze_many(cfg, patterns: Sequence, logs: Sequence[bytes], threads: int = 1, memo: bool = False,
                 table: Optional[PatternTable] = None):
    """Timing helper: independent requests over host threads. Returns (events, lines, checksum)."""
    t = table or PatternTable(patterns)
    c = make_cfg(cfg)
    bufs = [np.frombuffer(x, dtype=np.uint8) if not isinstance(x, np.ndarray) else x for x in logs]
    ptrs = (C.c_void_p * len(bufs))(*[b.ctypes.data for b in bufs])
    lens = (C.c_size_t * len(bufs))(*[len(b) for b in bufs])
    total = C.c_uint64(0)
    cks = C.c_double(0)
    err = C.create_string_buffer(1024)
    ev = lib().orc_analyze_many(C.byref(c), t.arr, t.n, ptrs, lens, len(bufs), threads,
                                OPT_MEMO if memo else 0, C.byref(total), C.byref(cks), err, 1024)
    if ev < 0:
        raise RuntimeError(err.value.decode("utf-8", "replace"))
    return ev, total.value, cks.value
