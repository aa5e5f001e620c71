"""ctypes binding of liblogparser_b200.so — the same calls, in the same order, that the Java 21
host makes through Panama FFM. There is no Python/CPU fallback: if the shared
library is missing or no CUDA device is usable, the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from .model import PatternSpec, ScoringConfig, severity_multiplier

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "liblogparser_b200.so")

LP_OK, LP_ERR_INVALID, LP_ERR_SYNTAX, LP_ERR_UNSUPPORTED, LP_ERR_CUDA, LP_ERR_LIMIT, LP_ERR_BACKTRACK = 0, -1, -2, -3, -4, -5, -6
OPT_FACTORS, OPT_LINE_OFFSETS, OPT_NO_RANK, OPT_SPANS, OPT_STRICT_BUDGET = 1, 2, 4, 8, 16


class LpError(RuntimeError):
    def __init__(self, code: int, msg: str, bad_pattern: int = -1):
        super().__init__(f"[{code}] {msg}")
        self.code = code
        self.bad_pattern = bad_pattern


class PatternSyntaxError(LpError):
    """java.util.regex.PatternSyntaxException (AnalysisService.java:64,70,80 -> HTTP 500)."""


class UnsupportedRegexError(LpError):
    pass


class ScoringCfg(C.Structure):
    _fields_ = [
        ("decay_constant", C.c_double), ("max_window", C.c_int32), ("time_window_hours", C.c_int32),
        ("early_bonus_threshold", C.c_double), ("max_early_bonus", C.c_double), ("penalty_threshold", C.c_double),
        ("max_context_factor", C.c_double), ("frequency_threshold", C.c_double), ("max_penalty", C.c_double),
        ("decay_lut", C.POINTER(C.c_double)),
    ]


class SecondaryDesc(C.Structure):
    _fields_ = [("regex", C.c_char_p), ("weight", C.c_double), ("proximity_window", C.c_int32)]


class SequenceDesc(C.Structure):
    _fields_ = [("event_regex", C.POINTER(C.c_char_p)), ("n_events", C.c_uint32), ("bonus_multiplier", C.c_double)]


class PatternDesc(C.Structure):
    _fields_ = [
        ("id", C.c_char_p), ("primary_regex", C.c_char_p), ("confidence", C.c_double),
        ("severity_multiplier", C.c_double), ("has_context", C.c_int32), ("lines_before", C.c_int32),
        ("lines_after", C.c_int32), ("n_secondaries", C.c_uint32), ("secondaries", C.POINTER(SecondaryDesc)),
        ("n_sequences", C.c_uint32), ("sequences", C.POINTER(SequenceDesc)),
    ]


class ResultView(C.Structure):
    _fields_ = [
        ("n_logs", C.c_uint64), ("n_events", C.c_uint64), ("log_n_lines", C.POINTER(C.c_uint64)),
        ("log_event_begin", C.POINTER(C.c_uint64)), ("line", C.POINTER(C.c_uint32)), ("pattern", C.POINTER(C.c_uint32)),
        ("score", C.POINTER(C.c_double)), ("rank", C.POINTER(C.c_uint32)), ("factors", C.POINTER(C.c_double)),
        ("freq_delta", C.POINTER(C.c_uint64)), ("n_patterns", C.c_uint64), ("line_start", C.POINTER(C.c_uint32)),
        ("line_end", C.POINTER(C.c_uint32)), ("total_lines", C.c_uint64),
        ("span_start", C.POINTER(C.c_uint32)), ("span_end", C.POINTER(C.c_uint32)), ("budget_exceeded", C.c_uint64),
    ]


class Timings(C.Structure):
    _fields_ = [
        ("total_ms", C.c_float), ("scan_ms", C.c_float), ("verify_ms", C.c_float), ("score_ms", C.c_float),
        ("h2d_ms", C.c_float), ("kernel_launches", C.c_uint64), ("scan_bytes", C.c_uint64),
        ("candidates", C.c_uint64), ("verify_tasks", C.c_uint64), ("dfa_scan_ms", C.c_float), ("lines_ms", C.c_float),
        ("dfa_scan_bytes", C.c_uint64),
    ]


# every symbol include/logparser_b200.h declares
EXPORTS = [
    "lp_create", "lp_destroy", "lp_last_error", "lp_version", "lp_load_patterns", "lp_analyze", "lp_analyze_device",
    "lp_analyze_batch", "lp_analyze_batch_device", "lp_batch_scan_device", "lp_batch_score", "lp_result_get",
    "lp_result_free", "lp_get_timings", "lp_stream", "lp_split_lines_noalloc",
    "lp_counts_by_id", "lp_batch_scan_device_async", "lp_batch_score_gathered", "lp_comm_unique_id", "lp_comm_init",
    "lp_analyze_batch_dist_device", "lp_analyze_batch_dist", "lp_host_alloc", "lp_host_free", "lp_host_register",
    "lp_host_unregister", "lp_range_halo_lines", "lp_range_summary_words", "lp_range_scan_device", "lp_range_score", "lp_analyze_range_dist_device", "lp_analyze_range_dist", "lp_create_multi",
]

_lib = None


def load_library() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(
            f"{SO_PATH} is missing: build it with `python -m log_parser_b200.build` "
            "(the B200 path has no CPU fallback)")
    L = C.CDLL(SO_PATH)
    vp, u8p, u64p = C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)
    L.lp_create.restype = vp
    L.lp_create.argtypes = [C.POINTER(ScoringCfg), C.c_int]
    L.lp_create_multi.restype = vp
    L.lp_create_multi.argtypes = [C.POINTER(ScoringCfg), C.POINTER(C.c_int), C.c_size_t]
    L.lp_destroy.argtypes = [vp]
    L.lp_last_error.restype = C.c_char_p
    L.lp_version.restype = C.c_char_p
    L.lp_load_patterns.argtypes = [vp, C.POINTER(PatternDesc), C.c_size_t, C.POINTER(C.c_int64)]
    L.lp_analyze.argtypes = [vp, u8p, C.c_size_t, u64p, C.c_int, C.POINTER(vp)]
    L.lp_analyze_device.argtypes = [vp, u8p, C.c_size_t, u64p, C.c_int, C.POINTER(vp)]
    L.lp_analyze_batch.argtypes = [vp, u8p, u64p, C.c_size_t, u64p, C.c_int, C.POINTER(vp)]
    L.lp_analyze_batch_device.argtypes = [vp, u8p, u64p, C.c_size_t, u64p, C.c_int, C.POINTER(vp)]
    L.lp_batch_scan_device.argtypes = [vp, u8p, u64p, C.c_size_t, C.c_int, u64p]
    L.lp_batch_score.argtypes = [vp, u64p, C.POINTER(vp)]
    L.lp_result_get.argtypes = [vp, C.POINTER(ResultView)]
    L.lp_result_free.argtypes = [vp]
    L.lp_get_timings.argtypes = [vp, C.POINTER(Timings)]
    L.lp_stream.restype = vp
    L.lp_stream.argtypes = [vp]
    L.lp_counts_by_id.argtypes = [vp, u64p, u64p]
    L.lp_batch_scan_device_async.argtypes = [vp, u8p, u64p, C.c_size_t, C.c_int, vp]
    L.lp_batch_score_gathered.argtypes = [vp, vp, C.c_uint32, C.c_uint32, u64p, C.POINTER(vp)]
    L.lp_comm_unique_id.argtypes = [vp]
    L.lp_comm_init.argtypes = [vp, vp, C.c_uint32, C.c_uint32]
    L.lp_analyze_batch_dist_device.argtypes = [vp, u8p, u64p, C.c_size_t, u64p, C.c_int, C.POINTER(vp)]
    L.lp_analyze_batch_dist.argtypes = [vp, u8p, u64p, C.c_size_t, u64p, C.c_int, C.POINTER(vp)]
    L.lp_host_alloc.restype = vp
    L.lp_host_alloc.argtypes = [C.c_size_t]
    L.lp_host_free.argtypes = [vp]
    L.lp_host_register.argtypes = [vp, C.c_size_t]
    L.lp_host_unregister.argtypes = [vp]
    L.lp_range_halo_lines.restype = C.c_uint32
    L.lp_range_halo_lines.argtypes = [vp]
    L.lp_range_summary_words.restype = C.c_size_t
    L.lp_range_summary_words.argtypes = [vp]
    L.lp_range_scan_device.argtypes = [vp, u8p, C.c_size_t, C.c_uint32, C.c_uint32, C.c_int, u64p]
    L.lp_range_score.argtypes = [vp, u64p, C.c_uint32, C.c_uint32, u64p, C.POINTER(vp)]
    L.lp_analyze_range_dist_device.argtypes = [vp, u8p, C.c_size_t, C.c_uint32, C.c_uint32, u64p, C.c_int, C.POINTER(vp)]
    L.lp_analyze_range_dist.argtypes = [vp, u8p, C.c_size_t, C.c_uint32, C.c_uint32, u64p, C.c_int, C.POINTER(vp)]
    L.lp_split_lines_noalloc.argtypes = [C.c_int, vp, C.c_size_t, vp, vp, C.c_size_t, vp, vp, C.c_char_p, C.c_size_t]
    _lib = L
    return L


def _last_error() -> str:
    return load_library().lp_last_error().decode("utf-8", "replace")


def _raise(code: int, bad: int = -1):
    msg = _last_error()
    if code == LP_ERR_SYNTAX:
        raise PatternSyntaxError(code, msg, bad)
    if code == LP_ERR_UNSUPPORTED:
        raise UnsupportedRegexError(code, msg, bad)
    raise LpError(code, msg, bad)


class PatternTable:
    """lp_pattern_desc[] built from PatternSpec objects in (set, pattern) load order."""

    def __init__(self, patterns: Sequence[PatternSpec]):
        self.n = len(patterns)
        self._keep = []
        self.arr = (PatternDesc * max(1, self.n))()
        for i, p in enumerate(patterns):
            secs = (SecondaryDesc * max(1, len(p.secondaries)))()
            for j, s in enumerate(p.secondaries):
                secs[j] = SecondaryDesc(s.regex.encode("utf-8"), s.weight, s.proximity_window)
            seqs = (SequenceDesc * max(1, len(p.sequences)))()
            for j, s in enumerate(p.sequences):
                ev = (C.c_char_p * max(1, len(s.events)))(*[e.encode("utf-8") for e in s.events])
                self._keep.append(ev)
                seqs[j] = SequenceDesc(C.cast(ev, C.POINTER(C.c_char_p)), len(s.events), s.bonus_multiplier)
            self._keep += [secs, seqs]
            self.arr[i] = PatternDesc(
                None if p.id is None else p.id.encode("utf-8"), p.primary_regex.encode("utf-8"), p.confidence,
                severity_multiplier(p.severity), 1 if p.has_context else 0,
                p.lines_before if p.has_context else 0, p.lines_after if p.has_context else 0,
                len(p.secondaries), C.cast(secs, C.POINTER(SecondaryDesc)),
                len(p.sequences), C.cast(seqs, C.POINTER(SequenceDesc)))


@dataclass
class Result:
    n_logs: int
    log_n_lines: np.ndarray       # uint64 [n_logs]
    log_event_begin: np.ndarray   # uint64 [n_logs+1]
    line: np.ndarray              # uint32, 1-based within the log
    pattern: np.ndarray           # uint32 flat pattern index
    score: np.ndarray             # float64
    rank: np.ndarray              # uint32
    freq_delta: np.ndarray        # uint64 [n_patterns]
    factors: Optional[np.ndarray] = None
    line_start: Optional[np.ndarray] = None
    line_end: Optional[np.ndarray] = None
    span_start: Optional[np.ndarray] = None     # OPT_SPANS: Matcher.start()/end() as byte offsets within the line
    span_end: Optional[np.ndarray] = None
    budget_exceeded: int = 0                    # (line, regex) evaluations abandoned by the backtracking verifier (count as no match)

    @property
    def n_lines(self) -> int:
        return int(self.log_n_lines[0])

    @property
    def n_events(self) -> int:
        return int(self.line.shape[0])


def _arr(ptr, n, dtype):
    if n == 0 or not ptr:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, (n,)).astype(dtype, copy=True)


class _ResultOwner:
    """Keeps an lp_result alive while numpy views of its columns exist; lp_result_free when the last one goes."""

    def __init__(self, L, handle):
        self.L, self.h = L, handle

    def __del__(self):
        try:
            if self.h:
                self.L.lp_result_free(self.h)
                self.h = None
        except Exception:
            pass


_CT = {np.dtype(np.uint32): C.c_uint32, np.dtype(np.uint64): C.c_uint64, np.dtype(np.float64): C.c_double}


def _view(ptr, n, dtype, owner):
    """Zero-copy column view (the memory is the page-locked buffer the device wrote into)."""
    n = int(n)
    if n == 0 or not ptr:
        return np.zeros(0, dtype)
    buf = (_CT[np.dtype(dtype)] * n).from_address(C.cast(ptr, C.c_void_p).value)
    buf._owner = owner          # the buffer object is the array's base: it pins the result
    return np.frombuffer(buf, dtype=dtype)


class LogParser:
    """One lp_ctx: scoring.* configuration + loaded pattern set bound to one CUDA device."""

    def __init__(self, cfg: Optional[ScoringConfig] = None, device: int = 0, decay_lut: Optional[np.ndarray] = None,
                 devices: Optional[Sequence[int]] = None):
        """device = -1: every visible device; devices = [...]: lp_create_multi (one process driving several GPUs:
        analyze_batch cuts a batch into one block of logs per device)."""
        self.L = load_library()
        cfg = cfg or ScoringConfig()
        self.cfg = cfg
        lut_p = None
        if decay_lut is not None:
            self._lut = np.ascontiguousarray(decay_lut, dtype=np.float64)
            assert self._lut.shape[0] >= cfg.max_window + 1
            lut_p = self._lut.ctypes.data_as(C.POINTER(C.c_double))
        sc = ScoringCfg(cfg.decay_constant, cfg.max_window, cfg.time_window_hours, cfg.early_bonus_threshold,
                        cfg.max_early_bonus, cfg.penalty_threshold, cfg.max_context_factor, cfg.frequency_threshold,
                        cfg.max_penalty, lut_p)
        if devices is not None:
            arr = (C.c_int * len(devices))(*[int(d) for d in devices])
            self.h = self.L.lp_create_multi(C.byref(sc), arr, len(devices))
        else:
            self.h = self.L.lp_create(C.byref(sc), device)
        if not self.h:
            raise LpError(LP_ERR_CUDA, _last_error())
        self.n_patterns = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.lp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_patterns(self, patterns: Sequence[PatternSpec]):
        t = PatternTable(patterns)
        bad = C.c_int64(-1)
        rc = self.L.lp_load_patterns(self.h, t.arr, t.n, C.byref(bad))
        if rc != LP_OK:
            _raise(rc, bad.value)
        self.n_patterns = t.n

    # ---- helpers
    def _prior(self, prior_count):
        if prior_count is None:
            return None, None
        a = np.ascontiguousarray(prior_count, dtype=np.uint64)
        assert a.shape[0] == self.n_patterns
        return a, a.ctypes.data_as(C.POINTER(C.c_uint64))

    def _take(self, out) -> Result:
        v = ResultView()
        rc = self.L.lp_result_get(out, C.byref(v))
        if rc != LP_OK:
            _raise(rc)
        m, nl = v.n_events, v.n_logs
        owner = _ResultOwner(self.L, out)       # columns are views into the result; it is freed with the last of them
        res = Result(
            n_logs=nl, log_n_lines=_arr(v.log_n_lines, nl, np.uint64), log_event_begin=_arr(v.log_event_begin, nl + 1, np.uint64),
            line=_view(v.line, m, np.uint32, owner), pattern=_view(v.pattern, m, np.uint32, owner), score=_view(v.score, m, np.float64, owner),
            rank=_view(v.rank, m, np.uint32, owner), freq_delta=_view(v.freq_delta, v.n_patterns, np.uint64, owner))
        res.budget_exceeded = int(v.budget_exceeded)
        if v.factors:
            res.factors = _view(v.factors, m * 7, np.float64, owner).reshape(m, 7)
        if v.span_start:
            res.span_start = _view(v.span_start, m, np.uint32, owner)
            res.span_end = _view(v.span_end, m, np.uint32, owner)
        if v.line_start:
            res.line_start = _arr(v.line_start, v.total_lines + 1, np.uint32)
            res.line_end = _arr(v.line_end, v.total_lines, np.uint32)
        return res

    # ---- the hot path
    def analyze(self, log, prior_count=None, options: int = 0) -> Result:
        """lp_analyze: host buffer in, events out (replaces AnalysisService.java:53-113)."""
        buf = np.frombuffer(log, dtype=np.uint8) if not isinstance(log, np.ndarray) else log
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_analyze(self.h, buf.ctypes.data if buf.size else None, buf.size, pp, options, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    def analyze_host_ptr(self, ptr: int, nbytes: int, prior_count=None, options: int = 0) -> Result:
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_analyze(self.h, ptr, nbytes, pp, options, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    def analyze_device(self, dev_ptr: int, nbytes: int, prior_count=None, options: int = 0) -> Result:
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_analyze_device(self.h, dev_ptr, nbytes, pp, options, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    @staticmethod
    def _offsets(offsets):
        a = np.ascontiguousarray(offsets, dtype=np.uint64)
        return a, a.ctypes.data_as(C.POINTER(C.c_uint64))

    def analyze_batch(self, logs: Sequence[bytes], prior_count=None, options: int = 0) -> Result:
        blob = b"".join(logs)
        off = np.zeros(len(logs) + 1, np.uint64)
        np.cumsum([len(x) for x in logs], out=off[1:])
        buf = np.frombuffer(blob, dtype=np.uint8)
        oa, op = self._offsets(off)
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_analyze_batch(self.h, buf.ctypes.data if buf.size else None, op, len(logs), pp, options, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    def analyze_batch_host_ptr(self, host_ptr: int, offsets, prior_count=None, options: int = 0) -> Result:
        """lp_analyze_batch on a caller-owned host buffer (page-locked for a truly asynchronous copy) that holds the concatenated logs."""
        oa, op = self._offsets(offsets)
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_analyze_batch(self.h, host_ptr, op, len(oa) - 1, pp, options, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    def analyze_batch_device(self, dev_ptr: int, offsets, prior_count=None, options: int = 0) -> Result:
        oa, op = self._offsets(offsets)
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_analyze_batch_device(self.h, dev_ptr, op, len(oa) - 1, pp, options, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    def batch_scan_device(self, dev_ptr: int, offsets, options: int = 0) -> np.ndarray:
        oa, op = self._offsets(offsets)
        counts = np.zeros(max(1, self.n_patterns), np.uint64)
        rc = self.L.lp_batch_scan_device(self.h, dev_ptr, op, len(oa) - 1, options, counts.ctypes.data_as(C.POINTER(C.c_uint64)))
        if rc != LP_OK:
            _raise(rc)
        return counts[: self.n_patterns]

    def batch_score(self, prior_count=None) -> Result:
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_batch_score(self.h, pp, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    def counts_by_id(self, counts) -> np.ndarray:
        """Sums per-pattern counts over the patterns that share an id (the key of FrequencyTrackingService)."""
        a = np.ascontiguousarray(counts, dtype=np.uint64)
        out = np.zeros(max(1, self.n_patterns), np.uint64)
        rc = self.L.lp_counts_by_id(self.h, a.ctypes.data_as(C.POINTER(C.c_uint64)), out.ctypes.data_as(C.POINTER(C.c_uint64)))
        if rc != LP_OK:
            _raise(rc)
        return out[: self.n_patterns]

    def batch_scan_device_async(self, dev_ptr: int, offsets, counts_dev_ptr: int, options: int = 0) -> None:
        oa, op = self._offsets(offsets)
        rc = self.L.lp_batch_scan_device_async(self.h, dev_ptr, op, len(oa) - 1, options, counts_dev_ptr)
        if rc != LP_OK:
            _raise(rc)

    def batch_score_gathered(self, all_counts_dev_ptr: int, rank: int, world: int, prior_count=None) -> Optional[Result]:
        """None = LP_RETRY: every rank repeats scan + exchange + score (some rank's buffers were regrown)."""
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_batch_score_gathered(self.h, all_counts_dev_ptr, rank, world, pp, C.byref(out))
        if rc == 1:
            return None
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    # ---- NCCL inside the library
    def comm_unique_id(self) -> bytes:
        buf = C.create_string_buffer(128)
        rc = self.L.lp_comm_unique_id(buf)
        if rc != LP_OK:
            _raise(rc)
        return buf.raw

    def comm_init(self, unique_id: bytes, rank: int, world: int) -> None:
        assert len(unique_id) == 128
        rc = self.L.lp_comm_init(self.h, C.create_string_buffer(unique_id, 128), rank, world)
        if rc != LP_OK:
            _raise(rc)

    def analyze_batch_dist_device(self, dev_ptr: int, offsets, prior_count=None, options: int = 0) -> Result:
        """Collective: this rank's block of a batch sharded by log (scan -> ncclAllGather -> score in one call)."""
        oa, op = self._offsets(offsets)
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_analyze_batch_dist_device(self.h, dev_ptr, op, len(oa) - 1, pp, options, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    def analyze_batch_dist_host_ptr(self, host_ptr: int, offsets, prior_count=None, options: int = 0) -> Result:
        """The same collective step with this rank's logs in (page-locked) host memory: the copy overlaps the scan."""
        oa, op = self._offsets(offsets)
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_analyze_batch_dist(self.h, host_ptr, op, len(oa) - 1, pp, options, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    # ---- one log sharded by line range (lp_range_*)
    def range_halo_lines(self) -> int:
        return int(self.L.lp_range_halo_lines(self.h))

    def range_summary_words(self) -> int:
        return int(self.L.lp_range_summary_words(self.h))

    def range_scan_device(self, dev_ptr: int, length: int, halo_before: int, halo_after: int, options: int = 0) -> np.ndarray:
        summary = np.zeros(max(1, self.range_summary_words()), np.uint64)
        rc = self.L.lp_range_scan_device(self.h, dev_ptr, length, halo_before, halo_after, options,
                                         summary.ctypes.data_as(C.POINTER(C.c_uint64)))
        if rc != LP_OK:
            _raise(rc)
        return summary

    def range_score(self, summaries: np.ndarray, rank: int, world: int, prior_count=None) -> Result:
        sa = np.ascontiguousarray(summaries, dtype=np.uint64).reshape(world, -1)
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_range_score(self.h, sa.ctypes.data_as(C.POINTER(C.c_uint64)), rank, world, pp, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    def analyze_range_dist_device(self, dev_ptr: int, length: int, halo_before: int, halo_after: int, prior_count=None,
                                  options: int = 0) -> Result:
        """Collective: this rank's piece of ONE log cut by line range; summaries exchanged and composed on the device."""
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_analyze_range_dist_device(self.h, dev_ptr, length, halo_before, halo_after, pp, options, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    def analyze_range_dist_host_ptr(self, host_ptr: int, length: int, halo_before: int, halo_after: int, prior_count=None,
                                    options: int = 0) -> Result:
        keep, pp = self._prior(prior_count)
        out = C.c_void_p()
        rc = self.L.lp_analyze_range_dist(self.h, host_ptr, length, halo_before, halo_after, pp, options, C.byref(out))
        if rc != LP_OK:
            _raise(rc)
        return self._take(out)

    def stream_ptr(self) -> int:
        return int(self.L.lp_stream(self.h) or 0)

    def timings(self) -> dict:
        t = Timings()
        self.L.lp_get_timings(self.h, C.byref(t))
        return {f[0]: getattr(t, f[0]) for f in Timings._fields_}


NOALLOC_MAX_BYTES = 1 << 30


def split_lines_noalloc(device: int, d_log: int, nbytes: int, d_line_start: int, d_line_end: int, capacity: int,
                        d_counts: int, stream: int = 0) -> None:
    """lp_split_lines_noalloc on caller-owned device buffers (raw pointers). For callers that must
    not trigger any GPU allocation: the library's scratch for this call is a fixed in-module array,
    so inputs above NOALLOC_MAX_BYTES are refused, and concurrent calls on one device must be
    serialised by the caller."""
    L = load_library()
    err = C.create_string_buffer(512)
    rc = L.lp_split_lines_noalloc(device, d_log, nbytes, d_line_start, d_line_end, capacity, d_counts, stream or None, err, 512)
    if rc != LP_OK:
        raise LpError(rc, err.value.decode("utf-8", "replace"))


_dbg = None


def load_debug_library() -> C.CDLL:
    """liblogparser_b200_debug.so: host-compiler introspection for the CPU test tier (not part of the product library)."""
    global _dbg
    if _dbg is not None:
        return _dbg
    from . import build as _build

    L = C.CDLL(_build.build_debug())
    vp = C.c_void_p
    L.lp_debug_rx_compile.restype = vp
    L.lp_debug_rx_compile.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_int)]
    L.lp_debug_rx_free.argtypes = [vp]
    L.lp_debug_rx_is_dfa.argtypes = [vp]
    L.lp_debug_rx_states.restype = C.c_uint32
    L.lp_debug_rx_states.argtypes = [vp]
    L.lp_debug_rx_match.argtypes = [vp, C.c_char_p, C.c_size_t]
    L.lp_debug_rx_literals.restype = C.c_size_t
    L.lp_debug_rx_literals.argtypes = [vp, C.c_char_p, C.c_size_t]
    L.lp_debug_rx_anchor.argtypes = [vp]
    L.lp_debug_rx_match_local.argtypes = [vp, C.c_char_p, C.c_size_t]
    L.lp_debug_rx_span.argtypes = [vp, C.c_char_p, C.c_size_t, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.lp_debug_rx_match_anchored.argtypes = [vp, C.c_char_p, C.c_size_t]
    L.lp_debug_plan_dump.restype = C.c_size_t
    L.lp_debug_plan_dump.argtypes = [vp, C.c_size_t, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    L.lp_debug_group_match.restype = C.c_uint32
    L.lp_debug_group_match.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.c_size_t, C.c_char_p, C.c_size_t]
    L.lp_debug_group_skip.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.c_size_t, C.POINTER(C.c_uint32)]
    _dbg = L
    return L


def host_alloc(nbytes: int) -> int:
    """Page-locked host memory from the library (lp_host_alloc); the address, 0 on failure."""
    return int(load_library().lp_host_alloc(nbytes) or 0)


def host_free(ptr: int) -> None:
    load_library().lp_host_free(ptr)


class DebugRegex:
    """Host-compiler introspection for the CPU-only test tier (never part of analyze)."""

    def __init__(self, pattern: str, flags: int = 0, force_vm: bool = False, min_literal: int = 0):
        self.L = load_debug_library()
        err = C.create_string_buffer(512)
        kind = C.c_int(0)
        self.h = self.L.lp_debug_rx_compile(pattern.encode("utf-8"), flags, (1 if force_vm else 0) | (min_literal << 8), err, 512, C.byref(kind))
        if not self.h:
            msg = err.value.decode("utf-8", "replace")
            if kind.value == 1:
                raise PatternSyntaxError(LP_ERR_SYNTAX, msg)
            if kind.value == 2:
                raise UnsupportedRegexError(LP_ERR_UNSUPPORTED, msg)
            raise LpError(LP_ERR_INVALID, msg)

    @property
    def is_dfa(self) -> bool:
        return bool(self.L.lp_debug_rx_is_dfa(self.h))

    @property
    def size(self) -> int:
        return self.L.lp_debug_rx_states(self.h)

    def match(self, text) -> int:
        b = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        return self.L.lp_debug_rx_match(self.h, b, len(b))

    @property
    def anchor(self) -> int:
        """0: verified by walking the whole line; 1 / 2: decided at the literal occurrences by the
        forward / backward anchored automaton."""
        return int(self.L.lp_debug_rx_anchor(self.h))

    def match_local(self, text) -> int:
        """1/0 as the boundary test at the literal occurrences decides; -2 if the regex is not of that form."""
        b = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        return int(self.L.lp_debug_rx_match_local(self.h, b, len(b)))

    def match_anchored(self, text) -> int:
        b = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        return int(self.L.lp_debug_rx_match_anchored(self.h, b, len(b)))

    def span(self, text):
        """(start, end) byte offsets of the first find(), or None."""
        b = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        s, e = C.c_uint32(0), C.c_uint32(0)
        r = self.L.lp_debug_rx_span(self.h, b, len(b), C.byref(s), C.byref(e))
        if r < 0:
            raise LpError(LP_ERR_BACKTRACK, "backtracking budget exceeded")
        return (s.value, e.value) if r else None

    def literals(self):
        n = self.L.lp_debug_rx_literals(self.h, None, 0)
        buf = C.create_string_buffer(n + 1)
        self.L.lp_debug_rx_literals(self.h, buf, n + 1)
        out = []
        for ln in buf.raw[:n].split(b"\n"):
            if ln:
                out.append((ln[:1] == b"i", ln[2:]))
        return out

    def __del__(self):
        try:
            if self.h:
                self.L.lp_debug_rx_free(self.h)
        except Exception:
            pass
