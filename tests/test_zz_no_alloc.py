"""lp_split_lines_noalloc — the entry point for callers that may not allocate GPU memory (a host
with a fixed device-memory plan, a plugin inside someone else's allocator): its scratch is a fixed
array inside the loaded module instead of a cudaMalloc'ed pool.

Checked here through the C ABI against the existing entry point that computes the same thing,
lp_analyze with LP_OPT_LINE_OFFSETS, and directly against the oracle's String.split restatement: line
starts, line ends and the Java-trimmed line count must be identical (integers: bit-exact, like the
other parity tests). Two calls in a row on the same device,
the second on a different, shorter log, prove that what the first call left in the shared scratch
does not leak into the next one. All device buffers are allocated by the test, outside the call.
This file sorts last on purpose: a run with -x has already executed every other test.
"""
import numpy as np
import pytest

from log_parser_b200 import binding, synth
from log_parser_b200.model import PatternSpec, ScoringConfig

pytestmark = pytest.mark.gpu


def _reference_offsets(lp, log: bytes):
    r = lp.analyze(log, options=binding.OPT_LINE_OFFSETS)
    total = r.line_end.shape[0]
    return r.line_start[:total].astype(np.uint32), r.line_end.astype(np.uint32), int(r.log_n_lines[0]), total


def _noalloc_offsets(log: bytes, capacity: int):
    import torch

    dev = torch.device("cuda", 0)
    src = torch.from_numpy(np.frombuffer(log, np.uint8).copy()).to(dev) if len(log) else torch.zeros(16, dtype=torch.uint8, device=dev)
    ls = torch.full((capacity,), -1, dtype=torch.int32, device=dev)     # poisoned: stale values would show
    le = torch.full((capacity,), -1, dtype=torch.int32, device=dev)
    counts = torch.zeros(3, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    binding.split_lines_noalloc(0, src.data_ptr(), len(log), ls.data_ptr(), le.data_ptr(), capacity, counts.data_ptr())
    c = counts.cpu().numpy()
    total = int(c[1])
    return ls.cpu().numpy().view(np.uint32)[:total], le.cpu().numpy().view(np.uint32)[:total], int(c[0]), total, int(c[2])


def test_noalloc_split_matches_analyze_twice_in_a_row():
    wl, big = synth.config_c1()
    small = b"first\r\nsecond\n\nlast line without newline"
    lp = binding.LogParser(ScoringConfig())
    lp.load_patterns([PatternSpec(id="p", primary_regex="never-matches-anything-here", confidence=1.0)])
    from oracle import oracle as O

    blank_tail = b"x\r\n" + b"\n" * 70000 + b"\r\n\n"      # the trailing-empty rule has to look back over 70k blank lines
    for log in (big, small, b"", b"\n\n\n", b"a\n\r\n\n", blank_tail):      # consecutive calls share the in-module scratch
        exp_s, exp_e, exp_n, exp_total = _reference_offsets(lp, log)
        got_s, got_e, got_n, got_total, overflow = _noalloc_offsets(log, capacity=exp_total + 7)
        assert overflow == 0
        assert got_total == exp_total
        assert got_n == exp_n
        np.testing.assert_array_equal(got_s, exp_s)
        np.testing.assert_array_equal(got_e, exp_e)
        ora = O.split(log)                              # the kept lines of String.split("\\r?\\n")
        assert got_n == len(ora)
        assert [(int(a), int(b)) for a, b in zip(got_s[:got_n], got_e[:got_n])] == [(int(a), int(b)) for a, b in ora]
    lp.close()


def test_noalloc_rejects_oversize_and_small_capacity():
    import torch

    dev = torch.device("cuda", 0)
    buf = torch.zeros(64, dtype=torch.uint8, device=dev)
    out = torch.zeros(8, dtype=torch.int32, device=dev)
    counts = torch.zeros(3, dtype=torch.int64, device=dev)
    with pytest.raises(binding.LpError) as ei:
        binding.split_lines_noalloc(0, buf.data_ptr(), binding.NOALLOC_MAX_BYTES + 1, out.data_ptr(), out.data_ptr(), 8, counts.data_ptr())
    assert ei.value.code == binding.LP_ERR_LIMIT and "limit" in str(ei.value)
    log = b"a\nb\nc\nd\n"
    src = torch.from_numpy(np.frombuffer(log, np.uint8).copy()).to(dev)
    with pytest.raises(binding.LpError) as ei:
        binding.split_lines_noalloc(0, src.data_ptr(), len(log), out.data_ptr(), out.data_ptr(), 2, counts.data_ptr())
    assert ei.value.code == binding.LP_ERR_LIMIT
    assert int(counts.cpu()[1]) == 5          # required capacity is reported
