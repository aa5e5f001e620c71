"""Multi-GPU host logic: one process per GPU, a batch of pod logs sharded by log.

The reference handles requests one after the other on a single thread
(AnalysisService.java:89-113); the only state that couples them is the per-pattern frequency count
(FrequencyTrackingService.java:41-93, read before increment at ScoringService.java:84-88). A batch
is therefore split into contiguous blocks of logs, one block per rank; every rank scans its block,
the per-pattern event counts are exchanged with ONE all-gather (P x int64, 4 KB at P = 500 —
latency-bound, NCCL over NVLink on GPUs, gloo in the CPU tests), each rank adds the counts of the
lower ranks to the window's prior counts, and scores its block. Concatenating the ranks' results in
rank order reproduces the reference processing the whole batch sequentially.

The compute engine is anything with `scan(logs) -> counts[P]` and `score(prior[P]) -> result`;
on GPUs it is `GpuBatchEngine` (lp_batch_scan_device / lp_batch_score).
"""
from __future__ import annotations

from typing import List, Optional, Protocol, Sequence, Tuple

import numpy as np


def shard_bounds(sizes: Sequence[int], world_size: int) -> List[Tuple[int, int]]:
    """Contiguous blocks of log indices per rank, balanced by bytes; every log belongs to one rank."""
    n = len(sizes)
    total = float(sum(sizes)) or 1.0
    bounds, start, acc = [], 0, 0.0
    for r in range(world_size):
        target = total * (r + 1) / world_size
        end = start
        while end < n and (acc + sizes[end] <= target or end == start) and n - end > world_size - r - 1:
            acc += sizes[end]
            end += 1
        if r == world_size - 1:
            end = n
        bounds.append((start, end))
        start = end
    return bounds


class Engine(Protocol):
    n_patterns: int

    def scan(self, logs: Sequence[bytes]) -> np.ndarray: ...   # per-pattern event counts of these logs
    def score(self, prior: np.ndarray): ...                     # result of the last scan, given the counts seen before it


def exclusive_prefix(counts: np.ndarray, rank: int, world_size: int, group=None) -> Tuple[np.ndarray, np.ndarray]:
    """All-gathers the per-rank counts; returns (sum over lower ranks, sum over all ranks)."""
    import torch
    import torch.distributed as dist

    if world_size == 1:
        return np.zeros_like(counts), counts.copy()
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    mine = torch.from_numpy(counts.astype(np.int64)).to(dev)
    gathered = torch.empty(world_size * mine.numel(), dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(gathered, mine, group=group)
    g = gathered.view(world_size, -1).cpu().numpy()
    return g[:rank].sum(axis=0).astype(np.uint64), g.sum(axis=0).astype(np.uint64)


def analyze_batch_sharded(engine: Engine, logs: Sequence[bytes], rank: int, world_size: int,
                          prior_count: Optional[np.ndarray] = None, group=None):
    """Runs this rank's share of `logs`. Returns (result, (first_log, end_log), total_counts)."""
    sizes = [len(x) for x in logs]
    lo, hi = shard_bounds(sizes, world_size)[rank]
    n_pat = engine.n_patterns
    counts = np.asarray(engine.scan(logs[lo:hi]), dtype=np.uint64) if hi > lo else np.zeros(n_pat, np.uint64)
    before, total = exclusive_prefix(counts, rank, world_size, group)      # every rank takes part, even with no logs
    if hasattr(engine, "by_id"):                                            # counts are kept per pattern id (FTS:46-50)
        before = np.asarray(engine.by_id(before), dtype=np.uint64)
    prior = before if prior_count is None else before + np.asarray(prior_count, dtype=np.uint64)
    return (engine.score(prior) if hi > lo else None), (lo, hi), total


class GpuBatchEngine:
    """lp_batch_scan_device / lp_batch_score on the current CUDA device."""

    def __init__(self, parser):
        self.lp = parser
        self.n_patterns = parser.n_patterns
        self._keep = None

    def scan(self, logs: Sequence[bytes]) -> np.ndarray:
        import torch

        blob = b"".join(logs)
        off = np.zeros(len(logs) + 1, np.uint64)
        np.cumsum([len(x) for x in logs], out=off[1:])
        host = np.frombuffer(blob, dtype=np.uint8) if blob else np.zeros(16, np.uint8)
        self._keep = torch.from_numpy(host.copy()).cuda()
        return self.lp.batch_scan_device(self._keep.data_ptr(), off)

    def score(self, prior: np.ndarray):
        return self.lp.batch_score(prior)

    def by_id(self, counts: np.ndarray) -> np.ndarray:
        return self.lp.counts_by_id(counts)



def init_library_comm(parser, rank: int, world_size: int, group=None) -> None:
    """Creates the library's own NCCL communicator (lp_comm_init): rank 0's ncclUniqueId travels to the
    other ranks through torch.distributed, after which lp_analyze_batch_dist_device needs no host-side
    collective at all."""
    import torch.distributed as dist

    box = [parser.comm_unique_id() if rank == 0 else None]
    if world_size > 1:
        dist.broadcast_object_list(box, src=0, group=group)
    parser.comm_init(box[0], rank, world_size)


class DeviceExchange:
    """The by-log exchange kept on the device (lp_batch_scan_device_async / lp_batch_score_gathered):
    this rank's [P counts, status word] are written by the scan phase straight into a device buffer,
    all-gathered with NCCL ordered on the library's stream, and turned into the exclusive prefix by a
    kernel of the score phase. No host round trip between the phases; one synchronisation per step.
    Call close() before the parser is closed (the buffers are tied to the parser's stream)."""

    def __init__(self, parser, rank: int, world_size: int, group=None):
        import torch

        self.lp, self.rank, self.world, self.group = parser, rank, world_size, group
        dev = torch.device("cuda", torch.cuda.current_device())
        n = parser.n_patterns + 1
        self.mine = torch.zeros(n, dtype=torch.int64, device=dev)
        self.all = torch.zeros(world_size * n, dtype=torch.int64, device=dev)
        self.stream = torch.cuda.ExternalStream(parser.stream_ptr(), device=dev)
        torch.cuda.synchronize()

    def step(self, dev_ptr: int, offsets, prior_count=None, options: int = 0, max_rounds: int = 16):
        import torch
        import torch.distributed as dist

        for _ in range(max_rounds):
            self.lp.batch_scan_device_async(dev_ptr, offsets, self.mine.data_ptr(), options)
            with torch.cuda.stream(self.stream):
                if self.world > 1:
                    dist.all_gather_into_tensor(self.all, self.mine, group=self.group)
                else:
                    self.all.copy_(self.mine)
            res = self.lp.batch_score_gathered(self.all.data_ptr(), self.rank, self.world, prior_count)
            if res is not None:
                return res                      # None: some rank regrew its buffers; every rank repeats the step
        raise RuntimeError("scratch estimates did not settle")

    def close(self):
        import torch

        torch.cuda.synchronize()
        self.mine = self.all = None
        torch.cuda.synchronize()


# ------------------------------------------------------------------------------------------------
# One log sharded by line range (second mode; lp_range_* in include/logparser_b200.h)
# ------------------------------------------------------------------------------------------------
def strip_trailing_separators(log: bytes) -> int:
    """Length of `log` without its trailing `\\r?\\n` separators — the trailing empty strings that
    String.split drops (AnalysisService.java:53). 0 for a non-empty log means it has no lines at all."""
    n = len(log)
    while n > 0 and log[n - 1] == 0x0A:
        n -= 1
        if n > 0 and log[n - 1] == 0x0D:
            n -= 1
    return n


def _line_starts_forward(log, pos: int, n_lines: int, end: int) -> Tuple[int, int]:
    """From line start `pos`, skips up to n_lines whole lines; positions are in the log extended by
    one virtual '\\n' at `end` (so `end + 1` is "after the last line"). Returns (position, lines skipped)."""
    k = 0
    while k < n_lines and pos <= end:
        j = log.find(b"\n", pos, end)
        pos = (j if j >= 0 else end) + 1
        k += 1
    return pos, k


def _line_starts_backward(log, pos: int, n_lines: int) -> Tuple[int, int]:
    k = 0
    while k < n_lines and pos > 0:
        j = log.rfind(b"\n", 0, pos - 1)
        pos = j + 1
        k += 1
    return pos, k


def split_line_ranges(log: bytes, world_size: int, halo_lines: int, length: Optional[int] = None):
    """Cuts log[:length] (no trailing separators) on line boundaries into world_size contiguous own
    ranges balanced by bytes, each widened by up to halo_lines whole lines on either side.
    Returns, per rank, (text_begin, text_end, halo_before_lines, halo_after_lines) where
    log[text_begin:text_end] is the text to hand to lp_range_scan_device: it ends with the content of
    its last line (the separator after it is not included)."""
    if halo_lines < 1:
        raise ValueError("halo_lines must be >= 1 (a shard without own lines is described by its halo)")
    end = strip_trailing_separators(log) if length is None else length
    cuts = [0]
    for r in range(1, world_size):
        target = max(cuts[-1], (end * r) // world_size)
        if target == 0:
            cuts.append(0)
            continue
        j = log.find(b"\n", target - 1, end)
        cuts.append((j if j >= 0 else end) + 1)       # end + 1: after the last line
    cuts.append(end + 1)
    out = []
    for r in range(world_size):
        own_b, own_e = cuts[r], cuts[r + 1]
        tb, before = _line_starts_backward(log, own_b, halo_lines)
        te, after = _line_starts_forward(log, own_e, halo_lines, end)
        # te is one past a (real or virtual) '\n'; drop that separator, and the '\r' of a "\r\n"
        stop = te - 1
        if stop < end and stop - 1 >= tb and log[stop - 1] == 0x0D:
            stop -= 1
        out.append((tb, max(tb, stop), before, after))
    return out


def merge_ranks(scores_per_rank: Sequence[np.ndarray]) -> List[np.ndarray]:
    """Whole-log ranking (score descending, discovery order on ties) from the shards' scores in rank order."""
    sizes = [len(s) for s in scores_per_rank]
    allv = np.concatenate([np.asarray(s, np.float64) for s in scores_per_rank]) if sizes else np.zeros(0)
    order = np.lexsort((np.arange(allv.size), -allv))
    rank = np.empty(allv.size, np.uint32)
    rank[order] = np.arange(allv.size, dtype=np.uint32)
    out, at = [], 0
    for n in sizes:
        out.append(rank[at:at + n])
        at += n
    return out


class GpuRangeEngine:
    """lp_range_scan_device / lp_range_score on the current CUDA device."""

    def __init__(self, parser):
        self.lp = parser
        self._keep = None

    def halo_lines(self) -> int:
        return self.lp.range_halo_lines()

    def scan(self, text: bytes, halo_before: int, halo_after: int) -> np.ndarray:
        import torch

        host = np.frombuffer(text, dtype=np.uint8) if len(text) else np.zeros(16, np.uint8)
        self._keep = torch.from_numpy(host.copy()).cuda()
        return self.lp.range_scan_device(self._keep.data_ptr(), len(text), halo_before, halo_after)

    def score(self, summaries: np.ndarray, rank: int, world_size: int, prior):
        return self.lp.range_score(summaries, rank, world_size, prior)


def gather_summaries(summary: np.ndarray, world_size: int, group=None) -> np.ndarray:
    """All-gathers the per-rank shard summaries -> [world_size, words] (NCCL on GPUs, gloo in CPU tests)."""
    import torch
    import torch.distributed as dist

    if world_size == 1:
        return summary.reshape(1, -1).copy()
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    mine = torch.from_numpy(summary.view(np.int64).copy()).to(dev)
    gathered = torch.empty(world_size * mine.numel(), dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(gathered, mine, group=group)
    return gathered.cpu().numpy().view(np.uint64).reshape(world_size, -1)


def analyze_range_sharded(engine, log: bytes, rank: int, world_size: int, prior_count=None, group=None):
    """This rank's share of ONE log cut by line range. Every rank holds (or can read) the log bytes but
    uploads only its own range plus halos. Returns the shard's result (global line numbers), or None
    when the log has no lines at all."""
    end = strip_trailing_separators(log)
    if end == 0 and len(log) > 0:
        return None                                   # only separators: String.split gives [] (AS:53)
    tb, te, before, after = split_line_ranges(log, world_size, engine.halo_lines(), end)[rank]
    summary = engine.scan(bytes(log[tb:te]), before, after)
    return engine.score(gather_summaries(summary, world_size, group), rank, world_size, prior_count)


def analyze_range_dist(parser, log: bytes, rank: int, world_size: int, prior_count=None, options: int = 0):
    """The same share of ONE log with the exchange inside the library (lp_analyze_range_dist_device over the
    communicator of init_library_comm): the shard summaries are all-gathered over NVLink and composed on the
    device, nothing returns to the host between the scan and the events. Returns None when the log has no lines."""
    import torch

    end = strip_trailing_separators(log)
    if end == 0 and len(log) > 0:
        return None
    tb, te, before, after = split_line_ranges(log, world_size, max(1, parser.range_halo_lines()), end)[rank]
    text = torch.from_numpy(np.frombuffer(bytes(log[tb:te]) or b"\0", dtype=np.uint8).copy()).cuda()
    return parser.analyze_range_dist_device(text.data_ptr(), te - tb, before, after, prior_count, options)
