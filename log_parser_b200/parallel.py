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
