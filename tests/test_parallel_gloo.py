"""The N>1 host path on CPU: world_size-2 `gloo` processes shard a batch by log, exchange the
per-pattern counts with one all-gather and score with the exclusive prefix. The compute engine
here is the oracle (this is a test of the sharding / exchange logic, which is what differs from the
single-GPU path); on GPUs the same `analyze_batch_sharded` drives lp_batch_scan_device/lp_batch_score."""
import os
import socket

import numpy as np
import pytest

from log_parser_b200 import parallel, synth
from log_parser_b200.model import ScoringConfig


def test_shard_bounds_cover_everything():
    for sizes, w in [([5, 5, 5, 5], 2), ([100, 1, 1, 1], 2), ([1] * 7, 4), ([3], 2), ([], 2), ([9, 1, 9, 1, 9], 3), ([1, 1], 8)]:
        b = parallel.shard_bounds(sizes, w)
        assert len(b) == w and b[0][0] == 0 and b[-1][1] == len(sizes)
        assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
        assert all(lo <= hi for lo, hi in b)


class _OracleEngine:
    """scan = per-pattern event counts of a block of logs; score = the block analysed request after
    request with the frequency window starting at `prior` (what the reference would do)."""

    def __init__(self, patterns):
        from oracle import oracle as O

        self.O, self.patterns, self.n_patterns = O, patterns, len(patterns)
        self.cfg = ScoringConfig()
        self.logs = None

    def scan(self, logs):
        self.logs = list(logs)
        total = np.zeros(self.n_patterns, np.uint64)
        for lg in self.logs:
            total += self.O.analyze(self.cfg, self.patterns, lg).freq_delta
        return total

    def score(self, prior):
        out, prior = [], np.asarray(prior, np.uint64).copy()
        for lg in self.logs:
            r = self.O.analyze(self.cfg, self.patterns, lg, prior_count=prior)
            out.append(r)
            prior = prior + r.freq_delta
        return out


def _worker(rank, world, port, path):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    wl = synth.make_patterns(25, 4321, noliteral=1)
    logs = [synth.make_log(wl, n, 900 + i, pool_blocks=300, hit_rate=0.08) for i, n in enumerate([120, 40, 300, 15, 200])]
    eng = _OracleEngine(wl.patterns)
    prior0 = np.arange(len(wl.patterns), dtype=np.uint64) % 3
    res, (lo, hi), total = parallel.analyze_batch_sharded(eng, logs, rank, world, prior_count=prior0)
    np.savez(os.path.join(path, f"rank{rank}.npz"), lo=lo, hi=hi, total=total,
             lines=np.concatenate([r.line for r in res]) if res else np.zeros(0), scores=np.concatenate([r.score for r in res]) if res else np.zeros(0),
             pats=np.concatenate([r.pattern for r in res]) if res else np.zeros(0))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_matches_sequential(tmp_path, oracle):
    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    # sequential reference over the whole batch
    wl = synth.make_patterns(25, 4321, noliteral=1)
    logs = [synth.make_log(wl, n, 900 + i, pool_blocks=300, hit_rate=0.08) for i, n in enumerate([120, 40, 300, 15, 200])]
    prior = np.arange(len(wl.patterns), dtype=np.uint64) % 3
    exp_lines, exp_scores, exp_pats = [], [], []
    total = np.zeros(len(wl.patterns), np.uint64)
    for lg in logs:
        r = oracle.analyze(ScoringConfig(), wl.patterns, lg, prior_count=prior)
        exp_lines.append(r.line); exp_scores.append(r.score); exp_pats.append(r.pattern)
        prior = prior + r.freq_delta
        total += r.freq_delta
    parts = [np.load(os.path.join(tmp_path, f"rank{r}.npz")) for r in range(2)]
    assert int(parts[0]["lo"]) == 0 and int(parts[0]["hi"]) == int(parts[1]["lo"]) and int(parts[1]["hi"]) == len(logs)
    assert 0 < int(parts[0]["hi"]) < len(logs)
    np.testing.assert_array_equal(np.concatenate([p["lines"] for p in parts]), np.concatenate(exp_lines))
    np.testing.assert_array_equal(np.concatenate([p["pats"] for p in parts]), np.concatenate(exp_pats))
    np.testing.assert_array_equal(np.concatenate([p["scores"] for p in parts]), np.concatenate(exp_scores))   # bit-exact
    for p in parts:
        np.testing.assert_array_equal(p["total"], total)
    assert (np.concatenate(exp_scores) > 0).all() and total.sum() > 20
