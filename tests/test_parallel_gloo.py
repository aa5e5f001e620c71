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


# ---- one log sharded by line range: host-side cutting and the summary exchange
def _java_split(log: bytes):
    import re
    parts = re.split(rb"\r?\n", log)
    while parts and parts[-1] == b"":
        parts.pop()
    return parts if (parts or log) else [b""]      # "" -> [""], only separators -> []


def test_split_line_ranges_reassemble_to_java_split():
    import random
    rnd = random.Random(77)
    alphabet = [b"a", b"bc", b"", b"x\r", b"\r", b"ERROR q", b"  at f(G.java:1)"]
    for trial in range(300):
        n = rnd.randrange(0, 40)
        log = b"".join(rnd.choice(alphabet) + rnd.choice([b"\n", b"\r\n"]) for _ in range(n))
        if rnd.random() < 0.5:
            log += rnd.choice([b"tail", b"\r", b"t\r"])
        log += rnd.choice([b"", b"\n", b"\r\n\n", b"\n\r\n"])
        expect = _java_split(log)
        end = parallel.strip_trailing_separators(log)
        if end == 0 and len(log) > 0:
            assert expect == []
            continue
        for world in (1, 2, 3, 5, 8):
            halo = rnd.randrange(1, 7)
            pieces = parallel.split_line_ranges(log, world, halo, end)
            assert len(pieces) == world
            own_all = []
            for tb, te, before, after in pieces:
                lines = log[tb:te].split(b"\n")
                lines = [ln[:-1] if ln.endswith(b"\r") and i + 1 < len(lines) else ln for i, ln in enumerate(lines)]
                # a '\r' ending the very last line of a text is content only when no '\n' followed it in the log
                assert before + after <= len(lines)
                assert before <= halo and after <= halo
                own = lines[before:len(lines) - after]
                own_all.extend(own)
                # halos are as long as the log allows
                first_own = len(own_all) - len(own)
                assert before == min(halo, first_own)
                assert after == min(halo, len(expect) - first_own - len(own))
            assert own_all == expect, (log, world, halo)


def _gather_worker(rank, world, port, path):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = (np.arange(11, dtype=np.uint64) + 1000 * rank) | (np.uint64(rank + 1) << np.uint64(63 - rank))
    allv = parallel.gather_summaries(mine, world)
    np.save(os.path.join(path, f"g{rank}.npy"), allv)
    dist.barrier()
    dist.destroy_process_group()


def test_gather_summaries_world_size_2(tmp_path):
    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_gather_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    exp = np.stack([(np.arange(11, dtype=np.uint64) + 1000 * r) | (np.uint64(r + 1) << np.uint64(63 - r)) for r in range(2)])
    for r in range(2):
        np.testing.assert_array_equal(np.load(os.path.join(tmp_path, f"g{r}.npy")), exp)   # rank order, all 64 bits kept
