#!/usr/bin/env python
"""Benchmark of the B200 hot path (BASELINE.json metric: log lines/s and GB/s matched+scored).

Workload at N=1: configs[1] (C2) - one 1M-line synthetic pod log vs 500 DFA-only patterns. At N>1 every
rank holds a batch of the configs[4] shape (C5): `--logs-per-gpu` distinct logs of `--lines-per-log` lines,
50 x 20 000 by default, i.e. the same 1M lines per GPU (weak scaling); the batch is sharded by log and the
per-pattern frequency counts are exchanged with one NCCL all-gather inside the library.

A step = one pass of the hot path over one GPU's input. `value` is timed with the input resident in HBM;
`e2e` goes through the host-buffer entry points (lp_analyze / lp_analyze_batch_dist) with page-locked host
buffers: host->device copy and result read-back inside the timed region. Before anything is timed, a
>= 20 000-line slice of the timed input goes through the same entry point and is compared with the oracle
(lines and patterns bit-exact, scores to 1e-9): `parity`.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "log lines/s matched+scored (1M-line log x 500 patterns)"


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.index = index
        self.sm, self.reasons, self.mx = [], set(), None
        self._stop = threading.Event()
        self.t = None
        self.err = None

    def start(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            self.nv = nv
            self.h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
        except Exception as e:   # noqa: BLE001
            self.err = "nvml unavailable: %s" % e
            return
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:   # noqa: BLE001
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception as e:   # noqa: BLE001
                self.err = str(e)
                break
            time.sleep(0.002)

    def stop(self):
        self._stop.set()
        if self.t:
            self.t.join(timeout=2)
        sm = sorted(self.sm)
        out = {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.mx, "reasons": sorted(self.reasons), "samples": len(sm)}
        if self.err:
            out["note"] = self.err
        return out


def build_workload(args, rank: int, batch: bool):
    """(workload, text, offsets): one log (C2), or this rank's C5-shaped batch cut from one generated text."""
    import numpy as np

    from log_parser_b200 import synth

    if not batch and args.config != "c2":      # the other single-log configs of BASELINE.json (parity cases; timed for profiles/)
        maker = {"c1": lambda: synth.config_c1(), "c3": lambda: synth.config_c3(args.lines, args.patterns),
                 "c4": lambda: synth.config_c4(args.lines, args.patterns)}[args.config]
        wl, log = maker()
        return wl, log, np.array([0, len(log)], dtype=np.uint64)
    wl = synth.make_patterns(args.patterns, synth.MASTER_SEED ^ 2, n_sets=5)
    if not batch:
        log = synth.make_log(wl, args.lines, synth.MASTER_SEED ^ 2 ^ 0x5252)
        return wl, log, np.array([0, len(log)], dtype=np.uint64)
    text = synth.make_log(wl, args.logs_per_gpu * args.lines_per_log, (synth.MASTER_SEED ^ 5 ^ 0x5252) + rank * 7919)
    nl = np.flatnonzero(np.frombuffer(text, dtype=np.uint8) == 10)
    cuts = nl[args.lines_per_log - 1::args.lines_per_log][:args.logs_per_gpu] + 1      # a log ends after its last line's newline
    off = np.concatenate([[0], cuts]).astype(np.uint64)
    off[-1] = len(text)
    return wl, text, off


def jvm_harness_run(wl, log, per_lines: int, steps: int, warmup: int):
    """oracle/RefHarness.java on the real java.util.regex, if a JDK exists on this box (none in the
    authoring image). One request on one thread, as the reference serves it. Returns None if unusable."""
    import base64
    import shutil
    import tempfile

    from log_parser_b200.model import severity_multiplier

    java = shutil.which("java")
    if not java:
        return None
    try:
        b = lambda x: base64.b64encode(x.encode("utf-8")).decode("ascii")
        rows = []
        for p in wl.patterns:
            f = [p.id or "", repr(p.confidence), repr(severity_multiplier(p.severity)), "1" if p.has_context else "0",
                 str(p.lines_before or 0), str(p.lines_after or 0), b(p.primary_regex), str(len(p.secondaries))]
            for s in p.secondaries:
                f += [b(s.regex), repr(s.weight), str(s.proximity_window)]
            f.append(str(len(p.sequences)))
            for q in p.sequences:
                f += [repr(q.bonus_multiplier), str(len(q.events))] + [b(e) for e in q.events]
            rows.append("\t".join(f))
        lines = log.split(b"\n")
        sample = b"\n".join(lines[:per_lines]) + b"\n"
        with tempfile.TemporaryDirectory() as d:
            open(os.path.join(d, "p.tsv"), "w").write("\n".join(rows) + "\n")
            open(os.path.join(d, "l.log"), "wb").write(sample)
            r = subprocess.run([java, os.path.join(ROOT, "oracle", "RefHarness.java"), os.path.join(d, "p.tsv"), os.path.join(d, "l.log"),
                                str(max(warmup, 3)), str(max(steps, 1))], capture_output=True, text=True, timeout=600)
        if r.returncode != 0:
            return None
        kv = dict(x.split("=") for x in r.stdout.strip().split())
        ms = float(kv["ms_per_step"])
        return {"value": float(kv["lines"]) / (ms / 1e3), "ms_per_step": ms, "cores": 1, "kind": "port",
                "sample": f"RefHarness.java on java.util.regex: 1 request x {per_lines} lines x {len(wl.patterns)} patterns, 1 thread",
                "sample_lines": per_lines, "bytes_per_line": len(sample) / max(1, per_lines)}
    except Exception:   # noqa: BLE001
        return None


def cpu_reference_run(args, wl, text, offsets, steps: int, warmup: int, batch: bool):
    """The reference's CPU path on a bounded sample of the workload: the reference serves one
    request on one thread (AnalysisService.java:89-113), so a single-log config is timed on ONE host core; the
    batched config runs one log per core on all host cores. Engine: RefHarness.java on java.util.regex when a
    JDK exists on the box, else the oracle's C++ restatement (no JVM in this image)."""
    from oracle import oracle as O
    from log_parser_b200.model import ScoringConfig

    per = args.ref_lines
    jvm = None if batch else jvm_harness_run(wl, text, per, steps, warmup)
    if jvm is not None:
        return jvm
    cores = os.cpu_count() or 1
    threads = max(1, min(cores, args.ref_threads or cores)) if batch else 1
    reqs = []
    for t in range(threads):
        k = t % (len(offsets) - 1)
        piece = text[int(offsets[k]):int(offsets[k + 1])]
        cut = -1
        for _ in range(per):                                   # the first `per` lines of log k
            cut = piece.find(b"\n", cut + 1)
            if cut < 0:
                break
        reqs.append(piece if cut < 0 else piece[:cut + 1])
    table = O.PatternTable(wl.patterns)
    cfg = ScoringConfig(max_window=1000) if args.config == "c4" and not batch else ScoringConfig()
    times, total = [], 0
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        ev, nl, ck = O.analyze_many(cfg, wl.patterns, reqs, threads=threads, memo=False, table=table)
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
            total += nl
    el = sum(times)
    shape = (f"{threads} requests at once, one per host core, each the first {per} lines of one 20k-line log" if batch
             else f"1 request of the first {per} lines of the log on 1 host core")
    return {
        "value": total / el, "ms_per_step": 1e3 * el / max(1, steps), "cores": threads, "kind": "port",
        "sample": f"{shape} x {len(wl.patterns)} patterns (regexes recompiled per request as AnalysisService.java:56-86 does; "
                  f"no memoisation; C++ restatement, JVM unavailable on this box)",
        "sample_lines": sum(r.count(b"\n") for r in reqs), "bytes_per_line": len(text) / max(1, text.count(b"\n", 0, 1 << 24) * len(text) / min(len(text), 1 << 24)),
    }


def parity_check(lp, wl, text, offsets, rank: int, world: int, batch: bool, scfg=None, min_lines: int = 20000):
    """A slice of the timed input through the timed entry point vs the oracle: lines and patterns bit-exact, scores
    to 1e-9 relative (they are bit-identical in practice). At N>1 every rank checks its own slice through the
    collective call, against the oracle primed with the counts of the lower ranks (FTS read-before-increment order)."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from log_parser_b200.model import ScoringConfig
    from oracle import oracle as O

    if batch:
        k = 1
        while k < len(offsets) - 1 and int(offsets[k]) < 1 or text.count(b"\n", 0, int(offsets[k])) < min_lines and k < len(offsets) - 1:
            k += 1
        sl, off = text[:int(offsets[k])], np.array(offsets[:k + 1], dtype=np.uint64)
    else:
        cut = -1
        for _ in range(min_lines):
            nxt = text.find(b"\n", cut + 1)
            if nxt < 0:
                break
            cut = nxt
        sl = text[:cut + 1] if cut >= 0 else text
        off = np.array([0, len(sl)], dtype=np.uint64)
    dev = torch.device("cuda", torch.cuda.current_device())
    d = torch.from_numpy(np.frombuffer(sl, dtype=np.uint8).copy()).to(dev)
    if world == 1:
        res = lp.analyze_device(d.data_ptr(), len(sl))
        prior = None
    else:
        res = lp.analyze_batch_dist_device(d.data_ptr(), off)
        mine = torch.from_numpy(res.freq_delta.astype(np.int64)).to(dev)
        allc = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allc, mine)
        prior = sum((c.cpu().numpy() for c in allc[:rank]), np.zeros(len(wl.patterns), np.int64)) if rank else None
    ok, lines, detail = True, 0, ""
    eb = res.log_event_begin
    cfg = scfg or ScoringConfig()
    table = O.PatternTable(wl.patterns)
    run_prior = None if prior is None else prior.astype(np.uint64)
    for j in range(len(off) - 1):       # the logs of the slice one after the other, as the reference would serve them
        ora = O.analyze(cfg, wl.patterns, sl[int(off[j]):int(off[j + 1])], prior_count=run_prior, table=table)
        lo, hi = int(eb[j]), int(eb[j + 1])
        lines += ora.n_lines
        same = (int(res.log_n_lines[j]) == ora.n_lines and hi - lo == len(ora.line) and np.array_equal(res.line[lo:hi], ora.line)
                and np.array_equal(res.pattern[lo:hi], ora.pattern)
                and bool(np.all(np.abs(res.score[lo:hi] - ora.score) <= 1e-9 * np.maximum(1.0, np.abs(ora.score)))))
        if not same:
            ok, detail = False, f"log {j} of the slice differs from the oracle"
            break
        run_prior = ora.freq_delta.astype(np.uint64) + (0 if run_prior is None else run_prior)
    bits = bool(ok and all(np.array_equal(res.score[int(eb[j]):int(eb[j + 1])].view(np.uint64), res.score[int(eb[j]):int(eb[j + 1])].view(np.uint64))
                           for j in range(len(off) - 1)))
    flag = torch.tensor([1 if ok else 0, lines], dtype=torch.int64, device=dev)
    if world > 1:
        allf = [torch.zeros_like(flag) for _ in range(world)]
        dist.all_gather(allf, flag)
        ok = all(int(f[0]) == 1 for f in allf)
        lines = sum(int(f[1]) for f in allf)
    out = {"checked_lines": int(lines), "events": int(res.n_events), "ok": bool(ok), "ranks_checked": world,
           "against": "oracle (C++ restatement of the reference path), through " + ("lp_analyze_device" if world == 1 else "lp_analyze_batch_dist_device (NCCL exchange)")}
    if detail:
        out["detail"] = detail
    del bits
    return out


def parity_check_lines(lp, wl, text, rank: int, world: int, scfg=None, min_lines: int = 20000):
    """Line-range mode: the first `min_lines` lines of the log are cut over the ranks like the timed input and go
    through the timed collective entry point; rank 0 puts the shards' events together and compares them with the
    oracle's answer for the slice as one log."""
    import numpy as np
    import torch.distributed as dist

    from log_parser_b200 import parallel
    from log_parser_b200.model import ScoringConfig
    from oracle import oracle as O

    cut = -1
    for _ in range(min_lines):
        nxt = text.find(b"\n", cut + 1)
        if nxt < 0:
            break
        cut = nxt
    sl = text[:cut + 1] if cut >= 0 else text
    res = parallel.analyze_range_dist(lp, sl, rank, world)
    mine = {"line": np.array(res.line), "pattern": np.array(res.pattern), "score": np.array(res.score), "n": int(res.log_n_lines[0])}
    parts = [None] * world
    dist.all_gather_object(parts, mine)
    out = {"checked_lines": 0, "events": 0, "ok": True, "ranks_checked": world,
           "against": "oracle (C++ restatement of the reference path), through lp_analyze_range_dist_device (NCCL exchange, shards put together in rank order)"}
    if rank == 0:
        ora = O.analyze(scfg or ScoringConfig(), wl.patterns, sl)
        line = np.concatenate([p["line"] for p in parts])
        pat = np.concatenate([p["pattern"] for p in parts])
        score = np.concatenate([p["score"] for p in parts])
        ok = (all(p["n"] == ora.n_lines for p in parts) and len(line) == len(ora.line) and np.array_equal(line, ora.line)
              and np.array_equal(pat, ora.pattern) and bool(np.all(np.abs(score - ora.score) <= 1e-9 * np.maximum(1.0, np.abs(ora.score)))))
        out.update(checked_lines=int(ora.n_lines), events=int(len(ora.line)), ok=bool(ok))
    box = [out]
    dist.broadcast_object_list(box, src=0)
    return box[0]


def _emit(line: dict) -> None:
    """The one JSON line, written to the process's original stdout."""
    os.write(_STDOUT_FD, (json.dumps(line) + "\n").encode())


_STDOUT_FD = 1


def main():
    # Libraries (NCCL's version banner, torchrun) write to fd 1: keep it for the JSON line alone.
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--lines", type=int, default=1_000_000, help="N=1 (and --shard lines): lines of the single log")
    ap.add_argument("--patterns", type=int, default=500)
    ap.add_argument("--logs-per-gpu", type=int, default=50, help="N>1: logs in each rank's batch")
    ap.add_argument("--lines-per-log", type=int, default=20_000, help="N>1: lines per log of the batch (configs[4])")
    ap.add_argument("--config", default="c2", choices=["c1", "c2", "c3", "c4"],
                    help="single-log workload: c2 = the headline config; c1 (10k lines x 50 patterns), c3 (look-around / back-reference "
                         "patterns), c4 (windows up to 1000 lines; use with --shard lines at N>1) for profiles/")
    ap.add_argument("--batch", action="store_true", help="use the batch workload at N=1 too (profiling the C5 shape on one GPU)")
    ap.add_argument("--ref-lines", type=int, default=4000, help="lines of one CPU request in the reference arm / cpu_baseline")
    ap.add_argument("--ref-threads", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiling runs)")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle comparison (profiling runs)")
    ap.add_argument("--shard", default="logs", choices=["logs", "lines"],
                    help="N>1 only. logs: a batch of logs per GPU (weak scaling, the default); lines: ONE log cut by line range "
                         "with halos over the GPUs (strong scaling, lp_range_*)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    by_lines = args.shard == "lines" and world > 1
    batch = (world > 1 and not by_lines) or args.batch
    lines_per_gpu = args.logs_per_gpu * args.lines_per_log if batch else args.lines
    if batch:
        config = {"workload": f"C5 shape: per GPU a batch of {args.logs_per_gpu} distinct synthetic Java pod logs x {args.lines_per_log} lines "
                              f"(~{{:.0f}} MB, the same {lines_per_gpu} lines per GPU as C2) vs {args.patterns} DFA-only patterns",
                  "patterns": args.patterns, "lines_per_log": args.lines_per_log, "logs_per_gpu": args.logs_per_gpu,
                  "parallelism": f"batch sharded by log over {world} GPU(s); per-pattern frequency counts all-gathered over NCCL inside the library"}
    else:
        what = {"c1": "C1: single 10000-line synthetic Java pod log (~{:.1f} MB) vs 50 patterns",
                "c2": f"C2: single {args.lines}-line synthetic Java pod log (~{{:.0f}} MB) vs {args.patterns} DFA-only patterns",
                "c3": f"C3: single {args.lines}-line log (~{{:.0f}} MB) vs {args.patterns} patterns, 12 % with look-around / back-references / possessive groups",
                "c4": f"C4: single {args.lines}-line log (~{{:.0f}} MB), dense secondary matches, proximity / context windows up to 1000 lines, {args.patterns} patterns"}[args.config]
        if args.config == "c1":
            args.lines = 10_000
        config = {"workload": what, "patterns": args.patterns, "lines_per_log": args.lines, "logs_per_gpu": 1, "parallelism": "1 GPU"}
    config["l2"] = "3 device copies of the input at distinct addresses, rotated per step (3 x 135 MB > 126 MB L2)"
    if by_lines:
        config["logs_per_gpu"] = 1.0 / world
        config["parallelism"] = (f"ONE log cut by line range over {world} GPUs with halos of lp_range_halo_lines() lines; "
                                 "one NCCL all-gather of the shard summaries (line counts, per-pattern counts, sequence-chain states) inside the library, composed on the device")
    if args.impl == "reference":
        if rank != 0:
            return 0
        wl, text, offsets = build_workload(args, 0, batch)
        config["workload"] = config["workload"].format(len(text) / 1e6)
        r = cpu_reference_run(args, wl, text, offsets, args.steps, args.warmup, batch)
        bpl = len(text) / max(1, lines_per_gpu)
        line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": "lines/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8/f64", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": r["value"], "unit": "lines/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
                "e2e": {"value": r["value"], "unit": "lines/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gb_per_s": r["value"] * bpl / 1e9,
                "note": "JVM absent in this image: the reference arm is the in-repo C++ restatement of the reference's java.util.regex + scoring path"}
        _emit(line)
        return 0

    import numpy as np
    import torch
    import torch.distributed as dist

    from log_parser_b200 import binding, build, parallel
    from log_parser_b200.model import ScoringConfig

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the B200 path has no CPU fallback")
    build.build()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # keep stdout for the one JSON line
        dist.init_process_group("nccl", device_id=dev)

    wl, text, offsets = build_workload(args, 0 if by_lines else rank, batch)
    nbytes = len(text)
    config["workload"] = config["workload"].format(nbytes / 1e6)
    scfg = ScoringConfig(max_window=1000) if args.config == "c4" and not batch else ScoringConfig()
    lp = binding.LogParser(scfg, device=local_rank)
    lp.load_patterns(wl.patterns)
    P = len(wl.patterns)
    if world > 1:
        parallel.init_library_comm(lp, rank, world)               # the library's own NCCL communicator

    log = text
    halo_b = halo_a = 0
    if by_lines:      # this rank keeps only its own line range + halos
        end = parallel.strip_trailing_separators(text)
        tb, te, halo_b, halo_a = parallel.split_line_ranges(text, world, lp.range_halo_lines(), end)[rank]
        log = bytes(text[tb:te])
        nbytes_all, nbytes = nbytes, len(log)
        offsets = np.array([0, nbytes], dtype=np.uint64)
    # page-locked host buffer from the library (what the JVM binding uses), wrapped as a tensor for the timed copies
    hptr = binding.host_alloc(max(nbytes, 1))
    if not hptr:
        raise SystemExit("lp_host_alloc failed")
    import ctypes
    host_np = np.ctypeslib.as_array((ctypes.c_uint8 * max(nbytes, 1)).from_address(hptr))
    host_np[:nbytes] = np.frombuffer(log, dtype=np.uint8)
    host = torch.from_numpy(host_np)
    copies = [host[:nbytes].to(dev, non_blocking=False).clone() for _ in range(3)]
    lib_stream = torch.cuda.ExternalStream(lp.stream_ptr(), device=dev)

    def step_device(i):
        d = copies[i % len(copies)]
        if by_lines:      # scan -> ncclAllGather of the shard summaries -> composition on the device -> score, one call
            return lp.analyze_range_dist_device(d.data_ptr(), nbytes, halo_b, halo_a)
        if world > 1:
            return lp.analyze_batch_dist_device(d.data_ptr(), offsets)   # scan -> ncclAllGather of (P + 1) x u64 -> score, one call
        if batch:
            return lp.analyze_batch_device(d.data_ptr(), offsets)
        return lp.analyze_device(d.data_ptr(), nbytes)

    def step_host():
        if by_lines:
            return lp.analyze_range_dist_host_ptr(hptr, nbytes, halo_b, halo_a)
        if world > 1:
            return lp.analyze_batch_dist_host_ptr(hptr, offsets)         # the copy is cut in pieces and overlaps the scan
        if batch:
            return lp.analyze_batch_host_ptr(hptr, offsets)
        return lp.analyze_host_ptr(hptr, nbytes)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up (also settles the scratch estimates), parity of a slice against the oracle, then K timed steps
    for i in range(args.warmup):
        res = step_device(i)
    parity = None
    if not args.no_parity and by_lines:
        parity = parity_check_lines(lp, wl, text, rank, world, scfg)
        for i in range(2):
            res = step_device(i)
    if not args.no_parity and not by_lines:
        parity = parity_check(lp, wl, text, offsets, rank, world, batch, scfg)
        for i in range(2):
            res = step_device(i)        # back to the estimates of the full-size input
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    keys = ["scan_ms", "scan_bytes", "total_ms", "verify_ms", "score_ms", "dfa_scan_ms", "lines_ms"]
    acc = {k: [] for k in keys}
    barrier()
    t0 = time.perf_counter()
    e0.record(lib_stream)
    for i in range(args.steps):
        res = step_device(i)
        t = lp.timings()
        launches += int(t["kernel_launches"])
        for k in keys:
            acc[k].append(t[k])
    e1.record(lib_stream)
    barrier()
    wall = time.perf_counter() - t0
    dev_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    n_events = res.n_events
    n_lines_seen = int(np.sum(res.log_n_lines))
    cand, tasks = t["candidates"], t["verify_tasks"]
    mean = {k: sum(v) / len(v) for k, v in acc.items()}

    el = torch.tensor([dev_ms / 1e3, wall], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(el, op=dist.ReduceOp.MAX)
    dev_s, wall_s = float(el[0]), float(el[1])
    lines_all = lines_per_gpu * (1 if by_lines else world)
    value = lines_all * args.steps / dev_s

    # ---- N>1, by log: the same per-GPU batch through the non-collective call on every GPU at once, so that the cost of
    # the exchange can be read off next to the weak-scaling value (the N=1 line of this bench is the single-log config
    # C2, a different workload: a batch of 50 logs takes ~0.1 ms longer per step than one log of the same size)
    alone = None
    if world > 1 and not by_lines:
        for i in range(3):
            lp.analyze_batch_device(copies[i % len(copies)].data_ptr(), offsets)
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(lib_stream)
        for i in range(args.steps):
            lp.analyze_batch_device(copies[i % len(copies)].data_ptr(), offsets)
        a1.record(lib_stream)
        barrier()
        al = torch.tensor([a0.elapsed_time(a1) / 1e3], dtype=torch.float64, device=dev)
        dist.all_reduce(al, op=dist.ReduceOp.MAX)
        alone = {"what": "the same per-GPU batch through lp_analyze_batch_device on all GPUs at once, no exchange (max over ranks)",
                 "ms_per_step": 1e3 * float(al[0]) / args.steps, "exchange_cost_ms_per_step": 1e3 * (dev_s - float(al[0])) / args.steps}
        for i in range(2):
            step_device(i)

    # ---- e2e: host buffer in, events out, through the host-buffer entry point (the call the JVM makes)
    e2e_steps = 0 if args.no_e2e else max(3, min(args.steps, 10))
    r2 = res
    for i in range(0 if args.no_e2e else 2):
        r2 = step_host()
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        r2 = step_host()
    barrier()
    e2e_wall = time.perf_counter() - t0
    el2 = torch.tensor([e2e_wall], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(el2, op=dist.ReduceOp.MAX)
    e2e_value = lines_all * e2e_steps / max(float(el2[0]), 1e-9)
    d2h = int(r2.n_events * (4 + 4 + 8 + 4) + P * 8 + 8 * 3)

    # the host-side ceiling of e2e: all ranks copy their input host->device at once, nothing else running
    h2d_ceiling = None
    if not args.no_e2e:
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        c0.record()
        for _ in range(5):
            copies[0].copy_(host[:nbytes], non_blocking=True)
        c1.record()
        barrier()
        tt = torch.tensor([c0.elapsed_time(c1) / 5e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        h2d_ceiling = {"lines_per_s": lines_all / float(tt[0]), "gb_per_s_per_gpu": nbytes / float(tt[0]) / 1e9,
                       "what": f"{world} concurrent pinned host->device copies of one step's input, nothing else running"}

    # ---- the same call from two request threads on the ONE context (what the reference's worker threads do with the
    # @ApplicationScoped service, AS:35): the context's workers let one request's host->device copy run under the other's
    # kernels. Informational; `e2e.value` stays the single synchronous caller.
    e2e_two = None
    if world == 1 and not args.no_e2e and not batch:
        hptr2 = binding.host_alloc(max(nbytes, 1))
        ctypes.memmove(hptr2, hptr, nbytes)
        for hp in (hptr, hptr2):
            for _ in range(3):
                lp.analyze_host_ptr(hp, nbytes)
        torch.cuda.synchronize()

        def worker(hp, n):
            for _ in range(n):
                lp.analyze_host_ptr(hp, nbytes)

        # both request threads at once, untimed: the context's second worker allocates its scratch and settles its
        # estimates only when two calls actually overlap (a single caller always lands on the first worker)
        warm = [threading.Thread(target=worker, args=(hp, 4)) for hp in (hptr, hptr2)]
        for th in warm:
            th.start()
        for th in warm:
            th.join()
        torch.cuda.synchronize()
        threads = [threading.Thread(target=worker, args=(hp, e2e_steps)) for hp in (hptr, hptr2)]
        t0 = time.perf_counter()
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        torch.cuda.synchronize()
        e2e_two = args.lines * len(threads) * e2e_steps / (time.perf_counter() - t0)
        binding.host_free(hptr2)

    if by_lines:
        nbytes = nbytes_all
    if rank == 0:
        peak, peak_src = _peaks()
        legacy = os.environ.get("LP_SCAN_MODE") == "legacy"
        kernel = ("k_filter_count (newline count + literal-gram prefilter over every byte)" if legacy else
                  "k_fscan (the one pass over the log: TMA-staged tiles, newline lists for String.split + literal-gram prefilter)")
        traffic, traffic_src = None, None          # DRAM bytes of one launch of that kernel from the committed ncu --set full capture (C2 only)
        tpath = os.path.join(ROOT, "profiles", "r01_v8_traffic.json" if legacy else "r02_fscan_traffic.json")
        if os.path.exists(tpath) and not batch and args.lines == 1_000_000 and args.patterns == 500 and not by_lines:
            tj = json.load(open(tpath))
            traffic, traffic_src = int(tj["dram_bytes_read"]) + int(tj["dram_bytes_write"]), tj["source"]
        scan_s = mean["scan_ms"] / 1e3
        achieved = mean["scan_bytes"] / scan_s / 1e9
        # SURVEY §8d scan phase: everything up to the match records (split + prefilter + verification), algorithmic bytes
        # B + 4 (N+1) + 8 M
        phase_bytes = nbytes + 4 * (n_lines_seen + 1) + 8 * int(n_events)
        phase_s = (mean["scan_ms"] + mean["verify_ms"]) / 1e3
        line = {
            "metric": METRIC, "value": value, "unit": "lines/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dev_s / args.steps, "higher_is_better": True, "scaling": "strong" if by_lines else "weak", "vs_baseline": None,
            "dtype": "u8/f64", "data": "synthetic", "config": config,
            "gb_per_s": value * nbytes / lines_per_gpu / 1e9,
            "wall_ms_per_step": 1e3 * wall_s / args.steps,
            "events_per_step_per_gpu": int(n_events), "candidates": int(cand), "verify_tasks": int(tasks),
            "phase_ms": {"scan": mean["scan_ms"], "verify": mean["verify_ms"], "score": mean["score_ms"],
                         "lines (inside verify)": mean["lines_ms"], "dfa_scan (inside verify)": mean["dfa_scan_ms"],
                         "device_total": mean["total_ms"]},
            "roofline": {"bound": "hbm", "kernel": kernel, "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": int(mean["scan_bytes"]),
                         "scan_phase": {"what": "SURVEY §8d scan phase (split + prefilter + verification, up to the match records): "
                                                "B + 4(N+1) + 8M bytes over scan + verify time",
                                        "bytes": int(phase_bytes), "ms": 1e3 * phase_s, "achieved": phase_bytes / phase_s / 1e9,
                                        "frac": phase_bytes / phase_s / 1e9 / peak}},
            "e2e": {"value": e2e_value, "unit": "lines/s", "h2d_bytes_per_step": int(nbytes), "d2h_bytes_per_step": d2h,
                    "gb_per_s": e2e_value * nbytes / lines_per_gpu / 1e9, "steps": e2e_steps,
                    "entry_point": "lp_analyze_range_dist" if by_lines else "lp_analyze_batch_dist" if world > 1 else "lp_analyze_batch" if batch else "lp_analyze",
                    "host_buffer": "page-locked, from lp_host_alloc", "h2d_ceiling": h2d_ceiling,
                    "two_request_threads": e2e_two},
            "parity": parity,
            "without_exchange": alone,
            "build": {"library": "log_parser_b200/liblogparser_b200.so", "mode": build.LAST_BUILD_MODE, "source_hash": build.source_hash(),
                      "matches_sources": not build.stale()},
            "gpu_launches": int(launches), "clocks": clocks,
        }
        if not args.no_cpu_baseline:
            r = cpu_reference_run(args, wl, text, offsets, steps=2, warmup=1, batch=batch)
            line["cpu_baseline"] = {"value": r["value"], "unit": "lines/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
        _emit(line)
    if world > 1:
        dist.destroy_process_group()
    lp.close()
    binding.host_free(hptr)
    return 0


if __name__ == "__main__":
    sys.exit(main())
