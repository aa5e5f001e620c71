#!/usr/bin/env python
"""Benchmark of the B200 hot path (BASELINE.json metric: log lines/s and GB/s matched+scored).

Workload at N=1: configs[1] — one 1M-line synthetic pod log vs 500 DFA-only patterns. At N>1 each
rank holds one such log (a batch sharded by log, weak scaling) and the per-pattern frequency
counts are exchanged with one NCCL all-gather between the scan and score phases.

A step = one pass of the hot path over one log. `value` is timed with the log resident in HBM;
`e2e` goes through lp_analyze with a pinned host buffer (H2D + result D2H inside the timed region).
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


def build_workload(args, rank: int):
    from log_parser_b200 import synth

    wl = synth.make_patterns(args.patterns, synth.MASTER_SEED ^ 2, n_sets=5)
    log = synth.make_log(wl, args.lines, (synth.MASTER_SEED ^ 2 ^ 0x5252) + rank * 7919)
    return wl, log


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


def cpu_reference_run(args, wl, log, steps: int, warmup: int):
    """The reference's CPU path (oracle restatement; no JVM in this image) on a bounded sample.
    One request per host thread: the reference's only parallelism (AnalysisService.java:89-113)."""
    from oracle import oracle as O
    from log_parser_b200.model import ScoringConfig

    jvm = jvm_harness_run(wl, log, args.ref_lines * 4, steps, warmup)
    if jvm is not None:
        return jvm
    cores = os.cpu_count() or 1
    threads = max(1, min(cores, args.ref_threads or cores))
    per = args.ref_lines
    lines = log.split(b"\n")
    n_avail = len(lines) - 1
    reqs = []
    for t in range(threads):
        lo = (t * per) % max(1, n_avail - per)
        reqs.append(b"\n".join(lines[lo:lo + per]) + b"\n")
    table = O.PatternTable(wl.patterns)
    cfg = ScoringConfig()
    times, total = [], 0
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        ev, nl, ck = O.analyze_many(cfg, wl.patterns, reqs, threads=threads, memo=False, table=table)
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
            total += nl
    el = sum(times)
    return {
        "value": total / el, "ms_per_step": 1e3 * el / max(1, steps), "cores": threads, "kind": "port",
        "sample": f"{threads} concurrent requests x {per} lines of the same log x {len(wl.patterns)} patterns "
                  f"(regexes recompiled per request as AnalysisService.java:56-86 does; no memoisation)",
        "sample_lines": threads * per, "bytes_per_line": len(log) / max(1, n_avail),
    }


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
    ap.add_argument("--lines", type=int, default=1_000_000)
    ap.add_argument("--patterns", type=int, default=500)
    ap.add_argument("--ref-lines", type=int, default=1500, help="lines per concurrent CPU request in the reference arm")
    ap.add_argument("--ref-threads", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiling runs)")
    ap.add_argument("--exchange", default="nccl", choices=["nccl", "torch"],
                    help="N>1, --shard logs: nccl = lp_analyze_batch_dist_device (NCCL inside the library), torch = all-gather via torch.distributed")
    ap.add_argument("--shard", default="logs", choices=["logs", "lines"],
                    help="N>1 only. logs: one C2 log per GPU (weak scaling, the default); lines: ONE C2 log cut by line range "
                         "with halos over the GPUs (strong scaling, lp_range_*)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    config = {"workload": f"C2: single {args.lines}-line synthetic Java pod log (~{{:.0f}} MB) vs {args.patterns} DFA-only patterns per GPU",
              "patterns": args.patterns, "lines_per_log": args.lines, "logs_per_gpu": 1,
              "parallelism": f"sharded by log over {world} GPU(s); per-pattern frequency counts all-gathered over NCCL" if world > 1 else "1 GPU",
              "l2": "3 device copies of the log at distinct addresses, rotated per step (420 MB > 126 MB L2)"}

    by_lines = args.shard == "lines" and world > 1
    if by_lines:
        config["logs_per_gpu"] = 1.0 / world
        config["parallelism"] = (f"ONE log cut by line range over {world} GPUs with halos of lp_range_halo_lines() lines; "
                                 "one NCCL all-gather of the shard summaries (line counts, per-pattern counts, sequence-chain states)")
    if args.impl == "reference":
        if rank != 0:
            return 0
        wl, log = build_workload(args, 0)
        config["workload"] = config["workload"].format(len(log) / 1e6)
        r = cpu_reference_run(args, wl, log, args.steps, args.warmup)
        line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": "lines/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8/f64", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": r["value"], "unit": "lines/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
                "e2e": {"value": r["value"], "unit": "lines/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gb_per_s": r["value"] * r["bytes_per_line"] / 1e9,
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
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # keep stdout for the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    wl, log = build_workload(args, 0 if by_lines else rank)
    nbytes = len(log)
    config["workload"] = config["workload"].format(nbytes / 1e6)
    lp = binding.LogParser(ScoringConfig(), device=local_rank)
    lp.load_patterns(wl.patterns)
    P = len(wl.patterns)

    halo_b = halo_a = 0
    if by_lines:      # this rank keeps only its own line range + halos
        end = parallel.strip_trailing_separators(log)
        tb, te, halo_b, halo_a = parallel.split_line_ranges(log, world, lp.range_halo_lines(), end)[rank]
        log = bytes(log[tb:te])
        nbytes_all, nbytes = nbytes, len(log)
    host = torch.from_numpy(np.frombuffer(log, dtype=np.uint8).copy()).pin_memory()
    copies = [host.to(f"cuda:{local_rank}", non_blocking=False).clone() for _ in range(3)]
    off = np.array([0, nbytes], dtype=np.uint64)
    lib_stream = torch.cuda.ExternalStream(lp.stream_ptr(), device=torch.device("cuda", local_rank))

    exchange = None
    if world > 1 and not by_lines:
        if args.exchange == "torch":
            exchange = parallel.DeviceExchange(lp, rank, world)       # all-gather issued from torch.distributed
        else:
            parallel.init_library_comm(lp, rank, world)               # the library's own NCCL communicator

    def step_device(i):
        d = copies[i % len(copies)]
        if world == 1:
            return lp.analyze_device(d.data_ptr(), nbytes)
        if by_lines:
            summary = lp.range_scan_device(d.data_ptr(), nbytes, halo_b, halo_a)
            return lp.range_score(parallel.gather_summaries(summary, world), rank, world)
        if exchange is not None:
            return exchange.step(d.data_ptr(), off)
        return lp.analyze_batch_dist_device(d.data_ptr(), off)   # scan -> ncclAllGather of (P + 1) x u64 -> score, one call

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up, then K timed device-resident steps
    for i in range(args.warmup):
        res = step_device(i)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    scan_ms, scan_bytes, tot_dev_ms, ver_ms, sc_ms, ds_ms, ds_bytes = [], [], [], [], [], [], []
    barrier()
    t0 = time.perf_counter()
    e0.record(lib_stream)
    for i in range(args.steps):
        res = step_device(i)
        t = lp.timings()
        launches += int(t["kernel_launches"])
        scan_ms.append(t["scan_ms"]); scan_bytes.append(t["scan_bytes"]); tot_dev_ms.append(t["total_ms"])
        ver_ms.append(t["verify_ms"]); sc_ms.append(t["score_ms"]); ds_ms.append(t["dfa_scan_ms"]); ds_bytes.append(t["dfa_scan_bytes"])
    e1.record(lib_stream)
    barrier()
    wall = time.perf_counter() - t0
    dev_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    n_events = res.n_events
    cand, tasks = t["candidates"], t["verify_tasks"]

    el = torch.tensor([dev_ms / 1e3, wall], dtype=torch.float64, device=f"cuda:{local_rank}")
    if world > 1:
        dist.all_reduce(el, op=dist.ReduceOp.MAX)
    dev_s, wall_s = float(el[0]), float(el[1])
    total_lines = args.lines * (1 if by_lines else world) * args.steps
    value = total_lines / dev_s

    # ---- e2e: host buffer in, events out, through lp_analyze (the call the JVM makes)
    e2e_steps = 0 if args.no_e2e else max(3, min(args.steps, 10))
    r2 = res

    def step_host():
        if world == 1:
            return lp.analyze_host_ptr(host.data_ptr(), nbytes)
        copies[0].copy_(host, non_blocking=False)             # this rank's share, pinned host -> device
        if not by_lines:
            return step_device(0)                             # the same exchange as the device-resident steps
        summary = lp.range_scan_device(copies[0].data_ptr(), nbytes, halo_b, halo_a)
        return lp.range_score(parallel.gather_summaries(summary, world), rank, world)

    for i in range(0 if args.no_e2e else 2):
        r2 = step_host()
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        r2 = step_host()
    barrier()
    e2e_wall = time.perf_counter() - t0
    el2 = torch.tensor([e2e_wall], dtype=torch.float64, device=f"cuda:{local_rank}")
    if world > 1:
        dist.all_reduce(el2, op=dist.ReduceOp.MAX)
    e2e_value = args.lines * (1 if by_lines else world) * e2e_steps / max(float(el2[0]), 1e-9)
    if by_lines:
        nbytes = nbytes_all
    d2h = int(r2.n_events * (4 + 4 + 8 + 4) + P * 8 + 8 * 3)

    # ---- the same call from two request threads, each with its own context (what the reference's worker
    # threads would do): one request's host->device copy overlaps the other's kernels. Informational;
    # `e2e.value` stays the single synchronous caller.
    e2e_two = None
    if world == 1 and not args.no_e2e:
        import threading
        lp2 = binding.LogParser(ScoringConfig(), device=local_rank)
        lp2.load_patterns(wl.patterns)
        parsers = [lp, lp2]
        for q in parsers:
            for _ in range(3):
                q.analyze_host_ptr(host.data_ptr(), nbytes)
        torch.cuda.synchronize()

        def worker(q):
            for _ in range(e2e_steps):
                q.analyze_host_ptr(host.data_ptr(), nbytes)

        threads = [threading.Thread(target=worker, args=(q,)) for q in parsers]
        t0 = time.perf_counter()
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        torch.cuda.synchronize()
        e2e_two = args.lines * len(parsers) * e2e_steps / (time.perf_counter() - t0)
        lp2.close()

    if rank == 0:
        peak, peak_src = _peaks()
        traffic, traffic_src = None, None          # DRAM bytes of one k_filter_count launch from the committed ncu --set full capture (C2 only)
        tpath = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01_v8_traffic.json")
        if os.path.exists(tpath) and args.lines == 1_000_000 and args.patterns == 500 and not by_lines:
            tj = json.load(open(tpath))
            traffic, traffic_src = int(tj["dram_bytes_read"]) + int(tj["dram_bytes_write"]), tj["source"]
        scan_s = sum(scan_ms) / 1e3 / len(scan_ms)
        achieved = (sum(scan_bytes) / len(scan_bytes)) / scan_s / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": "lines/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dev_s / args.steps, "higher_is_better": True, "scaling": "strong" if by_lines else "weak", "vs_baseline": None,
            "dtype": "u8/f64", "data": "synthetic", "config": config,
            "gb_per_s": value * nbytes / args.lines / 1e9,
            "wall_ms_per_step": 1e3 * wall_s / args.steps,
            "events_per_log": int(n_events), "candidates": int(cand), "verify_tasks": int(tasks),
            "phase_ms": {"scan": sum(scan_ms) / len(scan_ms), "verify": sum(ver_ms) / len(ver_ms), "score": sum(sc_ms) / len(sc_ms),
                         "dfa_scan (inside verify)": sum(ds_ms) / len(ds_ms), "device_total": sum(tot_dev_ms) / len(tot_dev_ms)},
            "roofline": {"bound": "hbm", "kernel": "k_filter_count (newline count + literal-gram prefilter over every byte)", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": int(sum(scan_bytes) / len(scan_bytes))},
            "e2e": {"value": e2e_value, "unit": "lines/s", "h2d_bytes_per_step": int(nbytes), "d2h_bytes_per_step": d2h,
                    "gb_per_s": e2e_value * nbytes / args.lines / 1e9, "steps": e2e_steps,
                    "two_request_threads": e2e_two},
            "gpu_launches": int(launches), "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_reference_run(args, wl, log, steps=1, warmup=0)
            line["cpu_baseline"] = {"value": r["value"], "unit": "lines/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
        _emit(line)
    if exchange is not None:
        exchange.close()
    if world > 1:
        dist.destroy_process_group()
    lp.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
