"""One process driving all visible GPUs (lp_create(cfg, -1)): throughput of lp_analyze_batch from a page-locked host buffer
against the same call on one device. Workload: the C5-shaped batch of bench.py (logs of 20 000 lines).
Usage: python scripts/multi_device_timing.py [logs] [steps]"""
import ctypes
import json
import sys
import time

sys.path.insert(0, ".")
import numpy as np  # noqa: E402
import torch  # noqa: E402

from log_parser_b200 import binding, synth  # noqa: E402
from log_parser_b200.model import ScoringConfig  # noqa: E402


def main():
    n_logs = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    lines_per_log = 20_000
    wl = synth.make_patterns(500, synth.MASTER_SEED ^ 2, n_sets=5)
    text = synth.make_log(wl, n_logs * lines_per_log, synth.MASTER_SEED ^ 0x77)
    # cut the text into logs of lines_per_log lines
    offs, pos = [0], 0
    for _ in range(n_logs):
        for _k in range(lines_per_log):
            pos = text.index(b"\n", pos) + 1
        offs.append(pos)
    text = text[:pos]
    off = np.array(offs, dtype=np.uint64)
    hptr = binding.host_alloc(len(text))
    ctypes.memmove(hptr, text, len(text))
    out = {"logs": n_logs, "lines": n_logs * lines_per_log, "bytes": len(text), "gpus": torch.cuda.device_count()}
    ref = None
    for name, kw in (("one_device", {"device": 0}), ("all_devices_one_process", {"device": -1})):
        lp = binding.LogParser(ScoringConfig(), **kw)
        lp.load_patterns(wl.patterns)
        for _ in range(4):
            r = lp.analyze_batch_host_ptr(hptr, off)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            r = lp.analyze_batch_host_ptr(hptr, off)
        dt = (time.perf_counter() - t0) / steps
        out[name] = {"ms_per_step": 1e3 * dt, "lines_per_s": n_logs * lines_per_log / dt, "events": int(r.n_events)}
        if ref is None:
            ref = (np.array(r.line), np.array(r.score))
        else:
            out["identical_to_one_device"] = bool(np.array_equal(ref[0], r.line) and np.array_equal(ref[1], r.score))
        lp.close()
    binding.host_free(hptr)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
