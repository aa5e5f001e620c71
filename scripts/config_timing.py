"""Device-resident step time of the non-headline configs (C3: non-regular patterns, C4: large windows,
C5: a batch of logs) at full or reduced size; prints one JSON object per config. Tuning / DESIGN.md aid."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from log_parser_b200 import binding, synth  # noqa: E402
from log_parser_b200.model import ScoringConfig  # noqa: E402


def run(name, cfg, patterns, logs, steps=10):
    lp = binding.LogParser(cfg)
    lp.load_patterns(patterns)
    blob = b"".join(logs)
    off = np.zeros(len(logs) + 1, np.uint64)
    np.cumsum([len(x) for x in logs], out=off[1:])
    dev = torch.from_numpy(np.frombuffer(blob, dtype=np.uint8).copy()).cuda()
    for _ in range(4):
        r = lp.analyze_batch_device(dev.data_ptr(), off)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        r = lp.analyze_batch_device(dev.data_ptr(), off)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    t = lp.timings()
    n_lines = int(np.sum(r.log_n_lines))
    print(json.dumps({"config": name, "logs": len(logs), "lines": n_lines, "bytes": len(blob), "events": r.n_events,
                      "ms_per_step": dt * 1e3, "lines_per_s": n_lines / dt, "phase_ms": {k: round(t[k], 3) for k in ("scan_ms", "verify_ms", "score_ms", "total_ms")},
                      "candidates": t["candidates"], "verify_tasks": t["verify_tasks"], "launches": t["kernel_launches"]}))
    lp.close()


def main():
    which = sys.argv[1:] or ["c3", "c4", "c5"]
    if "c3" in which:
        wl, log = synth.config_c3()
        run("C3 1M lines, 12% non-regular patterns", ScoringConfig(), wl.patterns, [log])
    if "c4" in which:
        wl, log = synth.config_c4()
        run("C4 1M lines, windows up to 1000", ScoringConfig(max_window=1000), wl.patterns, [log])
    if "c5" in which:
        wl, logs = synth.config_c5_logs(150)
        run("C5 slice: 150 logs x 20k lines", ScoringConfig(), wl.patterns, logs, steps=5)


if __name__ == "__main__":
    main()
