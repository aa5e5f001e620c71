"""Wall time of one device-resident lp_analyze_device call: the raw C call, the Python wrapper, and the
first-to-last-kernel time the library reports, on C2. Tuning aid (run on a GPU box)."""
import sys, time, ctypes as C
sys.path.insert(0, ".")
import numpy as np, torch
from log_parser_b200 import binding, synth
from log_parser_b200.model import ScoringConfig
wl = synth.make_patterns(500, synth.MASTER_SEED ^ 2, n_sets=5)
log = synth.make_log(wl, 1_000_000, synth.MASTER_SEED ^ 2 ^ 0x5252)
lp = binding.LogParser(ScoringConfig()); lp.load_patterns(wl.patterns)
dev = torch.from_numpy(np.frombuffer(log, dtype=np.uint8).copy()).cuda()
for _ in range(5): r = lp.analyze_device(dev.data_ptr(), len(log))
def timeit(f, n=30):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
L = lp.L
def raw():
    out = C.c_void_p()
    L.lp_analyze_device(lp.h, dev.data_ptr(), len(log), None, 0, C.byref(out))
    L.lp_result_free(out)
def wrapped(): lp.analyze_device(dev.data_ptr(), len(log))
def wrapped_t(): lp.analyze_device(dev.data_ptr(), len(log)); lp.timings()
print("raw C call us", timeit(raw)); print("wrapper us", timeit(wrapped)); print("wrapper+timings us", timeit(wrapped_t)); print("device_total us", lp.timings()["total_ms"] * 1e3)
