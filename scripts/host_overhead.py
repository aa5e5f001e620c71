"""Where does the per-call host time go? (diagnostic, run on the GPU box)"""
import ctypes as C, sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from log_parser_b200 import binding, synth
from log_parser_b200.model import ScoringConfig

wl, log = synth.config_c2(n_lines=int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000)
lp = binding.LogParser(ScoringConfig()); lp.load_patterns(wl.patterns)
d = torch.from_numpy(np.frombuffer(log, np.uint8).copy()).cuda()
for _ in range(4): lp.analyze_device(d.data_ptr(), d.numel())
torch.cuda.synchronize()
K = 20
t0 = time.perf_counter()
for _ in range(K): r = lp.analyze_device(d.data_ptr(), d.numel())
t1 = time.perf_counter()
out = C.c_void_p()
for _ in range(K):
    rc = lp.L.lp_analyze_device(lp.h, d.data_ptr(), d.numel(), None, 0, C.byref(out)); lp.L.lp_result_free(out)
t2 = time.perf_counter()
for _ in range(K):
    rc = lp.L.lp_analyze_device(lp.h, d.data_ptr(), d.numel(), None, binding.OPT_NO_RANK, C.byref(out)); lp.L.lp_result_free(out)
t3 = time.perf_counter()
tm = lp.timings()
print("full python call %.3f ms | raw C call %.3f ms | raw no-rank %.3f ms | device_total %.3f ms | launches %d" % (
    1e3 * (t1 - t0) / K, 1e3 * (t2 - t1) / K, 1e3 * (t3 - t2) / K, tm["total_ms"], tm["kernel_launches"]))
