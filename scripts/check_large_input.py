"""One call on a device-resident log of more than 2 GiB (offsets are 32-bit, the documented limit is 4 GiB - 16):
a block of synthetic log tiled 40 times must give 40 times the block's events, at the block's line numbers."""
import sys

sys.path.insert(0, ".")
import numpy as np  # noqa: E402
import torch  # noqa: E402
from log_parser_b200 import binding, synth  # noqa: E402
from log_parser_b200.model import ScoringConfig  # noqa: E402


def main():
    wl = synth.make_patterns(500, synth.MASTER_SEED ^ 2, n_sets=5)
    block = synth.make_log(wl, 480_000, 777)
    reps = 40
    lp = binding.LogParser(ScoringConfig())
    lp.load_patterns(wl.patterns)
    one = lp.analyze(block, options=binding.OPT_NO_RANK)
    n_block = one.n_lines
    dev_block = torch.from_numpy(np.frombuffer(block, dtype=np.uint8).copy()).cuda()
    big = dev_block.repeat(reps)
    print("bytes", big.numel(), "GiB", big.numel() / 2 ** 30)
    res = lp.analyze_device(big.data_ptr(), big.numel(), options=binding.OPT_NO_RANK)
    ok = res.n_lines == n_block * reps and res.n_events == one.n_events * reps
    print("lines", res.n_lines, n_block * reps, "events", res.n_events, one.n_events * reps)
    if ok:
        line0 = (res.line.astype(np.int64) - 1) % n_block + 1
        ok = np.array_equal(line0.reshape(reps, -1), np.tile(one.line.astype(np.int64), (reps, 1))) and \
            np.array_equal(res.pattern.reshape(reps, -1), np.tile(one.pattern, (reps, 1)))
    print("OK" if ok else "MISMATCH")
    lp.close()


if __name__ == "__main__":
    main()
