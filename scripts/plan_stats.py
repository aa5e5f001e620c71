"""Tuning aid (CPU only): which literals of the C2 pattern set produce the prefilter candidates and
verify tasks. Counts literal occurrences in a slice of the synthetic log with bytes.count."""
import ctypes as C
import collections
import sys

sys.path.insert(0, ".")
from log_parser_b200 import binding, synth  # noqa: E402


def plan_dump(patterns, sample: bytes) -> str:
    L = binding.load_debug_library()
    tab = binding.PatternTable(patterns)
    L.lp_debug_plan_dump.restype = C.c_size_t
    L.lp_debug_plan_dump.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    n = L.lp_debug_plan_dump(tab.arr, len(patterns), sample, len(sample), None, 0)
    buf = C.create_string_buffer(n + 1)
    L.lp_debug_plan_dump(tab.arr, len(patterns), sample, len(sample), buf, n + 1)
    return buf.value.decode("utf-8", "replace")


def main():
    lines = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
    wl = synth.make_patterns(500, synth.MASTER_SEED ^ 2, n_sets=5)
    log = synth.make_log(wl, lines, synth.MASTER_SEED ^ 2 ^ 0x5252)
    low = log.lower()
    dump = plan_dump(wl.patterns, log[: 1 << 20])
    per_rx = collections.Counter()
    per_mode = collections.Counter()
    rows = []
    for ln in dump.splitlines():
        if ln.startswith("#") or ln.startswith("!"):
            if ln.startswith("#regexes") or ln.startswith("#group") or ln.startswith("#always"):
                print(ln[:200])
            continue
        head, lit = ln.split("\t", 2)[0:2], ln.split("\t", 2)[2]
        gram, rest = head[0], head[1]
        icase, rxs = rest.split(":")
        n = (low if icase == "1" else log).count(lit.lower().encode() if icase == "1" else lit.encode())
        pairs = [p for p in rxs.split(",") if p]
        for p in pairs:
            rx, mode = p.split("/")
            per_rx[int(rx)] += n
            per_mode[{0: "line walk"}.get(int(mode), "boundary test" if int(mode) & 1 else "anchored fwd" if int(mode) & 8 else "anchored rev")] += n
        rows.append((n, lit, icase, rxs))
    rows.sort(reverse=True)
    print("occurrences x regexes by mode:", dict(per_mode), "per line:", {k: v / lines for k, v in per_mode.items()})
    for n, lit, icase, rxs in rows[:40]:
        print(f"{n:8d} {n / lines:7.4f}/line  i={icase} {lit!r:24} -> {rxs[:60]}")


if __name__ == "__main__":
    main()
