"""CPU-only fuzz of the always-on scan's product automata (what k_dfa_scan walks): random groups of regular regexes,
texts with non-ASCII characters, the serialised tables' hit mask against the oracle regex by regex.
Usage: python scripts/fuzz_scan_groups.py [seed] [rounds]"""
import ctypes as C
import random
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from log_parser_b200 import binding  # noqa: E402
from oracle import oracle  # noqa: E402
from rxgen import Gen  # noqa: E402

ALPHA = "abcERx_1 .-:\r\u00e9\u20ac\u0085\u2028"


def main():
    seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 600
    L = binding.load_debug_library()
    L.lp_debug_group_match.restype = C.c_uint32
    L.lp_debug_group_match.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.c_size_t, C.c_char_p, C.c_size_t]
    rng = random.Random(seed)
    bad = n = skipped = 0
    for _ in range(rounds):
        regs, oras = [], []
        want = rng.randint(1, 4)
        while len(regs) < want:
            j, _p = Gen(rng, regular_only=True, max_depth=2).regex()
            try:
                o = oracle.Regex(j)
                d = binding.DebugRegex(j)
            except Exception:
                continue
            if not d.is_dfa:
                continue
            regs.append(j)
            oras.append(o)
        arr = (C.c_char_p * len(regs))(*[r.encode() for r in regs])
        fl = (C.c_int * len(regs))(*([0] * len(regs)))
        for _k in range(12):
            t = "".join(rng.choice(ALPHA) for _ in range(rng.randint(0, 14)))
            tb = t.encode()
            m = L.lp_debug_group_match(arr, fl, len(regs), tb, len(tb))
            if m == 0xFFFFFFFF:
                skipped += 1
                break
            exp = 0
            for k, o in enumerate(oras):
                if o.find(t):
                    exp |= 1 << k
            n += 1
            if m != exp:
                bad += 1
                if bad < 6:
                    print("BAD", regs, repr(t), hex(m), hex(exp))
    print("seed", seed, "checked", n, "bad", bad, "groups not buildable", skipped)


if __name__ == "__main__":
    main()
