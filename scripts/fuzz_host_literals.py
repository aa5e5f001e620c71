"""CPU-only fuzz of the prefilter's soundness: every text a regex matches (oracle) must contain one of the literals the host
compiler extracted for it (rx_literal.cpp required_literals) - a missing literal would be a missed match on the device,
where only lines holding a literal's 4-byte gram reach verification. Random patterns from both generators glued to words
in several scripts, (?i) and (?iu) folding, literal lengths 1..4. Hunting tool.
Usage: python scripts/fuzz_host_literals.py <seed> <rounds>"""
import importlib.util
import random
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from log_parser_b200 import binding  # noqa: E402
from oracle import oracle  # noqa: E402
from rxgen import Gen  # noqa: E402

spec = importlib.util.spec_from_file_location("fz", "scripts/fuzz_host_scripts.py")
fz = importlib.util.module_from_spec(spec)
spec.loader.exec_module(fz)


def ascii_lower(b: bytes) -> bytes:      # icase literals are stored with ASCII letters in lower case; nothing else is folded
    return bytes(c + 32 if 65 <= c <= 90 else c for c in b)


WORDS = ["ERRx", "abc_1", "Exception", "x-1:E", "timeout ", "дата", "Жx1", "éa", "STRASSE", "straße", "İx", "Kelvin"]
rng = random.Random(int(sys.argv[1])); rounds = int(sys.argv[2])
n = matched = withlits = bad = 0
for it in range(rounds):
    k = rng.randrange(3)
    a = fz.pattern(rng) if k == 0 else Gen(rng, regular_only=rng.random() < 0.7, max_depth=3).regex()[0]
    w = rng.choice(WORDS); wj = "".join("\\" + c if c in ".-:" else c for c in w)
    j = [a, wj + a, a + wj, "(?:" + a + ")" + wj, "(?i)" + wj + "(?:" + a + ")?", "(?iu)" + wj + a, "(?:" + wj + "|" + a + ")", wj + "|" + rng.choice(WORDS)][rng.randrange(8)]
    ml = rng.choice([0, 0, 1, 2, 3, 4])
    try:
        o = oracle.Regex(j); d = binding.DebugRegex(j, min_literal=ml)
    except Exception:
        continue
    lits = d.literals()
    for _ in range(10):
        parts = []
        for _k in range(rng.randint(0, 4)):
            r = rng.random()
            parts.append(rng.choice(WORDS) if r < 0.4 else rng.choice(WORDS).upper() if r < 0.5 else rng.choice(WORDS).lower() if r < 0.6 else fz.text(rng, 6))
        t = "".join(parts)
        n += 1
        if d.match(t) < 0: continue
        if not o.find(t): continue
        matched += 1
        if not lits: continue
        withlits += 1
        tb = t.encode()
        if not any((ascii_lower(l) in ascii_lower(tb)) if ic else (l in tb) for ic, l in lits):
            bad += 1
            if bad < 8: print("UNSOUND", repr(j), repr(t), lits[:6], "min", ml)
print("checked", n, "matched", matched, "with literals", withlits, "bad", bad)
