"""Writes tests/golden/regex_pin_corpus.tsv.gz: (pattern, flags, text) -> found / span as the ORACLE answers them, for
  * tests/test_oracle_golden.py::test_pin_corpus_* (oracle and product host compiler must keep giving these answers), and
  * oracle/RegexPin.java, which replays the same file on the real java.util.regex of a box that has a JDK (none in
    the authoring image) and lists every disagreement: that run is what pins the oracle to the engine the reference uses
    (AnalysisService.java:64,93-95; the engine is the JDK's, pom.xml:10).
The cases come from the generators the fuzzers use (tests/rxgen.py, scripts/fuzz_host_scripts.py,
scripts/fuzz_host_lookahead.py shapes) with fixed seeds, plus the dialect fixtures of regex_dialect_cases.json.
Line format, TAB separated: base64(pattern UTF-8), flags (java.util.regex.Pattern bit values), base64(text UTF-8),
found (0/1), start, end (code points; -1 -1 when not found).     Run from the repo root:
    python tests/golden/make_regex_pin_corpus.py
"""
import base64
import gzip
import importlib.util
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from rxgen import Gen, gen_text  # noqa: E402

OUT = os.path.join(HERE, "regex_pin_corpus.tsv.gz")


def load_script(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "scripts", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def b64(s: str) -> str:
    return base64.b64encode(s.encode("utf-8")).decode("ascii")


def main():
    rows = []

    def add(pattern: str, flags: int, text: str):
        try:
            o = O.Regex(pattern, flags)
        except (O.OracleSyntaxError, O.OracleUnsupported):
            return False
        sp = o.span(text)
        rows.append("\t".join([b64(pattern), str(flags), b64(text), "1" if sp else "0", str(sp[0] if sp else -1), str(sp[1] if sp else -1)]))
        return True

    # 1. the ASCII-centred generator of the differential tests: regular and backtracking constructs
    rng = random.Random(20261)
    alphabet = "abcERx_1 .-:\ré€ "
    n = 0
    while n < 1500:
        j, _p = Gen(rng, regular_only=(n % 3 != 0), max_depth=3).regex()
        for _k in range(2):
            t = gen_text(rng) if rng.random() < 0.5 else "".join(rng.choice(alphabet) for _ in range(rng.randint(0, 16)))
            if add(j, 0, t):
                n += 1
    # 2. multi-script text, property classes, (?U), (?iu), scoped flags, line terminators
    fz = load_script("fuzz_host_scripts")
    rng = random.Random(20262)
    n = 0
    while n < 2500:
        j = fz.pattern(rng)
        for _k in range(2):
            if add(j, 0, fz.text(rng)):
                n += 1
    # 3. leading look-aheads with an unbounded walk in front of a rest with captures / back-references
    fl = load_script("fuzz_host_lookahead")
    rng = random.Random(20263)
    n = 0
    while n < 600:
        cond = rng.choice(["E", "x1", r"\d", "[a-c]:", r"\bab"])
        look = rng.choice(["(?=", "(?!"]) + rng.choice(fl.WALKS) + cond + ")"
        j = look + Gen(rng, regular_only=rng.random() < 0.5, max_depth=2).regex()[0]
        if add(j, 0, fl.text(rng, 24)):
            n += 1
    # 4. the context regexes of ContextAnalysisService.java:27-34 with their flags, on log-like words
    ctx = [(r"\b(ERROR|FATAL|CRITICAL|SEVERE)\b", 2), (r"\b(WARN|WARNING)\b", 2), (r"^\s*at\s+[\w\.\$]+\(.*\)\s*$", 0), (r"\b\w*Exception\b|\b\w*Error\b", 0)]
    words = ["ERROR", "error", "Fatal", "WARN", "warning", "xERROR", "ERRORS", "at", "\tat", "com.foo.Bar$1.run(Bar.java:12)", "NullPointerException",
             "OutOfMemoryError", "Errors", "(x)", " ", "é", "\r", "_", "-", "MyError:"]
    rng = random.Random(20264)
    for pat, flags in ctx:
        for _ in range(100):
            add(pat, flags, " ".join(rng.choice(words) for _ in range(rng.randint(0, 5))))
    # 5. the dialect fixtures (java.util.regex vs PCRE2 by design)
    for c in json.load(open(os.path.join(HERE, "regex_dialect_cases.json")))["cases"]:
        add(c["pattern"], 0, c["text"])
    # 6. word boundaries next to combining marks: Pattern.Bound counts a non-spacing mark after a letter or digit as a
    #    word character; the oracle does not. RegexPin.java is expected to list exactly these.
    for pat, text in [("a\\b", "a\u0301"), ("a\\B", "a\u0301"), ("\\bx", "a\u0301 x"), ("\u0301\\b", "a\u0301 "), ("\u0301\\B", "a\u0301 "),
                      ("\\b\u0301", "\u00e9\u0301"), ("\\B\u0301", "\u00e9\u0301")]:
        add(pat, 0, text)
    with gzip.GzipFile(OUT, "wb", mtime=0) as f:
        f.write(("\n".join(rows) + "\n").encode("ascii"))
    print("wrote", len(rows), "cases,", os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
