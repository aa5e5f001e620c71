"""Writes tests/golden/regex_dialect_cases.json: small patterns on which java.util.regex and PCRE2 differ by
design (see tests/test_oracle_pcre2_diff.py). `java` is the answer that follows from the node semantics of the
JDK's java/util/regex/Pattern.java (reasoning in `why`); the script refuses to write a case whose `java` answer
the oracle does not reproduce or whose `pcre2` answer libpcre2-8 does not give. No JVM exists in this image, so
`java` is derived, not observed.   Run from the repo root:  python tests/golden/make_regex_dialect_cases.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import pcre2  # noqa: E402
from oracle import oracle as O  # noqa: E402

CASES = [
    # ---- captures survive an abandoned independent sub-match; groups are cleared once per find()
    ("backref", r"(?:(?=(a))b|\1)", "a", (0, 1), None,
     "Pos look-ahead body ends in `accept`: GroupTail sets group 1 and never restores it; after `b` fails the second alternative's \\1 still sees 'a'. PCRE2 restores the capture when it backtracks out of the first alternative."),
    ("backref", r"(?:(?>(a))b|\1)", "a", (0, 1), None,
     "Atomic group: its body is matched against `accept`, so group 1 stays set after `b` fails."),
    ("backref", r"(?:(a)++b|\1)", "a", (0, 1), None,
     "Possessive Curly over a group: the atom (GroupHead..GroupTail) is matched in isolation per iteration; Curly returns false without touching groups."),
    ("backref", r"(?=(a))b|c\1", "aca", (1, 3), None,
     "Matcher.search clears groups once; Start tries position 0 (look-ahead sets group 1, `b` fails), then position 1 where `c\\1` uses the capture left over from position 0."),
    ("backref", r"(?!(a)b)\1", "ac", None, None,
     "Control case: the look-ahead body fails at `b`, GroupTail's continuation failed, so the capture IS restored: both engines agree that \\1 is unset."),
    # ---- possessive quantifier on a group: atomic per iteration (Java) vs. as a whole (PCRE2)
    ("poss-group", r"(\W{1,3}){2}+", ":.", None, (0, 2),
     "Curly(POSSESSIVE): iteration 1 takes ':.' and is final; iteration 2 finds nothing. PCRE2 backtracks inside iteration 1."),
    ("poss-group", r"(a{1,2}){2}+b", "aab", None, (0, 3),
     "Iteration 1 keeps 'aa'; iteration 2 finds no 'a'. PCRE2 gives one 'a' back inside the still-open atomic group."),
    ("poss-group", r"(a|ab){2}+c", "aabc", None, None,
     "Control case: once both iterations have matched 'a','a' neither engine reopens them for 'ab'."),
    # ---- a group iteration that consumed nothing ends the loop, even below the minimum count
    ("empty-iter", r"(\.{2,}|\D|^){2,}", "x1", (0, 0), (0, 1),
     "Loop.match: `if (i > locals[beginIndex])` is false after the empty iteration via ^, so the loop exits to `next` with count 1 < cmin."),
    ("empty-iter", r"(a|^){2}b", "ab", None, (0, 2),
     "Java never runs a second iteration after the empty one (and the first alternative 'a' leaves no 'a' for iteration 2); PCRE2 matches ^ then a."),
    ("empty-iter", r"(?:x|\b){2}y", " y", (1, 2), (1, 2),
     "Control case where both agree although the group can match empty: Java leaves the loop after the empty \\b iteration, PCRE2 runs a second empty one."),
]


def main():
    O.lib()
    out = []
    for cls, pat, text, java, pc, why in CASES:
        got = O.Regex(pat).span(text)
        got = None if got is None else tuple(got)
        assert got == java, (pat, text, "oracle", got, "expected java", java)
        pgot = pcre2.Regex(pat).search(text)
        assert pgot == pc, (pat, text, "pcre2", pgot, "expected", pc)
        out.append({"class": cls, "pattern": pat, "text": text, "java": java, "pcre2": pc, "why": why})
    with open(os.path.join(HERE, "regex_dialect_cases.json"), "w") as f:
        json.dump({"pcre2_version": pcre2.version(), "jdk": "java.util.regex of JDK 21 (pom.xml:10), derived from its source, not run",
                   "cases": out}, f, indent=1)
    print("wrote", len(out), "cases")


if __name__ == "__main__":
    main()
