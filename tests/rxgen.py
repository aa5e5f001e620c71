"""Seeded random generator of Java-dialect regexes (plus their CPython `re` translation) and texts.

Used by the differential tests of the oracle (vs CPython re / regex) and of the GPU compiler (vs the
oracle). The generated subset is the one all three engines share after translation:
literals, escapes, classes, `.`, anchors, \\b, groups, alternation, greedy/lazy/possessive
quantifiers, atomic groups, lookahead, fixed-width lookbehind, back-references, inline (?i).
"""
from __future__ import annotations

import random
import re

ALPHA = "abcERx_1 .-:\r"
JAVA_DOT = "[^\\n\\r\\x85\\u2028\\u2029]"
PY_EOL = "(?=[\\r\\x85\\u2028\\u2029]?\\Z)"


class Gen:
    def __init__(self, rng: random.Random, regular_only: bool = False, max_depth: int = 3,
                 no_backrefs: bool = False, no_group_possessive: bool = False):
        self.rng = rng
        self.regular_only = regular_only
        self.no_backrefs = no_backrefs                      # the PCRE2 differential keeps capture-state quirks out
        self.no_group_possessive = no_group_possessive      # Java: X{n,m}+ on a group is atomic per iteration, PCRE2: as a whole
        self.max_depth = max_depth
        self.ngroups = 0
        self.open_groups = []

    # each generator returns (java, python, fixed_width or None)
    def lit(self):
        c = self.rng.choice("abcERx_1 .-:")
        j = re.escape(c) if not c.isalnum() and c not in " _" else c
        if c == " ":
            j = " "
        return j, j, 1

    def cls(self):
        r = self.rng.random()
        if r < 0.25:
            k = self.rng.choice(["\\d", "\\w", "\\s", "\\D", "\\W", "\\S"])
            return k, k, 1
        if r < 0.4:
            return ".", JAVA_DOT, 1
        items = []
        for _ in range(self.rng.randint(1, 3)):
            t = self.rng.random()
            if t < 0.5:
                items.append(self.rng.choice("abcERx_1"))
            elif t < 0.8:
                items.append(self.rng.choice(["a-c", "A-Z", "0-9", "a-z"]))
            else:
                items.append(self.rng.choice(["\\d", "\\w", "\\s", "\\.", "\\-"]))
        neg = "^" if self.rng.random() < 0.25 else ""
        s = "[" + neg + "".join(items) + "]"
        return s, s, 1

    def anchor(self):
        k = self.rng.choice(["^", "$", "\\b", "\\B", "\\A", "\\z", "\\Z"])
        py = {"$": PY_EOL, "\\Z": PY_EOL, "\\z": "\\Z", "\\A": "\\A"}.get(k, k)
        return k, py, 0

    def atom(self, depth):
        r = self.rng.random()
        if depth >= self.max_depth or r < 0.45:
            return self.lit()
        if r < 0.6:
            return self.cls()
        if r < 0.68:
            return self.anchor()
        if r < 0.8:
            # capturing group
            self.ngroups += 1
            g = self.ngroups
            self.open_groups.append(g)
            j, p, w = self.alt(depth + 1)
            self.open_groups.pop()
            self.closed = getattr(self, "closed", []) + [g]
            return "(" + j + ")", "(" + p + ")", w
        if r < 0.86:
            j, p, w = self.alt(depth + 1)
            if self.rng.random() < 0.3:
                return "(?i:" + j + ")", "(?i:" + p + ")", w
            return "(?:" + j + ")", "(?:" + p + ")", w
        if self.regular_only:
            return self.lit()
        if r < 0.9:
            j, p, w = self.alt(depth + 1)
            return "(?>" + j + ")", "(?>" + p + ")", w
        if r < 0.94:
            j, p, _ = self.alt(depth + 1)
            k = self.rng.choice(["=", "!"])
            return "(?" + k + j + ")", "(?" + k + p + ")", 0
        if r < 0.97:
            # fixed-width lookbehind (CPython requires it)
            saved = (self.ngroups, list(getattr(self, "closed", [])))
            j, p, w = self.alt(depth + 1, fixed=True)
            if w is None:
                self.ngroups, self.closed = saved
                return self.lit()
            k = self.rng.choice(["=", "!"])
            return "(?<" + k + j + ")", "(?<" + k + p + ")", 0
        closed = [g for g in getattr(self, "closed", []) if g not in self.open_groups]
        if closed and not self.no_backrefs:
            g = self.rng.choice(closed)
            return "\\%d" % g, "(?:\\%d)" % g, None
        return self.lit()

    def quant(self, depth, fixed=False):
        j, p, w = self.atom(depth)
        r = self.rng.random()
        if r < 0.6 or (fixed and w is None):
            return j, p, w
        if fixed:
            n = self.rng.randint(1, 3)
            return j + "{%d}" % n, p + "{%d}" % n, (w * n if w is not None else None)
        q = self.rng.choice(["*", "+", "?", "{2}", "{1,3}", "{2,}", "{0,2}"])
        m = self.rng.random()
        suffix = ""
        if m < 0.2:
            suffix = "?"
        elif m < 0.3 and not self.regular_only and not (self.no_group_possessive and j.startswith("(")):
            suffix = "+"
        return j + q + suffix, p + q + suffix, None

    def seq(self, depth, fixed=False):
        n = self.rng.randint(1, 4 if depth == 0 else 3)
        js, ps, w = [], [], 0
        for _ in range(n):
            j, p, wi = self.quant(depth, fixed)
            js.append(j)
            ps.append(p)
            w = None if (w is None or wi is None) else w + wi
        return "".join(js), "".join(ps), w

    def alt(self, depth, fixed=False):
        n = 1 if self.rng.random() < 0.7 else self.rng.randint(2, 3)
        parts = [self.seq(depth, fixed) for _ in range(n)]
        ws = {p[2] for p in parts}
        w = parts[0][2] if len(ws) == 1 else None
        if fixed and w is None and n > 1:
            # keep lookbehind alternatives the same width: retry as a single sequence
            parts = parts[:1]
            w = parts[0][2]
        return "|".join(p[0] for p in parts), "|".join(p[1] for p in parts), w

    def regex(self):
        self.ngroups = 0
        self.open_groups = []
        self.closed = []
        j, p, _ = self.alt(0)
        if self.rng.random() < 0.1:
            j, p = "(?i)" + j, "(?i)" + p
        return j, p


def gen_text(rng: random.Random, maxlen: int = 24) -> str:
    n = rng.randint(0, maxlen)
    return "".join(rng.choice(ALPHA) for _ in range(n))
