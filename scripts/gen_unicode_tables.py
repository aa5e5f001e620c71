#!/usr/bin/env python
"""Generates the Unicode 15.0 property tables behind \\p{..} / (?u) / (?U) of the java.util.regex dialect:
general categories, scripts, blocks, binary properties, the java.lang.Character predicates and the simple case
mappings. Source of the data: perl's Unicode::UCD (perl 5.38 carries Unicode 15.0.0, the version of Java 21's
java.lang.Character). Written twice, identically: log_parser_b200/csrc/unicode_tables.inc (product host compiler)
and oracle/unicode_tables.inc (the checker keeps its own copy).      Run:  python scripts/gen_unicode_tables.py
"""
import json
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PERL = r'''
use strict; use warnings;
use Unicode::UCD qw(prop_invlist charscripts charblocks charinfo prop_values);
use JSON::PP;
my %out; $out{version} = Unicode::UCD::UnicodeVersion();
sub inv { my @l = prop_invlist($_[0]); return \@l; }
my @gc = qw(L LC Lu Ll Lt Lm Lo M Mn Mc Me N Nd Nl No P Pc Pd Ps Pe Pi Pf Po S Sm Sc Sk So Z Zs Zl Zp C Cc Cf Cs Co Cn);
$out{gc}{$_} = inv("General_Category=$_") for @gc;
my @bin = qw(Alphabetic Ideographic Lowercase Uppercase White_Space Hex_Digit Join_Control Noncharacter_Code_Point Bidi_Mirrored
             Emoji Emoji_Presentation Emoji_Modifier Emoji_Modifier_Base Emoji_Component Extended_Pictographic ID_Start ID_Continue);
$out{bin}{$_} = inv($_) for @bin;
my $s = charscripts(); for my $name (keys %$s) { my @r; push @r, [$_->[0], $_->[1]] for @{$s->{$name}}; $out{script}{$name} = \@r; }
my $b = charblocks(); for my $name (keys %$b) { my @r; push @r, [$_->[0], $_->[1]] for @{$b->{$name}}; $out{block}{$name} = \@r; }
# ISO 15924 aliases of the scripts (Character.UnicodeScript.forName accepts them)
for my $v (prop_values("Script")) { my @a = Unicode::UCD::prop_value_aliases("Script", $v); $out{script_alias}{$a[0]} = $a[1] if @a > 1; }
# simple case mappings (UnicodeData.txt fields 12-14), only where they differ from the code point itself
my @cased = prop_invlist("Changes_When_Casemapped");
for (my $i = 0; $i < @cased; $i += 2) {
    my $hi = $i + 1 < @cased ? $cased[$i + 1] : 0x110000;
    for my $cp ($cased[$i] .. $hi - 1) {
        my $ci = charinfo($cp) or next;
        my ($u, $l) = ($ci->{upper}, $ci->{lower});
        push @{$out{case}}, [$cp, $u eq "" ? $cp : hex($u), $l eq "" ? $cp : hex($l)] if $u ne "" || $l ne "";
    }
}
print JSON::PP->new->canonical->encode(\%out);
'''


def norm(name: str) -> str:
    return re.sub(r"[ _\-]", "", name).upper()


def inv_to_ranges(inv):
    out = []
    for i in range(0, len(inv), 2):
        lo = int(inv[i])
        hi = (int(inv[i + 1]) if i + 1 < len(inv) else 0x110000) - 1
        out.append((lo, hi))
    return out


def union(*sets):
    pts = sorted(r for s in sets for r in s)
    out = []
    for lo, hi in pts:
        if out and lo <= out[-1][1] + 1:
            out[-1] = (out[-1][0], max(out[-1][1], hi))
        else:
            out.append((lo, hi))
    return out


def minus(a, pts):
    out = []
    for lo, hi in a:
        cur = lo
        for p in sorted(pts):
            if cur <= p <= hi:
                if p > cur:
                    out.append((cur, p - 1))
                cur = p + 1
        if cur <= hi:
            out.append((cur, hi))
    return out


def main():
    d = json.loads(subprocess.run(["perl", "-e", PERL], capture_output=True, text=True, check=True).stdout)
    gc = {k: inv_to_ranges(v) for k, v in d["gc"].items()}
    bn = {k: inv_to_ranges(v) for k, v in d["bin"].items()}
    groups = {"gc": {}, "script": {}, "block": {}, "prop": {}, "java": {}}
    for k, v in gc.items():
        groups["gc"][k.upper()] = v
    for k, v in d["script"].items():
        groups["script"][norm(k)] = [(int(r[0]), int(r[1])) for r in v]
    for a0, a1 in d["script_alias"].items():
        n0, n1 = norm(a0), norm(a1)
        if n1 in groups["script"] and n0 not in groups["script"]:
            groups["script"][n0] = groups["script"][n1]
        elif n0 in groups["script"] and n1 not in groups["script"]:
            groups["script"][n1] = groups["script"][n0]
    for k, v in d["block"].items():
        groups["block"][norm(k)] = [(int(r[0]), int(r[1])) for r in v]
    # java.lang.Character.UnicodeBlock identifiers that differ from the normalised canonical names
    for java_name, canon in {"GREEK": "GREEKANDCOPTIC", "CYRILLICSUPPLEMENTARY": "CYRILLICSUPPLEMENT", "COMBININGMARKSFORSYMBOLS": "COMBININGDIACRITICALMARKSFORSYMBOLS"}.items():
        if canon in groups["block"]:
            groups["block"][java_name] = groups["block"][canon]
    L, Nd = gc["L"], gc["Nd"]
    marks = union(gc["Mn"], gc["Mc"], gc["Me"])
    P = groups["prop"]      # CharPredicates.forUnicodeProperty
    P["ALPHABETIC"] = bn["Alphabetic"]; P["IDEOGRAPHIC"] = bn["Ideographic"]; P["LETTER"] = L
    P["LOWERCASE"] = bn["Lowercase"]; P["UPPERCASE"] = bn["Uppercase"]; P["TITLECASE"] = gc["Lt"]
    P["WHITESPACE"] = bn["White_Space"]; P["CONTROL"] = gc["Cc"]; P["PUNCTUATION"] = gc["P"]
    P["HEXDIGIT"] = bn["Hex_Digit"]; P["ASSIGNED"] = _complement(gc["Cn"])
    P["NONCHARACTERCODEPOINT"] = bn["Noncharacter_Code_Point"]; P["DIGIT"] = Nd
    P["ALNUM"] = union(bn["Alphabetic"], Nd); P["BLANK"] = union(gc["Zs"], [(9, 9)])
    P["GRAPH"] = _complement(union(gc["Zs"], gc["Zl"], gc["Zp"], gc["Cc"], gc["Cs"], gc["Cn"]))
    P["PRINT"] = _minus_set(union(P["GRAPH"], P["BLANK"]), gc["Cc"])
    P["WORD"] = union(bn["Alphabetic"], marks, Nd, gc["Pc"], bn["Join_Control"]); P["JOINCONTROL"] = bn["Join_Control"]
    P["EMOJI"] = bn["Emoji"]; P["EMOJIPRESENTATION"] = bn["Emoji_Presentation"]; P["EMOJIMODIFIER"] = bn["Emoji_Modifier"]
    P["EMOJIMODIFIERBASE"] = bn["Emoji_Modifier_Base"]; P["EMOJICOMPONENT"] = bn["Emoji_Component"]; P["EXTENDEDPICTOGRAPHIC"] = bn["Extended_Pictographic"]
    J = groups["java"]      # java.lang.Character.isXxx
    J["JAVALOWERCASE"] = bn["Lowercase"]; J["JAVAUPPERCASE"] = bn["Uppercase"]; J["JAVATITLECASE"] = gc["Lt"]
    J["JAVAALPHABETIC"] = bn["Alphabetic"]; J["JAVAIDEOGRAPHIC"] = bn["Ideographic"]; J["JAVADIGIT"] = Nd; J["JAVALETTER"] = L
    J["JAVALETTERORDIGIT"] = union(L, Nd); J["JAVADEFINED"] = _complement(gc["Cn"]); J["JAVAISOCONTROL"] = [(0, 0x1F), (0x7F, 0x9F)]
    J["JAVASPACECHAR"] = union(gc["Zs"], gc["Zl"], gc["Zp"])
    J["JAVAWHITESPACE"] = union(minus(J["JAVASPACECHAR"], [0xA0, 0x2007, 0x202F]), [(9, 0xD), (0x1C, 0x1F)])
    J["JAVAMIRRORED"] = bn["Bidi_Mirrored"]
    ignorable = union([(0, 8), (0xE, 0x1B), (0x7F, 0x9F)], gc["Cf"])
    J["JAVAIDENTIFIERIGNORABLE"] = ignorable
    J["JAVAJAVAIDENTIFIERSTART"] = union(L, gc["Nl"], gc["Sc"], gc["Pc"])
    J["JAVAJAVAIDENTIFIERPART"] = union(J["JAVAJAVAIDENTIFIERSTART"], Nd, gc["Mn"], gc["Mc"], ignorable)
    J["JAVAUNICODEIDENTIFIERSTART"] = bn["ID_Start"]
    J["JAVAUNICODEIDENTIFIERPART"] = union(bn["ID_Continue"], ignorable)
    # case folding of UNICODE_CASE: fold(c) = toLowerCase(toUpperCase(c)) with the simple mappings; classes of equal fold
    up, lo = {}, {}
    for cp, u, l in d["case"]:
        up[int(cp)], lo[int(cp)] = int(u), int(l)
    fold = lambda c: lo.get(up.get(c, c), up.get(c, c))
    classes = {}
    for cp in set(up) | set(lo) | set(up.values()) | set(lo.values()):
        classes.setdefault(fold(cp), set()).add(cp)
    pairs = sorted((c, o) for members in classes.values() if len(members) > 1 for c in members for o in members if o != c)

    lines = ["// GENERATED by scripts/gen_unicode_tables.py from perl's Unicode::UCD (Unicode %s, the version of Java 21). Do not edit." % d["version"],
             "// Code-point ranges behind \\p{..} of java.util.regex.Pattern (general categories, scripts, blocks, binary properties,",
             "// java.lang.Character predicates) and the simple case-folding classes of UNICODE_CASE. Names are upper-cased with",
             "// spaces, '_' and '-' removed.", "struct UniRange { unsigned lo, hi; };",
             "struct UniProp { const char* name; const UniRange* r; unsigned n; };"]
    tables = []
    for grp, props in groups.items():
        seen = {}
        for name in sorted(props):
            rs = props[name]
            key = tuple(rs)
            if key not in seen:
                ident = "UR_%s_%d" % (grp, len(seen))
                seen[key] = ident
                body = ",".join("{0x%X,0x%X}" % r for r in rs) or "{1,0}"
                lines.append("static const UniRange %s[] = {%s};" % (ident, body))
            tables.append((grp, name, seen[key], len(rs)))
    for grp in groups:
        lines.append("static const UniProp UNI_%s[] = {" % grp.upper())
        lines += ['    {"%s", %s, %d},' % (n, ident, cnt) for g, n, ident, cnt in tables if g == grp]
        lines.append("    {nullptr, nullptr, 0}};")
    lines.append("// (c, other): `other` matches `c` under CASE_INSENSITIVE | UNICODE_CASE; sorted by c")
    lines.append("static const unsigned UNI_CASE_PAIRS[][2] = {" + ",".join("{0x%X,0x%X}" % p for p in pairs) + "};")
    lines.append("static const unsigned UNI_CASE_PAIRS_N = %d;" % len(pairs))
    text = "\n".join(lines) + "\n"
    for path in ("log_parser_b200/csrc/unicode_tables.inc", "oracle/unicode_tables.inc"):
        open(os.path.join(ROOT, path), "w").write(text)
    print("unicode", d["version"], {g: len(p) for g, p in groups.items()}, "case pairs", len(pairs), "bytes", len(text))


def _complement(rs):
    out, cur = [], 0
    for lo, hi in sorted(rs):
        if lo > cur:
            out.append((cur, lo - 1))
        cur = hi + 1
    if cur <= 0x10FFFF:
        out.append((cur, 0x10FFFF))
    return out


def _minus_set(a, b):
    return _complement(union(_complement(a), b))


if __name__ == "__main__":
    main()
