// java.util.regex (JDK 21) dialect -> AST for the device compilers.
// Stands in for Pattern.compile at AnalysisService.java:64,70,80 and
// ContextAnalysisService.java:27-34 (flags: none, or CASE_INSENSITIVE for ERROR/WARN).
// Every consuming atom becomes a code-point set so that the NFA builder, the literal
// extractor and the bytecode emitter all see one representation.
#include <algorithm>
#include <climits>
#include <map>

#include "rx.h"

namespace lpx {

void cpset_normalize(CpSet& s) {
    std::sort(s.begin(), s.end(), [](const CpRange& a, const CpRange& b) { return a.lo < b.lo; });
    CpSet out;
    for (const auto& r : s) {
        if (r.lo > r.hi) continue;
        if (!out.empty() && (uint64_t)r.lo <= (uint64_t)out.back().hi + 1) out.back().hi = std::max(out.back().hi, r.hi);
        else out.push_back(r);
    }
    s.swap(out);
}
CpSet cpset_negate(const CpSet& in) {
    CpSet s = in;
    cpset_normalize(s);
    CpSet out;
    uint32_t next = 0;
    for (const auto& r : s) {
        if (r.lo > next) out.push_back({next, r.lo - 1});
        next = r.hi + 1;
    }
    if (next <= 0x10FFFF) out.push_back({next, 0x10FFFF});
    return out;
}
CpSet cpset_intersect(const CpSet& a0, const CpSet& b0) {
    CpSet a = a0, b = b0, out;
    cpset_normalize(a);
    cpset_normalize(b);
    size_t i = 0, j = 0;
    while (i < a.size() && j < b.size()) {
        uint32_t lo = std::max(a[i].lo, b[j].lo), hi = std::min(a[i].hi, b[j].hi);
        if (lo <= hi) out.push_back({lo, hi});
        if (a[i].hi < b[j].hi) ++i; else ++j;
    }
    return out;
}
bool cpset_has(const CpSet& s, uint32_t c) {
    for (const auto& r : s) if (c >= r.lo && c <= r.hi) return true;
    return false;
}

namespace {

inline bool is_alpha(uint32_t c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z'); }
inline bool is_digit(uint32_t c) { return c >= '0' && c <= '9'; }
inline uint32_t flip_case(uint32_t c) { return is_alpha(c) ? (c ^ 0x20) : c; }

// ASCII case closure of literal members of a class (Pattern.range / caseInsensitiveRangeFor).
void close_case(CpSet& s) {
    cpset_normalize(s);
    CpSet extra;
    for (uint32_t c = 'A'; c <= 'z'; ++c) {
        if (!is_alpha(c)) continue;
        if (!cpset_has(s, c) && cpset_has(s, flip_case(c))) extra.push_back({c, c});
    }
    s.insert(s.end(), extra.begin(), extra.end());
    cpset_normalize(s);
}

#include "unicode_tables.inc"

// names are compared upper-cased with ' ', '_' and '-' removed (UnicodeScript.forName / UnicodeBlock.forName are lenient)
std::string uni_norm(const std::string& name) {
    std::string o;
    for (unsigned char ch : name) {
        if (ch == ' ' || ch == '_' || ch == '-') continue;
        o.push_back((char)((ch >= 'a' && ch <= 'z') ? ch - 32 : ch));
    }
    return o;
}
bool uni_find(const UniProp* tab, const std::string& norm, CpSet& out) {
    for (; tab->name; ++tab) {
        if (norm != tab->name) continue;
        out.clear();
        for (unsigned i = 0; i < tab->n; ++i) out.push_back({tab->r[i].lo, tab->r[i].hi});
        return true;
    }
    return false;
}
CpSet uni_get(const UniProp* tab, const char* norm) {
    CpSet s;
    if (!uni_find(tab, norm, s)) throw std::logic_error(std::string("missing Unicode table ") + norm);
    return s;
}
// UNICODE_CASE closure: every code point whose simple case folding equals that of a member joins the set
void close_case_unicode(CpSet& s) {
    cpset_normalize(s);
    CpSet extra;
    for (unsigned i = 0; i < UNI_CASE_PAIRS_N; ++i)
        if (cpset_has(s, UNI_CASE_PAIRS[i][0]) && !cpset_has(s, UNI_CASE_PAIRS[i][1])) extra.push_back({UNI_CASE_PAIRS[i][1], UNI_CASE_PAIRS[i][1]});
    s.insert(s.end(), extra.begin(), extra.end());
    cpset_normalize(s);
}

std::vector<uint32_t> utf8_to_cps(const std::string& in) {
    std::vector<uint32_t> out;
    const unsigned char* s = reinterpret_cast<const unsigned char*>(in.data());
    size_t n = in.size(), i = 0;
    while (i < n) {
        uint32_t c = s[i];
        int need = c < 0x80 ? 0 : (c & 0xE0) == 0xC0 ? 1 : (c & 0xF0) == 0xE0 ? 2 : (c & 0xF8) == 0xF0 ? 3 : -1;
        if (need == 0) { out.push_back(c); ++i; continue; }
        if (need < 0 || i + (size_t)need >= n) throw RegexSyntaxError("pattern is not valid UTF-8", (int)i);
        uint32_t cp = need == 1 ? (c & 0x1F) : need == 2 ? (c & 0x0F) : (c & 0x07);
        for (int k = 1; k <= need; ++k) {
            if ((s[i + k] & 0xC0) != 0x80) throw RegexSyntaxError("pattern is not valid UTF-8", (int)i);
            cp = (cp << 6) | (s[i + k] & 0x3F);
        }
        out.push_back(cp);
        i += need + 1;
    }
    return out;
}

class P {
public:
    std::vector<uint32_t> src;
    size_t at = 0;
    int fl = 0;
    int groups = 1;
    bool regular = true;
    std::map<std::string, int> named;

    [[noreturn]] void bad(const std::string& m) const { throw RegexSyntaxError(m, (int)at); }
    [[noreturn]] void nope(const std::string& m) const { throw RegexUnsupported(m); }
    bool on(int f) const { return (fl & f) != 0; }
    static constexpr uint32_t END = 0xFFFFFFFFu;

    void ws() {
        if (!on(RXF_COMMENTS)) return;
        while (at < src.size()) {
            uint32_t c = src[at];
            if (c == ' ' || (c >= 9 && c <= 13)) { ++at; continue; }
            if (c == '#') { while (at < src.size() && src[at] != '\n' && src[at] != '\r') ++at; continue; }
            break;
        }
    }
    uint32_t look() { ws(); return at < src.size() ? src[at] : END; }
    uint32_t raw(size_t k = 0) const { return at + k < src.size() ? src[at + k] : END; }
    uint32_t take() { ws(); return at < src.size() ? src[at++] : END; }

    // (?U) UNICODE_CHARACTER_CLASS implies UNICODE_CASE (Pattern.java: "flags |= UNICODE_CASE" wherever it is switched on)
    void flags_ok() { if (on(RXF_UNICODE_CLASS)) fl |= RXF_UNICODE_CASE; }
    bool ucase() const { return on(RXF_ICASE) && on(RXF_UNICODE_CASE); }
    void fold(CpSet& s) const { if (ucase()) close_case_unicode(s); else if (on(RXF_ICASE)) close_case(s); }

    // \Q..\E -> backslash-escaped literals, as Pattern.RemoveQEQuoting does before parsing
    void unquote() {
        bool any = false;
        for (size_t i = 0; i + 1 < src.size(); ++i) {
            if (src[i] == '\\') { if (src[i + 1] == 'Q') { any = true; break; } ++i; }
        }
        if (!any) return;
        std::vector<uint32_t> out;
        bool q = false;
        for (size_t i = 0; i < src.size();) {
            uint32_t c = src[i];
            if (!q) {
                if (c == '\\' && i + 1 < src.size()) {
                    if (src[i + 1] == 'Q') { q = true; i += 2; continue; }
                    out.push_back(c); out.push_back(src[i + 1]); i += 2; continue;
                }
                out.push_back(c); ++i;
            } else {
                if (c == '\\' && i + 1 < src.size() && src[i + 1] == 'E') { q = false; i += 2; continue; }
                if (c < 128 && !is_alpha(c) && !is_digit(c)) out.push_back('\\');
                out.push_back(c); ++i;
            }
        }
        src.swap(out);
    }

    static AstP node(AstKind k) { AstP a(new Ast()); a->kind = k; return a; }
    AstP set_node(CpSet s) { cpset_normalize(s); AstP a = node(AstKind::Set); a->set = std::move(s); return a; }
    AstP literal(uint32_t c) {
        CpSet s{{c, c}};
        if (ucase()) close_case_unicode(s);
        else if (on(RXF_ICASE) && is_alpha(c)) s.push_back({flip_case(c), flip_case(c)});
        return set_node(std::move(s));
    }
    AstP assertion(AssertKind k) { AstP a = node(AstKind::Assert); a->assert_kind = k; return a; }
    // (?U)\b / (?U)\B: Pattern.Bound with UNICODE_CHARACTER_CLASS asks CharPredicates.WORD() of the code points on both
    // sides (a non-spacing mark is a word character there whatever precedes it). The automata keep one "previous byte
    // is an ASCII word byte" bit, so this form goes to the backtracking engine, written with what it already has:
    //   \b = (?<=\w)(?!\w) | (?<!\w)(?=\w)      \B = (?<=\w)(?=\w) | (?<!\w)(?!\w)
    AstP unicode_word_boundary(bool negated) {
        regular = false;
        auto side = [&](bool behind, bool neg) {
            AstP l = node(AstKind::Look); l->behind = behind; l->neg = neg;
            l->kids.push_back(set_node(uni_get(UNI_PROP, "WORD")));
            return l;
        };
        AstP alt = node(AstKind::Alt);
        for (int after_word = 1; after_word >= 0; --after_word) {
            AstP cat = node(AstKind::Cat);
            cat->kids.push_back(side(true, !after_word));
            cat->kids.push_back(side(false, negated ? !after_word : (bool)after_word));
            alt->kids.push_back(std::move(cat));
        }
        return alt;
    }

    // -------- expression / sequence
    AstP expr() {
        std::vector<AstP> branches;
        branches.push_back(sequence());
        while (look() == '|') { take(); branches.push_back(sequence()); }
        if (branches.size() == 1) return std::move(branches.front());
        AstP a = node(AstKind::Alt);
        a->kids = std::move(branches);
        return a;
    }
    AstP sequence() {
        AstP cat = node(AstKind::Cat);
        for (;;) {
            uint32_t c = look();
            if (c == END || c == '|' || c == ')') break;
            bool was_flag_group = false;
            AstP item = atom(was_flag_group);
            if (was_flag_group) continue;
            cat->kids.push_back(quantified(std::move(item)));
        }
        if (cat->kids.empty()) return node(AstKind::Empty);
        if (cat->kids.size() == 1) return std::move(cat->kids.front());
        return cat;
    }

    long number() {
        long v = 0;
        while (is_digit(raw())) { v = v * 10 + (src[at++] - '0'); if (v > INT_MAX) bad("Illegal repetition range"); }
        return v;
    }
    AstP quantified(AstP item) {
        uint32_t c = look();
        int lo, hi;
        if (c == '?') { take(); lo = 0; hi = 1; }
        else if (c == '*') { take(); lo = 0; hi = -1; }
        else if (c == '+') { take(); lo = 1; hi = -1; }
        else if (c == '{') {
            take();
            if (!is_digit(raw())) bad("Illegal repetition");
            lo = (int)number(); hi = lo;
            if (raw() == ',') {
                ++at; hi = -1;
                if (raw() != '}') { if (!is_digit(raw())) bad("Unclosed counted closure"); hi = (int)number(); }
            }
            if (raw() != '}') bad("Unclosed counted closure");
            ++at;
            if (hi >= 0 && hi < lo) bad("Illegal repetition range");
        } else return item;
        AstP r = node(AstKind::Repeat);
        r->min = lo; r->max = hi;
        uint32_t m = look();
        if (m == '?') { take(); r->mode = 1; }
        else if (m == '+') { take(); r->mode = 2; regular = false; }
        r->kids.push_back(std::move(item));
        return r;
    }

    // -------- groups
    int inline_flags(int cur) {
        bool adding = true;
        for (;;) {
            int bit;
            switch (raw()) {
                case 'i': bit = RXF_ICASE; break;
                case 'd': bit = RXF_UNIX_LINES; break;
                case 'm': bit = RXF_MULTILINE; break;
                case 's': bit = RXF_DOTALL; break;
                case 'u': bit = RXF_UNICODE_CASE; break;
                case 'x': bit = RXF_COMMENTS; break;
                case 'U': bit = RXF_UNICODE_CLASS; break;
                case 'c': nope("CANON_EQ");
                case '-': if (!adding) bad("Unknown inline modifier"); adding = false; ++at; continue;
                default: return cur;
            }
            cur = adding ? (cur | bit) : (cur & ~bit);
            ++at;
        }
    }
    AstP group(bool& was_flag_group) {
        const int outer = fl;
        AstP res;
        if (raw() != '?') {
            AstP g = node(AstKind::Group);
            g->group = groups++;
            g->kids.push_back(expr());
            res = std::move(g);
        } else {
            ++at;
            uint32_t c = raw();
            if (c == ':') { ++at; res = expr(); }
            else if (c == '=' || c == '!') {
                ++at; regular = false;
                AstP l = node(AstKind::Look); l->neg = c == '!';
                l->kids.push_back(expr()); res = std::move(l);
            } else if (c == '>') {
                ++at; regular = false;
                AstP a = node(AstKind::Atomic); a->kids.push_back(expr()); res = std::move(a);
            } else if (c == '<' && (raw(1) == '=' || raw(1) == '!')) {
                bool neg = raw(1) == '!';
                at += 2; regular = false;
                AstP l = node(AstKind::Look); l->neg = neg; l->behind = true;
                l->kids.push_back(expr()); res = std::move(l);
            } else if (c == '<') {
                ++at;
                if (!is_alpha(raw())) bad("capturing group name does not start with a Latin letter");
                std::string name;
                while (raw() != '>') {
                    uint32_t ch = raw();
                    if (ch == END || !(is_alpha(ch) || is_digit(ch))) bad("named capturing group is missing trailing '>'");
                    name.push_back((char)ch); ++at;
                }
                ++at;
                if (named.count(name)) bad("Named capturing group <" + name + "> is already defined");
                AstP g = node(AstKind::Group);
                g->group = groups++;
                named[name] = g->group;
                g->kids.push_back(expr());
                res = std::move(g);
            } else {
                int nf = inline_flags(fl);
                if (raw() == ')') { ++at; fl = nf; flags_ok(); was_flag_group = true; return nullptr; }
                if (raw() != ':') bad("Unknown inline modifier");
                ++at; fl = nf; flags_ok();
                res = expr();
            }
        }
        if (look() != ')') bad("Unclosed group");
        take();
        fl = outer;
        return res;
    }

    // -------- escapes
    static CpSet digits() { return {{'0', '9'}}; }
    static CpSet word() { return {{'0', '9'}, {'A', 'Z'}, {'_', '_'}, {'a', 'z'}}; }
    static CpSet space() { return {{9, 13}, {' ', ' '}}; }
    static CpSet hspace() {
        return {{9, 9}, {' ', ' '}, {0xA0, 0xA0}, {0x1680, 0x1680}, {0x180E, 0x180E}, {0x2000, 0x200A},
                {0x202F, 0x202F}, {0x205F, 0x205F}, {0x3000, 0x3000}};
    }
    static CpSet vspace() { return {{10, 13}, {0x85, 0x85}, {0x2028, 0x2029}}; }

    bool shorthand(uint32_t c, CpSet& out) {
        const bool U = on(RXF_UNICODE_CLASS);      // (?U): \d = Nd, \w = Alphabetic|M|Nd|Pc|Join_Control, \s = White_Space
        switch (c) {
            case 'd': out = U ? uni_get(UNI_PROP, "DIGIT") : digits(); return true;
            case 'D': out = cpset_negate(U ? uni_get(UNI_PROP, "DIGIT") : digits()); return true;
            case 'w': out = U ? uni_get(UNI_PROP, "WORD") : word(); return true;
            case 'W': out = cpset_negate(U ? uni_get(UNI_PROP, "WORD") : word()); return true;
            case 's': out = U ? uni_get(UNI_PROP, "WHITESPACE") : space(); return true;
            case 'S': out = cpset_negate(U ? uni_get(UNI_PROP, "WHITESPACE") : space()); return true;
            case 'h': out = hspace(); return true;
            case 'H': out = cpset_negate(hspace()); return true;
            case 'v': out = vspace(); return true;
            case 'V': out = cpset_negate(vspace()); return true;
            default: return false;
        }
    }
    CpSet posix(bool negate) {
        std::string name;
        if (raw() == '{') {
            ++at;
            while (raw() != '}') { if (raw() == END) bad("Unclosed character family"); name.push_back((char)src[at++]); }
            ++at;
        } else {
            if (raw() == END) bad("Illegal character family");
            name.push_back((char)src[at++]);
        }
        CpSet s = family(name);
        cpset_normalize(s);
        return negate ? cpset_negate(s) : s;
    }
    // Pattern.family(): \\p{InBlock}, \\p{IsScript|IsCategory|IsBinaryProperty}, \\p{sc=..|blk=..|gc=..}, general categories,
    // POSIX names (ASCII, or their Unicode versions under (?U)), java.lang.Character predicates
    CpSet family(const std::string& name) {
        const bool ci = on(RXF_ICASE);
        CpSet s;
        const size_t eq = name.find('=');
        if (eq != std::string::npos) {
            const std::string key = uni_norm(name.substr(0, eq)), value = name.substr(eq + 1);
            if (key == "SC" || key == "SCRIPT") { if (!uni_find(UNI_SCRIPT, uni_norm(value), s)) bad("Unknown character script name {" + value + "}"); }
            else if (key == "BLK" || key == "BLOCK") { if (!uni_find(UNI_BLOCK, uni_norm(value), s)) bad("Unknown character block name {" + value + "}"); }
            else if (key == "GC" || key == "GENERALCATEGORY") { if (!property(value, ci, s)) bad("Unknown Unicode property {name=<" + name.substr(0, eq) + ">, value=<" + value + ">}"); }
            else bad("Unknown Unicode property {name=<" + name.substr(0, eq) + ">, value=<" + value + ">}");
            return s;
        }
        if (name.rfind("In", 0) == 0) {
            if (!uni_find(UNI_BLOCK, uni_norm(name.substr(2)), s)) bad("Unknown character block name {" + name.substr(2) + "}");
            return s;
        }
        if (name.rfind("Is", 0) == 0) {
            const std::string sh = name.substr(2);
            if (unicode_property(sh, ci, s) || property(sh, ci, s) || uni_find(UNI_SCRIPT, uni_norm(sh), s)) return s;
            bad("Unknown character property name {" + name + "}");
        }
        if (on(RXF_UNICODE_CLASS) && posix_unicode(name, ci, s)) return s;
        if (property(name, ci, s)) return s;
        bad("Unknown character property name {" + name + "}");
    }
    // CharPredicates.forUnicodeProperty: binary properties, name matched without regard to case
    bool unicode_property(const std::string& name, bool ci, CpSet& s) {
        std::string n;
        for (unsigned char ch : name) n.push_back((char)((ch >= 'a' && ch <= 'z') ? ch - 32 : ch));
        static const char* const known[] = {"ALPHABETIC", "ASSIGNED", "CONTROL", "EMOJI", "EMOJI_PRESENTATION", "EMOJI_MODIFIER", "EMOJI_MODIFIER_BASE",
                                            "EMOJI_COMPONENT", "EXTENDED_PICTOGRAPHIC", "HEXDIGIT", "HEX_DIGIT", "IDEOGRAPHIC", "JOINCONTROL", "JOIN_CONTROL",
                                            "LETTER", "LOWERCASE", "NONCHARACTERCODEPOINT", "NONCHARACTER_CODE_POINT", "TITLECASE", "PUNCTUATION", "UPPERCASE",
                                            "WHITESPACE", "WHITE_SPACE", "WORD", "DIGIT", "ALNUM", "BLANK", "GRAPH", "PRINT"};
        bool ok = false;
        for (const char* k : known) ok = ok || n == k;
        if (!ok) return false;
        if (ci && (n == "LOWERCASE" || n == "UPPERCASE" || n == "TITLECASE")) {       // Pattern: case-insensitive -> any cased letter
            s = uni_get(UNI_PROP, "LOWERCASE");
            CpSet u = uni_get(UNI_PROP, "UPPERCASE"), t = uni_get(UNI_PROP, "TITLECASE");
            s.insert(s.end(), u.begin(), u.end());
            s.insert(s.end(), t.begin(), t.end());
            return true;
        }
        return uni_find(UNI_PROP, uni_norm(n), s);
    }
    // CharPredicates.forPOSIXName under UNICODE_CHARACTER_CLASS
    bool posix_unicode(const std::string& name, bool ci, CpSet& s) {
        static const char* const map[][2] = {{"ALPHA", "ALPHABETIC"}, {"LOWER", "LOWERCASE"}, {"UPPER", "UPPERCASE"}, {"SPACE", "WHITESPACE"},
                                             {"PUNCT", "PUNCTUATION"}, {"XDIGIT", "HEXDIGIT"}, {"ALNUM", "ALNUM"}, {"CNTRL", "CONTROL"},
                                             {"DIGIT", "DIGIT"}, {"BLANK", "BLANK"}, {"GRAPH", "GRAPH"}, {"PRINT", "PRINT"}};
        std::string n;
        for (unsigned char ch : name) n.push_back((char)((ch >= 'a' && ch <= 'z') ? ch - 32 : ch));
        for (const auto& m : map) if (n == m[0]) return unicode_property(m[1], ci, s);
        return false;
    }
    // CharPredicates.forProperty: general categories, POSIX names (ASCII), L1/all/LC/LD, javaXxx; names are case-sensitive
    bool property(const std::string& name, bool ci, CpSet& s) {
        static const CpSet lower{{'a', 'z'}}, upper{{'A', 'Z'}}, alpha{{'A', 'Z'}, {'a', 'z'}};
        static const CpSet punct{{0x21, 0x2F}, {0x3A, 0x40}, {0x5B, 0x60}, {0x7B, 0x7E}};
        auto add = [&](const CpSet& o) { s.insert(s.end(), o.begin(), o.end()); };
        s.clear();
        if (name == "Lower") add(ci ? alpha : lower);
        else if (name == "Upper") add(ci ? alpha : upper);
        else if (name == "ASCII") s = {{0, 0x7F}};
        else if (name == "Alpha") add(alpha);
        else if (name == "Digit") add(digits());
        else if (name == "Alnum") { add(alpha); add(digits()); }
        else if (name == "Punct") add(punct);
        else if (name == "Graph") { add(alpha); add(digits()); add(punct); }
        else if (name == "Print") { add(alpha); add(digits()); add(punct); s.push_back({' ', ' '}); }
        else if (name == "Blank") s = {{9, 9}, {' ', ' '}};
        else if (name == "Cntrl") s = {{0, 0x1F}, {0x7F, 0x7F}};
        else if (name == "XDigit") s = {{'0', '9'}, {'A', 'F'}, {'a', 'f'}};
        else if (name == "Space") add(space());
        else if (name == "L1") s = {{0, 0xFF}};
        else if (name == "all") s = {{0, 0x10FFFF}};
        else if (name == "LD") { add(uni_get(UNI_GC, "L")); add(uni_get(UNI_GC, "ND")); }
        else if (ci && (name == "Lu" || name == "Ll" || name == "Lt")) { add(uni_get(UNI_GC, "LU")); add(uni_get(UNI_GC, "LL")); add(uni_get(UNI_GC, "LT")); }
        else if (ci && (name == "javaLowerCase" || name == "javaUpperCase" || name == "javaTitleCase")) {
            add(uni_get(UNI_JAVA, "JAVALOWERCASE")); add(uni_get(UNI_JAVA, "JAVAUPPERCASE")); add(uni_get(UNI_JAVA, "JAVATITLECASE"));
        } else if (name.rfind("java", 0) == 0) return uni_find(UNI_JAVA, uni_norm(name), s);
        else {
            // general categories: exactly "L", "Lu", "LC", ... (a lower-case or longer spelling is an error in Java)
            const bool shape = (name.size() == 1 && name[0] >= 'A' && name[0] <= 'Z') ||
                               (name.size() == 2 && name[0] >= 'A' && name[0] <= 'Z' && ((name[1] >= 'a' && name[1] <= 'z') || name == "LC"));
            if (!shape) return false;
            return uni_find(UNI_GC, uni_norm(name), s);
        }
        return true;
    }
    static int hex(uint32_t c) {
        if (is_digit(c)) return (int)c - '0';
        if (c >= 'a' && c <= 'f') return (int)c - 'a' + 10;
        if (c >= 'A' && c <= 'F') return (int)c - 'A' + 10;
        return -1;
    }
    // escapes that stand for exactly one code point; `c` = char after the backslash (consumed)
    bool one_cp_escape(uint32_t c, uint32_t& out) {
        switch (c) {
            case 'a': out = 7; return true;
            case 'e': out = 27; return true;
            case 'f': out = 12; return true;
            case 'n': out = 10; return true;
            case 'r': out = 13; return true;
            case 't': out = 9; return true;
            case 'c':
                if (raw() == END) bad("Illegal control escape sequence");
                out = src[at++] ^ 64; return true;
            case '0': {
                uint32_t d1 = raw();
                if (!(d1 >= '0' && d1 <= '7')) bad("Illegal octal escape sequence");
                ++at; out = d1 - '0';
                uint32_t d2 = raw();
                if (d2 >= '0' && d2 <= '7') {
                    ++at; out = out * 8 + (d2 - '0');
                    uint32_t d3 = raw();
                    if (d3 >= '0' && d3 <= '7' && d1 <= '3') { ++at; out = out * 8 + (d3 - '0'); }
                }
                return true;
            }
            case 'x': {
                if (raw() == '{') {
                    ++at; uint64_t v = 0; int n = 0;
                    while (raw() != '}') {
                        int h = raw() == END ? -1 : hex(raw());
                        if (h < 0) bad(n ? "Unclosed hexadecimal escape sequence" : "Illegal hexadecimal escape sequence");
                        v = v * 16 + h; ++n; ++at;
                        if (v > 0x10FFFF) bad("Hexadecimal codepoint is too big");
                    }
                    if (!n) bad("Illegal hexadecimal escape sequence");
                    ++at; out = (uint32_t)v; return true;
                }
                int a = raw() == END ? -1 : hex(raw()), b = raw(1) == END ? -1 : hex(raw(1));
                if (a < 0 || b < 0) bad("Illegal hexadecimal escape sequence");
                at += 2; out = a * 16 + b; return true;
            }
            case 'u': {
                auto four = [&](size_t off, uint32_t& v) {
                    v = 0;
                    for (size_t k = 0; k < 4; ++k) { int h = raw(off + k) == END ? -1 : hex(raw(off + k)); if (h < 0) return false; v = v * 16 + h; }
                    return true;
                };
                uint32_t hi;
                if (!four(0, hi)) bad("Illegal Unicode escape sequence");
                at += 4;
                if (hi >= 0xD800 && hi <= 0xDBFF && raw() == '\\' && raw(1) == 'u') {
                    uint32_t lo;
                    if (four(2, lo) && lo >= 0xDC00 && lo <= 0xDFFF) { at += 6; out = 0x10000 + ((hi - 0xD800) << 10) + (lo - 0xDC00); return true; }
                }
                if (hi >= 0xD800 && hi <= 0xDFFF) nope("unpaired surrogate escape");
                out = hi; return true;
            }
            default: return false;
        }
    }

    AstP escape_atom() {
        if (at >= src.size()) bad("Unexpected internal error");
        uint32_t c = src[at++];
        CpSet cs;
        uint32_t cp;
        if (shorthand(c, cs)) return set_node(std::move(cs));
        if (one_cp_escape(c, cp)) return literal(cp);
        switch (c) {
            case 'p': return set_node(posix(false));
            case 'P': return set_node(posix(true));
            case 'b':
                if (raw() == '{') nope("\\b{g}");
                if (on(RXF_UNICODE_CLASS)) return unicode_word_boundary(false);
                return assertion(AssertKind::WordB);
            case 'B':
                if (on(RXF_UNICODE_CLASS)) return unicode_word_boundary(true);
                return assertion(AssertKind::NotWordB);
            case 'A': return assertion(AssertKind::Begin);
            case 'z': return assertion(AssertKind::End);
            case 'Z': return assertion(on(RXF_UNIX_LINES) ? AssertKind::End : AssertKind::Eol);
            case 'G': return assertion(AssertKind::Begin);     // end of the previous match: for the one find() per line, position 0
            case 'X': nope("\\X");
            case 'N': nope("\\N{name}");
            case 'R': {  // \r\n | [\n\x0B\f\r\x85  ]
                AstP alt = node(AstKind::Alt);
                AstP crlf = node(AstKind::Cat);
                crlf->kids.push_back(set_node({{13, 13}}));
                crlf->kids.push_back(set_node({{10, 10}}));
                alt->kids.push_back(std::move(crlf));
                alt->kids.push_back(set_node(vspace()));
                return alt;
            }
            case 'k': {
                if (raw() != '<') bad("\\k is not followed by '<' for named capturing group");
                ++at;
                std::string name;
                while (raw() != '>') { if (raw() == END) bad("named capturing group is missing trailing '>'"); name.push_back((char)src[at++]); }
                ++at;
                auto it = named.find(name);
                if (it == named.end()) bad("named capturing group <" + name + "> does not exist");
                regular = false;
                if (ucase()) nope("back-reference under (?iu)");
                AstP b = node(AstKind::Backref); b->group = it->second; b->icase = on(RXF_ICASE); return b;
            }
            default: break;
        }
        if (c >= '1' && c <= '9') {
            int ref = (int)c - '0';
            if (ref >= groups) bad("Illegal/unsupported escape sequence");
            while (is_digit(raw())) {
                int wider = ref * 10 + (int)(raw() - '0');
                if (groups - 1 < wider) break;
                ref = wider; ++at;
            }
            regular = false;
            if (ucase()) nope("back-reference under (?iu)");
            AstP b = node(AstKind::Backref); b->group = ref; b->icase = on(RXF_ICASE); return b;
        }
        if (is_alpha(c)) bad("Illegal/unsupported escape sequence");
        return literal(c);
    }

    // -------- character classes
    // one member: a single code point (returns true) or a whole set (false)
    bool member(uint32_t& cp, CpSet& cs) {
        uint32_t c = take();
        if (c == END) bad("Unclosed character class");
        if (c != '\\') { cp = c; return true; }
        if (at >= src.size()) bad("Unclosed character class");
        uint32_t e = src[at++];
        if (shorthand(e, cs)) return false;
        if (one_cp_escape(e, cp)) return true;
        if (e == 'p') { cs = posix(false); return false; }
        if (e == 'P') { cs = posix(true); return false; }
        if (is_alpha(e) || (e >= '1' && e <= '9')) bad("Illegal/unsupported escape sequence");
        cp = e; return true;
    }
    bool at_and() { return look() == '&' && raw(1) == '&'; }
    CpSet members(bool& seen) {
        CpSet plain, sets;
        for (;;) {
            uint32_t c = look();
            if (c == END) bad("Unclosed character class");
            if ((c == ']' && seen) || at_and()) break;
            if (c == '[') { take(); CpSet inner = bracket(); sets.insert(sets.end(), inner.begin(), inner.end()); seen = true; continue; }
            uint32_t a; CpSet cs;
            if (!member(a, cs)) { sets.insert(sets.end(), cs.begin(), cs.end()); seen = true; continue; }
            seen = true;
            if (look() == '-' && raw(1) != ']' && raw(1) != END && raw(1) != '[') {
                take();
                uint32_t b; CpSet cs2;
                if (!member(b, cs2)) bad("Illegal character range");      // Pattern.range(): a class escape as the end point gives -1 < ch
                if (b < a) bad("Illegal character range");
                plain.push_back({a, b});
            } else plain.push_back({a, a});
        }
        fold(plain);
        plain.insert(plain.end(), sets.begin(), sets.end());
        cpset_normalize(plain);
        return plain;
    }
    CpSet bracket() {  // after '[' ; eats the closing ']'
        bool negated = false;
        if (look() == '^') { take(); negated = true; }
        bool seen = false;
        CpSet acc = members(seen);
        while (at_and()) {
            at += 2;
            CpSet rhs; bool have = false;
            while (look() != ']' && !at_and()) {
                if (look() == END) bad("Unclosed character class");
                CpSet part;
                if (look() == '[') { take(); part = bracket(); }
                else { bool s2 = true; part = members(s2); }
                rhs = have ? cpset_intersect(rhs, part) : part;
                have = true;
            }
            if (!have) { if (!seen) bad("Bad class syntax"); continue; }
            acc = seen ? cpset_intersect(acc, rhs) : rhs;
            seen = true;
        }
        if (look() != ']') bad("Unclosed character class");
        take();
        return negated ? cpset_negate(acc) : acc;
    }

    AstP atom(bool& was_flag_group) {
        uint32_t c = take();
        switch (c) {
            case '(': return group(was_flag_group);
            case '[': return set_node(bracket());
            case '.': {
                if (on(RXF_DOTALL)) return set_node({{0, 0x10FFFF}});
                if (on(RXF_UNIX_LINES)) return set_node(cpset_negate({{10, 10}}));
                return set_node(cpset_negate({{10, 10}, {13, 13}, {0x85, 0x85}, {0x2028, 0x2029}}));
            }
            case '^': {
                if (!on(RXF_MULTILINE)) return assertion(AssertKind::Begin);
                regular = false;
                AstP a = assertion(AssertKind::BolMulti); a->unix_lines = on(RXF_UNIX_LINES); return a;
            }
            case '$': {
                if (on(RXF_MULTILINE)) { regular = false; AstP a = assertion(AssertKind::EolMulti); a->unix_lines = on(RXF_UNIX_LINES); return a; }
                return assertion(on(RXF_UNIX_LINES) ? AssertKind::End : AssertKind::Eol);
            }
            case '\\': return escape_atom();
            case '*': case '+': case '?': --at; bad(std::string("Dangling meta character '") + (char)c + "'");
            case '{': --at; return node(AstKind::Empty);
            default: return literal(c);
        }
    }
};

void lengths(Ast* a) {
    for (auto& k : a->kids) lengths(k.get());
    auto sat = [](long v) { return (int)std::min<long>(v, INT_MAX / 4); };
    switch (a->kind) {
        case AstKind::Set: a->min_len = a->max_len = 1; break;
        case AstKind::Group: case AstKind::Atomic: a->min_len = a->kids[0]->min_len; a->max_len = a->kids[0]->max_len; break;
        case AstKind::Alt: {
            int lo = INT_MAX, hi = 0;
            for (auto& k : a->kids) { lo = std::min(lo, k->min_len); hi = (hi < 0 || k->max_len < 0) ? -1 : std::max(hi, k->max_len); }
            a->min_len = lo; a->max_len = hi; break;
        }
        case AstKind::Cat: {
            long lo = 0, hi = 0;
            for (auto& k : a->kids) { lo += k->min_len; hi = (hi < 0 || k->max_len < 0) ? -1 : hi + k->max_len; }
            a->min_len = sat(lo); a->max_len = hi < 0 ? -1 : (hi > INT_MAX / 4 ? -1 : (int)hi); break;
        }
        case AstKind::Repeat: {
            const Ast* k = a->kids[0].get();
            a->min_len = sat((long)k->min_len * a->min);
            if (a->max < 0) a->max_len = k->max_len == 0 ? 0 : -1;
            else if (k->max_len < 0) a->max_len = -1;
            else { long hi = (long)k->max_len * a->max; a->max_len = hi > INT_MAX / 4 ? -1 : (int)hi; }
            break;
        }
        case AstKind::Backref: a->min_len = 0; a->max_len = -1; break;
        default: a->min_len = a->max_len = 0; break;
    }
}
void lookbehind_bounded(const Ast* a) {
    if (a->kind == AstKind::Look && a->behind && a->kids[0]->max_len < 0)
        throw RegexSyntaxError("Look-behind group does not have an obvious maximum length", 0);
    for (auto& k : a->kids) lookbehind_bounded(k.get());
}

bool has_assert(const Ast* a) {
    if (a->kind == AstKind::Assert) return true;
    for (auto& k : a->kids) if (has_assert(k.get())) return true;
    return false;
}
// Loop.match sends an iteration that consumed nothing straight on to what follows the loop, even
// when iterations are still owed. When the body can only match empty where an assertion holds
// (e.g. (?:^x*){2}a), that prunes matches a pure regular-language reading would find, so such
// loops are decided by the backtracking verifier, which follows the same rule.
bool has_assertion_guarded_empty_loop(const Ast* a) {
    if (a->kind == AstKind::Repeat && a->kids[0]->kind != AstKind::Set && (a->max < 0 || a->max > 1) &&
        a->kids[0]->min_len == 0 && has_assert(a->kids[0].get()))
        return true;
    for (auto& k : a->kids) if (has_assertion_guarded_empty_loop(k.get())) return true;
    return false;
}

}  // namespace

ParsedRegex parse_java_regex(const std::string& utf8, int flags) {
    P p;
    p.src = utf8_to_cps(utf8);
    p.fl = flags;
    p.flags_ok();
    p.unquote();
    AstP root = p.expr();
    if (p.at < p.src.size()) {
        if (p.src[p.at] == ')') throw RegexSyntaxError("Unmatched closing ')'", (int)p.at);
        throw RegexSyntaxError("Unexpected internal error", (int)p.at);
    }
    lengths(root.get());
    lookbehind_bounded(root.get());
    ParsedRegex out;
    out.root = std::move(root);
    out.n_groups = p.groups;
    out.regular = p.regular && !has_assertion_guarded_empty_loop(out.root.get());
    return out;
}

}  // namespace lpx
