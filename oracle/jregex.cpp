// ORACLE — TEST INFRASTRUCTURE ONLY (see jregex.hpp for scope, citations and parity status).
//
// Restates java.util.regex.Pattern (JDK 21) parsing and Matcher.find() as used at
// AnalysisService.java:64,70,80,93-95; ScoringService.java:281,300,330;
// ContextAnalysisService.java:27-34,64-79. Algorithm source: the published behaviour of
// java.util.regex.Pattern (JDK 21 javadoc + OpenJDK Pattern.java structure: Begin/Caret/
// Dollar/Bound/Curly/Loop/GroupCurly nodes). Not copied from the reference (it contains none).
#include "jregex.hpp"

#include <algorithm>
#include <climits>
#include <map>
#include <stdexcept>

namespace orx {

// ------------------------------------------------------------------ CharSet
void CharSet::add(int lo, int hi) {
    if (lo > hi) return;
    r.push_back({lo, hi});
}
void CharSet::add(const CharSet& o) {
    for (auto& x : o.r) r.push_back(x);
}
void CharSet::normalize() {
    std::sort(r.begin(), r.end(), [](const Range& a, const Range& b) { return a.lo < b.lo; });
    std::vector<Range> out;
    for (auto& x : r) {
        if (!out.empty() && x.lo <= out.back().hi + 1) out.back().hi = std::max(out.back().hi, x.hi);
        else out.push_back(x);
    }
    r.swap(out);
    ascii[0] = ascii[1] = 0;
    for (auto& x : r)
        for (int c = x.lo; c <= x.hi && c < 128; ++c) ascii[c >> 6] |= 1ull << (c & 63);
}
void CharSet::negate() {
    normalize();
    std::vector<Range> out;
    int prev = 0;
    for (auto& x : r) {
        if (x.lo > prev) out.push_back({prev, x.lo - 1});
        prev = x.hi + 1;
    }
    if (prev <= 0x10FFFF) out.push_back({prev, 0x10FFFF});
    r.swap(out);
    normalize();
}
void CharSet::intersect(const CharSet& o) {
    CharSet a = *this, b = o;
    a.normalize();
    b.normalize();
    std::vector<Range> out;
    size_t i = 0, j = 0;
    while (i < a.r.size() && j < b.r.size()) {
        int lo = std::max(a.r[i].lo, b.r[j].lo), hi = std::min(a.r[i].hi, b.r[j].hi);
        if (lo <= hi) out.push_back({lo, hi});
        if (a.r[i].hi < b.r[j].hi) ++i; else ++j;
    }
    r.swap(out);
    normalize();
}
static inline int ascii_lower(int c) { return (c >= 'A' && c <= 'Z') ? c + 32 : c; }
static inline int ascii_upper(int c) { return (c >= 'a' && c <= 'z') ? c - 32 : c; }
void CharSet::close_ascii_case() {
    normalize();
    std::vector<Range> extra;
    for (int c = 0; c < 128; ++c) {
        if (has(c)) continue;
        if (has(ascii_upper(c)) || has(ascii_lower(c))) extra.push_back({c, c});
    }
    for (auto& x : extra) r.push_back(x);
    normalize();
}
bool CharSet::has(int c) const {
    if (c >= 0 && c < 128) return (ascii[c >> 6] >> (c & 63)) & 1;
    size_t lo = 0, hi = r.size();
    while (lo < hi) {
        size_t mid = (lo + hi) / 2;
        if (r[mid].hi < c) lo = mid + 1; else hi = mid;
    }
    return lo < r.size() && r[lo].lo <= c;
}

void decode_utf8(const uint8_t* s, size_t n, std::vector<int>& out) {
    out.clear();
    out.reserve(n);
    size_t i = 0;
    while (i < n) {
        unsigned c = s[i];
        if (c < 0x80) { out.push_back((int)c); ++i; continue; }
        int need = 0; unsigned cp = 0, minv = 0;
        if ((c & 0xE0) == 0xC0) { need = 1; cp = c & 0x1F; minv = 0x80; }
        else if ((c & 0xF0) == 0xE0) { need = 2; cp = c & 0x0F; minv = 0x800; }
        else if ((c & 0xF8) == 0xF0) { need = 3; cp = c & 0x07; minv = 0x10000; }
        else { out.push_back(0xFFFD); ++i; continue; }
        bool ok = i + (size_t)need < n;
        if (!ok) { out.push_back(0xFFFD); ++i; continue; }
        for (int k = 1; k <= need; ++k) {
            unsigned cc = s[i + k];
            if ((cc & 0xC0) != 0x80) { ok = false; break; }
            cp = (cp << 6) | (cc & 0x3F);
        }
        if (!ok || cp < minv || cp > 0x10FFFF || (cp >= 0xD800 && cp <= 0xDFFF)) {
            out.push_back(0xFFFD); ++i; continue;
        }
        out.push_back((int)cp);
        i += need + 1;
    }
}

static void decode_pattern(const std::string& p, std::vector<int>& out) {
    decode_utf8(reinterpret_cast<const uint8_t*>(p.data()), p.size(), out);
}

// ------------------------------------------------------------------ Parser
namespace {

#include "unicode_tables.inc"

// java.lang.Character.UnicodeScript.forName / UnicodeBlock.forName ignore case; blocks also accept the name with its
// spaces removed: compare upper-cased, without ' ', '_' and '-'
static std::string squash(const std::string& n) {
    std::string o;
    for (unsigned char ch : n) if (ch != ' ' && ch != '_' && ch != '-') o.push_back((char)toupper(ch));
    return o;
}
static bool table_lookup(const UniProp* t, const std::string& squashed, CharSet& out) {
    for (; t->name; ++t)
        if (squashed == t->name) {
            out = CharSet();
            for (unsigned i = 0; i < t->n; ++i) out.add((int)t->r[i].lo, (int)t->r[i].hi);
            return true;
        }
    return false;
}
static CharSet table_get(const UniProp* t, const char* name) {
    CharSet s;
    if (!table_lookup(t, name, s)) throw std::logic_error(std::string("oracle: missing table ") + name);
    return s;
}
// UNICODE_CASE: ch matches c when toLowerCase(toUpperCase(ch)) == toLowerCase(toUpperCase(c)) (Pattern.SingleU); the table
// lists, per code point, the others of its class
static void close_unicode_case(CharSet& s) {
    s.normalize();
    CharSet extra;
    for (unsigned i = 0; i < UNI_CASE_PAIRS_N; ++i)
        if (s.has((int)UNI_CASE_PAIRS[i][0]) && !s.has((int)UNI_CASE_PAIRS[i][1])) extra.add((int)UNI_CASE_PAIRS[i][1], (int)UNI_CASE_PAIRS[i][1]);
    s.add(extra);
    s.normalize();
}

struct Parser {
    std::vector<int> p;
    int pos = 0;
    int flags = 0;
    int ngroups = 1;
    std::map<std::string, int> names;

    [[noreturn]] void err(const std::string& m) { throw SyntaxError(m, pos); }
    [[noreturn]] void unsup(const std::string& m) { throw Unsupported(m, pos); }
    bool has(int f) const { return (flags & f) != 0; }
    int len() const { return (int)p.size(); }

    void skip_comments() {
        if (!has(F_COMMENTS)) return;
        for (;;) {
            while (pos < len() && (p[pos] == ' ' || (p[pos] >= 9 && p[pos] <= 13))) ++pos;
            if (pos < len() && p[pos] == '#') {
                while (pos < len() && p[pos] != '\n' && p[pos] != '\r') ++pos;
                continue;
            }
            break;
        }
    }
    int peek() { skip_comments(); return pos < len() ? p[pos] : -1; }
    int raw_peek(int off = 0) const { return pos + off < len() ? p[pos + off] : -1; }
    int next() { skip_comments(); return pos < len() ? p[pos++] : -1; }

    // Pattern.java switches UNICODE_CASE on together with UNICODE_CHARACTER_CLASS
    void check_flags() { if (has(F_UNICODE_CLASS)) flags |= F_UNICODE_CASE; }
    bool unicode_ci() const { return has(F_CASE_INSENSITIVE) && has(F_UNICODE_CASE); }

    // RemoveQEQuoting: \Q...\E becomes escaped literals (Pattern.RemoveQEQuoting).
    void remove_qe() {
        std::vector<int> out;
        int n = len();
        int i = 0;
        bool found = false;
        while (i < n) {
            if (p[i] == '\\' && i + 1 < n && p[i + 1] == 'Q') { found = true; break; }
            if (p[i] == '\\' && i + 1 < n) { i += 2; continue; }
            ++i;
        }
        if (!found) return;
        i = 0;
        bool inq = false;
        while (i < n) {
            int c = p[i];
            if (!inq) {
                if (c == '\\' && i + 1 < n && p[i + 1] == 'Q') { inq = true; i += 2; continue; }
                if (c == '\\' && i + 1 < n) { out.push_back(c); out.push_back(p[i + 1]); i += 2; continue; }
                out.push_back(c); ++i;
            } else {
                if (c == '\\' && i + 1 < n && p[i + 1] == 'E') { inq = false; i += 2; continue; }
                bool alnum = (c < 128) && ((c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z'));
                if (c < 128 && !alnum) out.push_back('\\');
                out.push_back(c); ++i;
            }
        }
        p.swap(out);
    }

    NodeP mk(Kind k) { NodeP n(new Node()); n->kind = k; return n; }
    // Pattern.Bound: isWord() is the ASCII word class, or CharPredicates.WORD() under UNICODE_CHARACTER_CLASS (the node
    // then carries that set; a non-spacing mark belongs to it, whatever precedes the mark).
    NodeP bound(Kind k) {
        NodeP n = mk(k);
        if (has(F_UNICODE_CLASS)) { n->flags = F_UNICODE_CLASS; n->set = table_get(UNI_PROP, "WORD"); n->set.normalize(); }
        return n;
    }

    NodeP parse_alt() {
        std::vector<NodeP> alts;
        alts.push_back(parse_seq());
        while (peek() == '|') { next(); alts.push_back(parse_seq()); }
        if (alts.size() == 1) return std::move(alts[0]);
        NodeP n = mk(N_ALT);
        n->kids = std::move(alts);
        return n;
    }

    NodeP parse_seq() {
        NodeP cat = mk(N_CAT);
        for (;;) {
            int ch = peek();
            if (ch == -1 || ch == '|' || ch == ')') break;
            bool directive = false;
            NodeP a = parse_atom(directive);
            if (directive) continue;
            a = parse_closure(std::move(a));
            cat->kids.push_back(std::move(a));
        }
        if (cat->kids.empty()) return mk(N_EMPTY);
        if (cat->kids.size() == 1) return std::move(cat->kids[0]);
        return cat;
    }

    NodeP lit(int c) {
        if (unicode_ci()) {       // SingleU: any code point of the same case-folding class
            CharSet s; s.add(c, c);
            close_unicode_case(s);
            if (s.r.size() != 1 || s.r[0].lo != s.r[0].hi) return setnode(std::move(s));
        }
        NodeP n = mk(N_CHAR);
        n->c = c;
        n->icase = has(F_CASE_INSENSITIVE) && c < 128 && ascii_lower(c) != ascii_upper(c);
        return n;
    }
    NodeP setnode(CharSet s) {
        s.normalize();
        NodeP n = mk(N_SET);
        n->set = std::move(s);
        return n;
    }

    NodeP parse_closure(NodeP atom) {
        int ch = peek();
        int mn, mx;
        if (ch == '?') { next(); mn = 0; mx = 1; }
        else if (ch == '*') { next(); mn = 0; mx = -1; }
        else if (ch == '+') { next(); mn = 1; mx = -1; }
        else if (ch == '{') {
            next();
            int d = raw_peek();
            if (!(d >= '0' && d <= '9')) err("Illegal repetition");
            long v = 0;
            while (raw_peek() >= '0' && raw_peek() <= '9') { v = v * 10 + (p[pos++] - '0'); if (v > INT_MAX) err("Illegal repetition range"); }
            mn = (int)v; mx = mn;
            if (raw_peek() == ',') {
                ++pos; mx = -1;
                if (raw_peek() != '}') {
                    long w = 0;
                    if (!(raw_peek() >= '0' && raw_peek() <= '9')) err("Unclosed counted closure");
                    while (raw_peek() >= '0' && raw_peek() <= '9') { w = w * 10 + (p[pos++] - '0'); if (w > INT_MAX) err("Illegal repetition range"); }
                    mx = (int)w;
                }
            }
            if (raw_peek() != '}') err("Unclosed counted closure");
            ++pos;
            if (mx >= 0 && mx < mn) err("Illegal repetition range");
        } else return atom;
        int mode = 0;
        int c2 = peek();
        if (c2 == '?') { next(); mode = 1; }
        else if (c2 == '+') { next(); mode = 2; }
        NodeP r = mk(N_REPEAT);
        r->min = mn; r->max = mx; r->mode = mode;
        r->kids.push_back(std::move(atom));
        return r;
    }

    int parse_flag_letters(int cur, bool& saw_any) {
        bool on = true;
        for (;;) {
            int ch = raw_peek();
            int f = 0;
            switch (ch) {
                case 'i': f = F_CASE_INSENSITIVE; break;
                case 'd': f = F_UNIX_LINES; break;
                case 'm': f = F_MULTILINE; break;
                case 's': f = F_DOTALL; break;
                case 'u': f = F_UNICODE_CASE; break;
                case 'x': f = F_COMMENTS; break;
                case 'U': f = F_UNICODE_CLASS; break;
                case 'c': unsup("CANON_EQ flag");
                case '-': if (!on) err("Unknown inline modifier"); on = false; ++pos; continue;
                default: return cur;
            }
            saw_any = true;
            if (on) cur |= f; else cur &= ~f;
            ++pos;
        }
    }

    NodeP parse_group(bool& directive) {
        // '(' already consumed
        int saved = flags;
        NodeP out;
        if (raw_peek() == '?') {
            ++pos;
            int ch = raw_peek();
            if (ch == ':') { ++pos; out = parse_alt(); }
            else if (ch == '=' || ch == '!') {
                ++pos;
                NodeP l = mk(N_LOOK); l->neg = (ch == '!'); l->behind = false;
                l->kids.push_back(parse_alt()); out = std::move(l);
            } else if (ch == '>') {
                ++pos;
                NodeP a = mk(N_ATOMIC); a->kids.push_back(parse_alt()); out = std::move(a);
            } else if (ch == '<') {
                ++pos;
                int c2 = raw_peek();
                if (c2 == '=' || c2 == '!') {
                    ++pos;
                    NodeP l = mk(N_LOOK); l->neg = (c2 == '!'); l->behind = true;
                    l->kids.push_back(parse_alt()); out = std::move(l);
                } else {
                    std::string name;
                    if (!((c2 >= 'a' && c2 <= 'z') || (c2 >= 'A' && c2 <= 'Z'))) err("capturing group name does not start with a Latin letter");
                    while (raw_peek() != '>' && raw_peek() != -1) {
                        int c = p[pos++];
                        if (!((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9'))) err("named capturing group is missing trailing '>'");
                        name.push_back((char)c);
                    }
                    if (raw_peek() != '>') err("named capturing group is missing trailing '>'");
                    ++pos;
                    if (names.count(name)) err("Named capturing group <" + name + "> is already defined");
                    int g = ngroups++;
                    names[name] = g;
                    NodeP gn = mk(N_GROUP); gn->group = g;
                    gn->kids.push_back(parse_alt()); out = std::move(gn);
                }
            } else {
                bool saw = false;
                int nf = parse_flag_letters(flags, saw);
                int c3 = raw_peek();
                if (c3 == ')') {
                    ++pos;
                    flags = nf;  // applies to the rest of the enclosing group
                    check_flags();
                    directive = true;
                    return nullptr;
                } else if (c3 == ':') {
                    ++pos;
                    flags = nf;
                    check_flags();
                    out = parse_alt();
                } else err("Unknown inline modifier");
            }
        } else {
            int g = ngroups++;
            NodeP gn = mk(N_GROUP); gn->group = g;
            gn->kids.push_back(parse_alt()); out = std::move(gn);
        }
        if (peek() != ')') err("Unclosed group");
        next();
        flags = saved;
        return out;
    }

    static CharSet cls_digit() { CharSet s; s.add('0', '9'); return s; }
    static CharSet cls_word() { CharSet s; s.add('a', 'z'); s.add('A', 'Z'); s.add('0', '9'); s.add('_', '_'); return s; }
    static CharSet cls_space() { CharSet s; s.add(' ', ' '); s.add(9, 13); return s; }
    static CharSet cls_hspace() {
        CharSet s; s.add(' ', ' '); s.add(9, 9); s.add(0xA0, 0xA0); s.add(0x1680, 0x1680); s.add(0x180e, 0x180e);
        s.add(0x2000, 0x200a); s.add(0x202f, 0x202f); s.add(0x205f, 0x205f); s.add(0x3000, 0x3000); return s;
    }
    static CharSet cls_vspace() { CharSet s; s.add(10, 13); s.add(0x85, 0x85); s.add(0x2028, 0x2029); return s; }

    bool predefined(int ch, CharSet& out) {
        bool neg = false;
        const bool U = has(F_UNICODE_CLASS);      // CharPredicates.DIGIT() / WORD() / WHITE_SPACE() instead of the ASCII ones
        switch (ch) {
            case 'd': out = U ? table_get(UNI_PROP, "DIGIT") : cls_digit(); break;
            case 'D': out = U ? table_get(UNI_PROP, "DIGIT") : cls_digit(); neg = true; break;
            case 'w': out = U ? table_get(UNI_PROP, "WORD") : cls_word(); break;
            case 'W': out = U ? table_get(UNI_PROP, "WORD") : cls_word(); neg = true; break;
            case 's': out = U ? table_get(UNI_PROP, "WHITESPACE") : cls_space(); break;
            case 'S': out = U ? table_get(UNI_PROP, "WHITESPACE") : cls_space(); neg = true; break;
            case 'h': out = cls_hspace(); break;
            case 'H': out = cls_hspace(); neg = true; break;
            case 'v': out = cls_vspace(); break;
            case 'V': out = cls_vspace(); neg = true; break;
            default: return false;
        }
        if (neg) out.negate(); else out.normalize();
        return true;
    }

    CharSet property(bool neg) {
        // after \p or \P
        std::string name;
        if (raw_peek() == '{') {
            ++pos;
            while (raw_peek() != '}' && raw_peek() != -1) name.push_back((char)p[pos++]);
            if (raw_peek() != '}') err("Unclosed character family");
            ++pos;
        } else {
            if (raw_peek() == -1) err("Illegal character family");
            name.push_back((char)p[pos++]);
        }
        CharSet s;
        const bool ci = has(F_CASE_INSENSITIVE);
        const size_t eq = name.find('=');
        if (eq != std::string::npos) {                       // \\p{name=value}
            const std::string key = squash(name.substr(0, eq)), val = name.substr(eq + 1);
            bool ok = false;
            if (key == "SC" || key == "SCRIPT") { ok = table_lookup(UNI_SCRIPT, squash(val), s); if (!ok) err("Unknown character script name {" + val + "}"); }
            else if (key == "BLK" || key == "BLOCK") { ok = table_lookup(UNI_BLOCK, squash(val), s); if (!ok) err("Unknown character block name {" + val + "}"); }
            else if (key == "GC" || key == "GENERALCATEGORY") ok = for_property(val, ci, s);
            if (!ok) err("Unknown Unicode property {name=<" + name.substr(0, eq) + ">, value=<" + val + ">}");
        } else if (name.compare(0, 2, "In") == 0) {          // \\p{InBlockName}
            if (!table_lookup(UNI_BLOCK, squash(name.substr(2)), s)) err("Unknown character block name {" + name.substr(2) + "}");
        } else if (name.compare(0, 2, "Is") == 0) {          // \\p{IsGeneralCategory}, \\p{IsBinaryProperty}, \\p{IsScriptName}
            const std::string sh = name.substr(2);
            if (!for_unicode_property(sh, ci, s) && !for_property(sh, ci, s) && !table_lookup(UNI_SCRIPT, squash(sh), s))
                err("Unknown character property name {" + name + "}");
        } else {
            bool ok = has(F_UNICODE_CLASS) && for_posix_unicode(name, ci, s);
            if (!ok) ok = for_property(name, ci, s);
            if (!ok) err("Unknown character property name {" + name + "}");
        }
        if (neg) s.negate(); else s.normalize();
        return s;
    }
    static std::string upper_ascii(const std::string& n) { std::string o; for (unsigned char ch : n) o.push_back((char)toupper(ch)); return o; }
    // CharPredicates.forUnicodeProperty (the name is upper-cased first)
    bool for_unicode_property(const std::string& name, bool ci, CharSet& s) {
        const std::string n = upper_ascii(name);
        static const char* const names[] = {"ALPHABETIC", "LETTER", "IDEOGRAPHIC", "LOWERCASE", "UPPERCASE", "TITLECASE", "WHITE_SPACE", "WHITESPACE",
                                            "CONTROL", "PUNCTUATION", "HEX_DIGIT", "HEXDIGIT", "ASSIGNED", "NONCHARACTER_CODE_POINT", "NONCHARACTERCODEPOINT",
                                            "DIGIT", "ALNUM", "BLANK", "GRAPH", "PRINT", "WORD", "JOIN_CONTROL", "JOINCONTROL", "EMOJI", "EMOJI_PRESENTATION",
                                            "EMOJI_MODIFIER", "EMOJI_MODIFIER_BASE", "EMOJI_COMPONENT", "EXTENDED_PICTOGRAPHIC"};
        bool known = false;
        for (const char* k : names) if (n == k) known = true;
        if (!known) return false;
        if (ci && (n == "LOWERCASE" || n == "UPPERCASE" || n == "TITLECASE")) {
            s = table_get(UNI_PROP, "LOWERCASE"); s.add(table_get(UNI_PROP, "UPPERCASE")); s.add(table_get(UNI_PROP, "TITLECASE"));
            return true;
        }
        return table_lookup(UNI_PROP, squash(n), s);
    }
    // CharPredicates.forPOSIXName: the POSIX names under UNICODE_CHARACTER_CLASS
    bool for_posix_unicode(const std::string& name, bool ci, CharSet& s) {
        const std::string n = upper_ascii(name);
        const char* to = n == "ALPHA" ? "ALPHABETIC" : n == "LOWER" ? "LOWERCASE" : n == "UPPER" ? "UPPERCASE" : n == "SPACE" ? "WHITE_SPACE"
                       : n == "PUNCT" ? "PUNCTUATION" : n == "XDIGIT" ? "HEX_DIGIT" : n == "ALNUM" ? "ALNUM" : n == "CNTRL" ? "CONTROL"
                       : n == "DIGIT" ? "DIGIT" : n == "BLANK" ? "BLANK" : n == "GRAPH" ? "GRAPH" : n == "PRINT" ? "PRINT" : nullptr;
        return to && for_unicode_property(to, ci, s);
    }
    // CharPredicates.forProperty: the `props` map (general categories, POSIX ASCII classes, L1, all, LC, LD, javaXxx); exact names
    bool for_property(const std::string& name, bool ci, CharSet& s) {
        s = CharSet();
        auto alpha = [&]() { s.add('a', 'z'); s.add('A', 'Z'); };
        auto punct = [&]() { s.add(0x21, 0x2F); s.add(0x3A, 0x40); s.add(0x5B, 0x60); s.add(0x7B, 0x7E); };
        if (name == "Lower") { if (ci) alpha(); else s.add('a', 'z'); }
        else if (name == "Upper") { if (ci) alpha(); else s.add('A', 'Z'); }
        else if (name == "ASCII") s.add(0, 0x7F);
        else if (name == "Alpha") alpha();
        else if (name == "Digit") s.add('0', '9');
        else if (name == "Alnum") { alpha(); s.add('0', '9'); }
        else if (name == "Punct") punct();
        else if (name == "Graph") { alpha(); s.add('0', '9'); punct(); }
        else if (name == "Print") { alpha(); s.add('0', '9'); punct(); s.add(' ', ' '); }
        else if (name == "Blank") { s.add(' ', ' '); s.add(9, 9); }
        else if (name == "Cntrl") { s.add(0, 0x1F); s.add(0x7F, 0x7F); }
        else if (name == "XDigit") { s.add('0', '9'); s.add('a', 'f'); s.add('A', 'F'); }
        else if (name == "Space") { s.add(' ', ' '); s.add(9, 13); }
        else if (name == "L1") s.add(0, 0xFF);
        else if (name == "all") s.add(0, 0x10FFFF);
        else if (name == "LD") { s = table_get(UNI_GC, "L"); s.add(table_get(UNI_GC, "ND")); }
        else if (ci && (name == "Lu" || name == "Ll" || name == "Lt")) { s = table_get(UNI_GC, "LU"); s.add(table_get(UNI_GC, "LL")); s.add(table_get(UNI_GC, "LT")); }
        else if (ci && (name == "javaLowerCase" || name == "javaUpperCase" || name == "javaTitleCase")) {
            s = table_get(UNI_JAVA, "JAVALOWERCASE"); s.add(table_get(UNI_JAVA, "JAVAUPPERCASE")); s.add(table_get(UNI_JAVA, "JAVATITLECASE"));
        } else if (name.compare(0, 4, "java") == 0) return table_lookup(UNI_JAVA, squash(name), s);
        else {
            const bool cat = (name.size() == 1 && isupper((unsigned char)name[0])) ||
                             (name.size() == 2 && isupper((unsigned char)name[0]) && (islower((unsigned char)name[1]) || name == "LC"));
            return cat && table_lookup(UNI_GC, squash(name), s);
        }
        return true;
    }

    int hexval(int c) {
        if (c >= '0' && c <= '9') return c - '0';
        if (c >= 'a' && c <= 'f') return c - 'a' + 10;
        if (c >= 'A' && c <= 'F') return c - 'A' + 10;
        return -1;
    }

    // Escapes that denote one code point. `ch` is the char after the backslash (already consumed).
    // Returns -1 if `ch` is not such an escape.
    int single_escape(int ch) {
        switch (ch) {
            case '0': {
                int a = raw_peek();
                if (!(a >= '0' && a <= '7')) err("Illegal octal escape sequence");
                ++pos; int v = a - '0';
                int b = raw_peek();
                if (b >= '0' && b <= '7') {
                    ++pos; v = v * 8 + (b - '0');
                    int c = raw_peek();
                    if (c >= '0' && c <= '7' && a <= '3') { ++pos; v = v * 8 + (c - '0'); }
                }
                return v;
            }
            case 'a': return 7;
            case 'e': return 27;
            case 'f': return 12;
            case 'n': return 10;
            case 'r': return 13;
            case 't': return 9;
            case 'c': {
                int c = raw_peek();
                if (c == -1) err("Illegal control escape sequence");
                ++pos; return c ^ 64;
            }
            case 'x': {
                if (raw_peek() == '{') {
                    ++pos; long v = 0; int nd = 0;
                    while (raw_peek() != '}' && raw_peek() != -1) {
                        int h = hexval(p[pos]); if (h < 0) err("Illegal hexadecimal escape sequence");
                        v = v * 16 + h; ++pos; ++nd; if (v > 0x10FFFF) err("Hexadecimal codepoint is too big");
                    }
                    if (raw_peek() != '}' || nd == 0) err("Unclosed hexadecimal escape sequence");
                    ++pos; return (int)v;
                }
                int a = hexval(raw_peek()), b = hexval(raw_peek(1));
                if (a < 0 || b < 0) err("Illegal hexadecimal escape sequence");
                pos += 2; return a * 16 + b;
            }
            case 'u': {
                int v = 0;
                for (int k = 0; k < 4; ++k) { int h = hexval(raw_peek(k)); if (h < 0) err("Illegal Unicode escape sequence"); v = v * 16 + h; }
                pos += 4;
                if (v >= 0xD800 && v <= 0xDBFF && raw_peek() == '\\' && raw_peek(1) == 'u') {
                    int w = 0; bool ok = true;
                    for (int k = 0; k < 4; ++k) { int h = hexval(raw_peek(2 + k)); if (h < 0) { ok = false; break; } w = w * 16 + h; }
                    if (ok && w >= 0xDC00 && w <= 0xDFFF) { pos += 6; return 0x10000 + ((v - 0xD800) << 10) + (w - 0xDC00); }
                }
                if (v >= 0xD800 && v <= 0xDFFF) unsup("unpaired surrogate escape");
                return v;
            }
            default: return -1;
        }
    }

    NodeP linebreak() {  // \R
        NodeP alt = mk(N_ALT);
        NodeP crlf = mk(N_CAT);
        NodeP cr = mk(N_CHAR); cr->c = 13;
        NodeP lf = mk(N_CHAR); lf->c = 10;
        crlf->kids.push_back(std::move(cr)); crlf->kids.push_back(std::move(lf));
        alt->kids.push_back(std::move(crlf));
        alt->kids.push_back(setnode(cls_vspace()));
        return alt;
    }

    NodeP parse_escape_atom() {
        // backslash consumed
        if (pos >= len()) err("Unexpected internal error");  // trailing backslash
        int ch = p[pos++];
        CharSet cs;
        if (predefined(ch, cs)) return setnode(cs);
        int v = single_escape(ch);
        if (v >= 0) return lit(v);
        switch (ch) {
            case 'p': return setnode(property(false));
            case 'P': return setnode(property(true));
            case 'b': {
                if (raw_peek() == '{') unsup("\\b{g}");
                return bound(N_WORDB);
            }
            case 'B':
                return bound(N_NWORDB);
            case 'A': return mk(N_BEGIN);
            case 'z': return mk(N_END);
            case 'Z': { NodeP n = mk(N_EOL); n->flags = flags & F_UNIX_LINES; return n; }
            case 'G': return mk(N_BEGIN);        // LastMatch: oldLast is the search start for the first find() of a Matcher
            case 'X': unsup("\\X");
            case 'N': unsup("\\N{name}");
            case 'R': return linebreak();
            case 'k': {
                if (raw_peek() != '<') err("\\k is not followed by '<' for named capturing group");
                ++pos; std::string name;
                while (raw_peek() != '>' && raw_peek() != -1) name.push_back((char)p[pos++]);
                if (raw_peek() != '>') err("named capturing group is missing trailing '>'");
                ++pos;
                auto it = names.find(name);
                if (it == names.end()) err("named capturing group <" + name + "> does not exist");
                if (unicode_ci()) unsup("back-reference under (?iu)");
                NodeP n = mk(N_BACKREF); n->group = it->second; n->icase = has(F_CASE_INSENSITIVE); return n;
            }
            default: break;
        }
        if (ch >= '1' && ch <= '9') {
            int ref = ch - '0';
            if (ref >= ngroups) err("Illegal/unsupported escape sequence (back reference to a group that does not exist yet)");
            for (;;) {
                int d = raw_peek();
                if (!(d >= '0' && d <= '9')) break;
                int nr = ref * 10 + (d - '0');
                if (ngroups - 1 < nr) break;
                ref = nr; ++pos;
            }
            if (unicode_ci()) unsup("back-reference under (?iu)");
            NodeP n = mk(N_BACKREF); n->group = ref; n->icase = has(F_CASE_INSENSITIVE); return n;
        }
        if ((ch >= 'a' && ch <= 'z') || (ch >= 'A' && ch <= 'Z')) err("Illegal/unsupported escape sequence");
        return lit(ch);
    }

    // One class member that is a single code point (literal or escape) or a predefined set.
    // Returns true and sets `cp` for a single code point; false and fills `cs` for a set.
    bool class_item(int& cp, CharSet& cs) {
        int ch = next();
        if (ch == -1) err("Unclosed character class");
        if (ch != '\\') { cp = ch; return true; }
        if (pos >= len()) err("Unclosed character class");
        int e = p[pos++];
        if (predefined(e, cs)) return false;
        int v = single_escape(e);
        if (v >= 0) { cp = v; return true; }
        if (e == 'p') { cs = property(false); return false; }
        if (e == 'P') { cs = property(true); return false; }
        if ((e >= 'a' && e <= 'z') || (e >= 'A' && e <= 'Z') || (e >= '1' && e <= '9')) err("Illegal/unsupported escape sequence");
        cp = e; return true;
    }

    // Union of members up to (not consuming) ']' or "&&".
    CharSet class_union(bool& have) {
        CharSet lits, preds;
        for (;;) {
            int ch = peek();
            if (ch == -1) err("Unclosed character class");
            if (ch == ']' && have) break;
            if (ch == '&' && raw_peek(1) == '&') break;
            if (ch == '[') { next(); CharSet n = parse_class(); preds.add(n); have = true; continue; }
            int cp = 0; CharSet cs;
            if (!class_item(cp, cs)) { preds.add(cs); have = true; continue; }
            have = true;
            if (peek() == '-' && raw_peek(1) != ']' && raw_peek(1) != -1 && raw_peek(1) != '[') {
                next();
                int cp2 = 0; CharSet cs2;
                if (!class_item(cp2, cs2)) err("Illegal character range");     // Pattern.range(): escape(..) returns -1 for a class, and -1 < ch
                if (cp2 < cp) err("Illegal character range");
                lits.add(cp, cp2);
            } else lits.add(cp, cp);
        }
        if (unicode_ci()) close_unicode_case(lits); else if (has(F_CASE_INSENSITIVE)) lits.close_ascii_case();
        lits.add(preds);
        lits.normalize();
        return lits;
    }

    // after '[' ; consumes the closing ']'
    CharSet parse_class() {
        bool neg = false;
        if (peek() == '^') { next(); neg = true; }
        bool have = false;
        CharSet cur = class_union(have);
        while (peek() == '&' && raw_peek(1) == '&') {
            pos += 2;
            CharSet right; bool have_r = false;
            while (peek() != ']' && !(peek() == '&' && raw_peek(1) == '&')) {
                if (peek() == -1) err("Unclosed character class");
                CharSet part;
                if (peek() == '[') { next(); part = parse_class(); }
                else { bool h2 = true; part = class_union(h2); }
                if (have_r) right.intersect(part); else { right = part; have_r = true; }
            }
            if (!have_r) { if (!have) err("Bad class syntax"); continue; }
            if (have) cur.intersect(right); else { cur = right; have = true; }
        }
        if (peek() != ']') err("Unclosed character class");
        next();
        if (neg) cur.negate(); else cur.normalize();
        return cur;
    }

    NodeP parse_atom(bool& directive) {
        int ch = next();
        switch (ch) {
            case '(': return parse_group(directive);
            case '[': return setnode(parse_class());
            case '.': { NodeP n = mk(N_ANY); n->flags = flags; return n; }
            case '^': {
                if (has(F_MULTILINE)) { NodeP n = mk(N_BOL); n->flags = flags; return n; }
                return mk(N_BEGIN);
            }
            case '$': { NodeP n = mk(N_EOL); n->flags = flags & (F_MULTILINE | F_UNIX_LINES); return n; }
            case '\\': return parse_escape_atom();
            case '*': case '+': case '?': --pos; err(std::string("Dangling meta character '") + (char)ch + "'");
            case '{': --pos; return mk(N_EMPTY);  // closure() then raises "Illegal repetition" unless {n..}
            default: return lit(ch);
        }
    }
};

// min/max length in code points (max<0 = unbounded)
void measure(Node* n) {
    for (auto& k : n->kids) measure(k.get());
    auto addmax = [](int a, int b) { return (a < 0 || b < 0) ? -1 : (a > INT_MAX - b ? -1 : a + b); };
    switch (n->kind) {
        case N_CHAR: case N_SET: case N_ANY: n->minLen = n->maxLen = 1; break;
        case N_GROUP: case N_ATOMIC: n->minLen = n->kids[0]->minLen; n->maxLen = n->kids[0]->maxLen; break;
        case N_ALT: {
            int mn = INT_MAX, mx = 0;
            for (auto& k : n->kids) { mn = std::min(mn, k->minLen); mx = (mx < 0 || k->maxLen < 0) ? -1 : std::max(mx, k->maxLen); }
            n->minLen = mn; n->maxLen = mx; break;
        }
        case N_CAT: {
            int mn = 0, mx = 0;
            for (auto& k : n->kids) { mn = std::min<long>((long)mn + k->minLen, INT_MAX / 2); mx = addmax(mx, k->maxLen); }
            n->minLen = mn; n->maxLen = mx; break;
        }
        case N_REPEAT: {
            Node* k = n->kids[0].get();
            long mn = (long)k->minLen * n->min;
            n->minLen = (int)std::min<long>(mn, INT_MAX / 2);
            if (n->max < 0) n->maxLen = (k->maxLen == 0) ? 0 : -1;
            else if (k->maxLen < 0) n->maxLen = -1;
            else { long mx = (long)k->maxLen * n->max; n->maxLen = mx > INT_MAX / 2 ? -1 : (int)mx; }
            break;
        }
        case N_BACKREF: n->minLen = 0; n->maxLen = -1; break;
        default: n->minLen = n->maxLen = 0; break;
    }
}

void check_lookbehind(const Node* n) {
    if (n->kind == N_LOOK && n->behind && n->kids[0]->maxLen < 0)
        throw SyntaxError("Look-behind group does not have an obvious maximum length", 0);
    for (auto& k : n->kids) check_lookbehind(k.get());
}

// first-code-point analysis: returns nullable; accumulates into fs / any
bool first_of(const Node* n, CharSet& fs, bool& any) {
    switch (n->kind) {
        case N_CHAR:
            fs.add(n->c, n->c);
            if (n->icase) { fs.add(ascii_lower(n->c), ascii_lower(n->c)); fs.add(ascii_upper(n->c), ascii_upper(n->c)); }
            return false;
        case N_SET: fs.add(n->set); return false;
        case N_ANY: any = true; return false;
        case N_GROUP: case N_ATOMIC: return first_of(n->kids[0].get(), fs, any);
        case N_ALT: { bool nul = false; for (auto& k : n->kids) nul |= first_of(k.get(), fs, any); return nul; }
        case N_CAT: { for (auto& k : n->kids) if (!first_of(k.get(), fs, any)) return false; return true; }
        case N_REPEAT: { bool nul = first_of(n->kids[0].get(), fs, any); return nul || n->min == 0; }
        case N_BACKREF: any = true; return true;
        default: return true;  // zero-width
    }
}

// ------------------------------------------------------------------ Matcher
struct FR {
    void* obj;
    bool (*fn)(void*, int);
    bool operator()(int i) const { return fn(obj, i); }
};
template <class F>
FR mkfr(F& f) {
    return FR{&f, [](void* o, int i) { return (*static_cast<F*>(o))(i); }};
}

struct Ctx {
    const int* t;
    int n;
    std::vector<int> cap;
};

inline bool is_word(int c) {
    return c < 128 && ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9') || c == '_');
}
inline bool is_term(int c) { return c == '\n' || c == '\r' || c == 0x85 || c == 0x2028 || c == 0x2029; }

inline bool one(const Node* nd, int c) {
    switch (nd->kind) {
        case N_CHAR: return c == nd->c || (nd->icase && c < 128 && ascii_lower(c) == ascii_lower(nd->c));
        case N_SET: return nd->set.has(c);
        case N_ANY:
            if (nd->flags & F_DOTALL) return true;
            if (nd->flags & F_UNIX_LINES) return c != '\n';
            return !is_term(c);
        default: return false;
    }
}
inline bool is_single(const Node* nd) { return nd->kind == N_CHAR || nd->kind == N_SET || nd->kind == N_ANY; }

bool m(const Node* nd, int i, Ctx& cx, FR k);

bool m_cat(const Node* nd, size_t idx, int i, Ctx& cx, FR k) {
    if (idx == nd->kids.size()) return k(i);
    if (idx + 1 == nd->kids.size()) return m(nd->kids[idx].get(), i, cx, k);
    auto cont = [&](int j) { return m_cat(nd, idx + 1, j, cx, k); };
    return m(nd->kids[idx].get(), i, cx, mkfr(cont));
}

bool rep(const Node* nd, int count, int i, Ctx& cx, FR k) {
    // Loop / LazyLoop: an iteration that consumed nothing goes straight on to `next`
    // ("if (i > matcher.locals[beginIndex])" in Loop.match), keeping what it captured.
    const Node* body = nd->kids[0].get();
    auto iter = [&](int j) {
        if (j == i) return k(j);
        return rep(nd, count + 1, j, cx, k);
    };
    if (count < nd->min) return m(body, i, cx, mkfr(iter));
    bool more = nd->max < 0 || count < nd->max;
    if (nd->mode == 0) {
        if (more && m(body, i, cx, mkfr(iter))) return true;
        return k(i);
    }
    if (k(i)) return true;
    if (more) return m(body, i, cx, mkfr(iter));
    return false;
}

bool m(const Node* nd, int i, Ctx& cx, FR k) {
    switch (nd->kind) {
        case N_EMPTY: return k(i);
        case N_CHAR: case N_SET: case N_ANY:
            return i < cx.n && one(nd, cx.t[i]) && k(i + 1);
        case N_BEGIN: return i == 0 && k(i);
        case N_END: return i == cx.n && k(i);
        case N_BOL: {  // Caret / UnixCaret (MULTILINE)
            if (i == cx.n) return false;
            if (i > 0) {
                int ch = cx.t[i - 1];
                if (nd->flags & F_UNIX_LINES) { if (ch != '\n') return false; }
                else {
                    if (!is_term(ch)) return false;
                    if (ch == '\r' && cx.t[i] == '\n') return false;
                }
            }
            return k(i);
        }
        case N_EOL: {  // Dollar / UnixDollar; \Z is the non-multiline form
            bool ml = (nd->flags & F_MULTILINE) != 0;
            int n = cx.n;
            if (nd->flags & F_UNIX_LINES) {
                if (i < n) {
                    if (cx.t[i] != '\n') return false;
                    if (!ml && i != n - 1) return false;
                }
                return k(i);
            }
            if (!ml) {
                if (i < n - 2) return false;
                if (i == n - 2) { if (cx.t[i] != '\r' || cx.t[i + 1] != '\n') return false; }
            }
            if (i < n) {
                int ch = cx.t[i];
                if (ch == '\n') { if (i > 0 && cx.t[i - 1] == '\r') return false; }
                else if (ch == '\r' || ch == 0x85 || ch == 0x2028 || ch == 0x2029) {}
                else return false;
            }
            return k(i);
        }
        case N_WORDB: case N_NWORDB: {
            const bool uni = (nd->flags & F_UNICODE_CLASS) != 0;
            bool l = i > 0 && (uni ? nd->set.has(cx.t[i - 1]) : is_word(cx.t[i - 1]));
            bool r = i < cx.n && (uni ? nd->set.has(cx.t[i]) : is_word(cx.t[i]));
            bool b = l != r;
            return (nd->kind == N_WORDB ? b : !b) && k(i);
        }
        case N_GROUP: {
            int g = nd->group;
            if (g < 0) return m(nd->kids[0].get(), i, cx, k);
            int os = cx.cap[2 * g], oe = cx.cap[2 * g + 1];
            auto cont = [&](int j) {
                int ps = cx.cap[2 * g], pe = cx.cap[2 * g + 1];
                cx.cap[2 * g] = i; cx.cap[2 * g + 1] = j;
                if (k(j)) return true;
                cx.cap[2 * g] = ps; cx.cap[2 * g + 1] = pe;
                return false;
            };
            if (m(nd->kids[0].get(), i, cx, mkfr(cont))) return true;
            cx.cap[2 * g] = os; cx.cap[2 * g + 1] = oe;
            return false;
        }
        case N_ALT:
            for (auto& kid : nd->kids) if (m(kid.get(), i, cx, k)) return true;
            return false;
        case N_CAT: return m_cat(nd, 0, i, cx, k);
        case N_REPEAT: {
            const Node* body = nd->kids[0].get();
            if (is_single(body)) {  // Curly on a one-character atom
                int mx = nd->max < 0 ? INT_MAX : nd->max;
                if (nd->mode == 1) {
                    int j = i, cnt = 0;
                    while (cnt < nd->min) { if (j < cx.n && one(body, cx.t[j])) { ++j; ++cnt; } else return false; }
                    for (;;) {
                        if (k(j)) return true;
                        if (cnt >= mx || j >= cx.n || !one(body, cx.t[j])) return false;
                        ++j; ++cnt;
                    }
                }
                int j = i, cnt = 0;
                while (cnt < mx && j < cx.n && one(body, cx.t[j])) { ++j; ++cnt; }
                if (cnt < nd->min) return false;
                if (nd->mode == 2) return k(j);
                for (; cnt >= nd->min; --cnt, --j) if (k(j)) return true;
                return false;
            }
            if (nd->mode == 2) {  // possessive on a group: each iteration is committed
                int j = i, cnt = 0;
                int mx = nd->max < 0 ? INT_MAX : nd->max;
                while (cnt < mx) {
                    int e = -1;
                    auto take = [&](int x) { e = x; return true; };
                    if (!m(body, j, cx, mkfr(take))) break;
                    ++cnt;
                    if (e == j) { if (cnt < nd->min) cnt = nd->min; break; }
                    j = e;
                }
                if (cnt < nd->min) return false;
                return k(j);
            }
            return rep(nd, 0, i, cx, k);
        }
        case N_ATOMIC: {
            int e = -1;
            auto take = [&](int x) { e = x; return true; };
            if (!m(nd->kids[0].get(), i, cx, mkfr(take))) return false;
            return k(e);
        }
        case N_LOOK: {
            const Node* body = nd->kids[0].get();
            // Pos/Neg/Behind/NotBehind run the condition as a stand-alone match and never roll its
            // captures back, whatever happens afterwards (groups set inside a look-around that
            // matched stay set even when the enclosing path later fails).
            bool ok = false;
            if (!nd->behind) {
                auto yes = [&](int) { return true; };
                ok = m(body, i, cx, mkfr(yes));
            } else {
                int lo = body->maxLen < 0 ? 0 : std::max(0, i - body->maxLen);
                for (int j = i - body->minLen; j >= lo && !ok; --j) {
                    auto at = [&](int e) { return e == i; };
                    ok = m(body, j, cx, mkfr(at));
                }
            }
            return (ok != nd->neg) && k(i);
        }
        case N_BACKREF: {
            int s = cx.cap[2 * nd->group], e = cx.cap[2 * nd->group + 1];
            if (s < 0 || e < 0) return false;
            int len = e - s;
            if (i + len > cx.n) return false;
            for (int x = 0; x < len; ++x) {
                int a = cx.t[s + x], b = cx.t[i + x];
                if (a != b) {
                    if (!(nd->icase && a < 128 && b < 128 && ascii_lower(a) == ascii_lower(b))) return false;
                }
            }
            return k(i + len);
        }
    }
    return false;
}

}  // namespace

// ------------------------------------------------------------------ Regex
Regex::Regex(const std::string& pattern, int flags) {
    Parser ps;
    decode_pattern(pattern, ps.p);
    ps.flags = flags;
    ps.check_flags();
    ps.remove_qe();
    NodeP r = ps.parse_alt();
    if (ps.pos < ps.len()) {
        if (ps.p[ps.pos] == ')') throw SyntaxError("Unmatched closing ')'", ps.pos);
        throw SyntaxError("Unexpected internal error", ps.pos);
    }
    measure(r.get());
    check_lookbehind(r.get());
    ngroups_ = ps.ngroups;
    {   // a back-reference can see a capture left behind by an attempt at an EARLIER start position (groups are
        // cleared once per find()), so no start position may be skipped when the pattern has one
        std::vector<const Node*> todo{r.get()};
        while (!todo.empty()) {
            const Node* x = todo.back();
            todo.pop_back();
            if (x->kind == N_BACKREF) has_backref_ = true;
            for (const auto& k : x->kids) todo.push_back(k.get());
        }
    }
    bool any = false;
    CharSet fs;
    can_empty_ = first_of(r.get(), fs, any);
    first_any_ = any;
    fs.normalize();
    first_ = fs;
    // Begin-anchored: every alternative starts with \A / non-multiline ^ (cheap check: top node or first of CAT)
    const Node* t = r.get();
    while (t->kind == N_GROUP) t = t->kids[0].get();
    if (t->kind == N_BEGIN) anchored_begin_ = true;
    else if (t->kind == N_CAT && t->kids[0]->kind == N_BEGIN) anchored_begin_ = true;
    root_ = std::move(r);
}
Regex::~Regex() = default;

bool Regex::find_cp(const int* t, int n, int* start, int* end) const {
    Ctx cx;
    cx.t = t;
    cx.n = n;
    cx.cap.assign(2 * (size_t)ngroups_, -1);
    int e = -1;
    auto fin = [&](int j) { e = j; return true; };
    const bool skip = first_opt_ && !has_backref_;
    int last = anchored_begin_ && skip ? 0 : n;
    for (int s = 0; s <= last; ++s) {
        if (skip && !can_empty_ && !first_any_) {
            if (s >= n) break;
            if (!first_.has(t[s])) continue;
        }
        // Matcher.search() clears the groups once per find(), not per start position
        if (m(root_.get(), s, cx, mkfr(fin))) {
            if (start) *start = s;
            if (end) *end = e;
            return true;
        }
    }
    return false;
}

bool Regex::find_span(const uint8_t* s, size_t n, int* start, int* end) const {
    // ASCII fast path avoids the decode allocation for the common case
    std::vector<int> cps;
    decode_utf8(s, n, cps);
    return find_cp(cps.data(), (int)cps.size(), start, end);
}

bool Regex::find(const uint8_t* s, size_t n) const { return find_span(s, n, nullptr, nullptr); }

}  // namespace orx
