// Required-literal extraction for the device prefilter: a set L of byte strings such that every
// match of the regex contains some member of L. Lines holding no member of L cannot satisfy
// Matcher.find() (AnalysisService.java:95), so only lines where the gram filter fires are handed
// to the DFA / backtracking verifiers. Lookaround bodies are ignored (sound: they consume nothing).
//
// Per node the analysis keeps: the exact language when it is a small finite set; a set every match
// must START with and one it must END with (so that `\d{8,}ms` yields "00ms".."99ms"); and the best
// factor set seen anywhere inside. Small character classes (digits, [45], [Ff]) are expanded.
#include <algorithm>
#include <cmath>

#include "rx.h"

namespace lpx {
namespace {

constexpr size_t MAX_SET = 256;    // alternatives kept in one literal set
constexpr size_t MAX_LEN = 48;     // bytes kept per literal
constexpr size_t MAX_CLASS = 16;   // a class with at most this many code points is expanded

using LitSet = std::vector<Literal>;

struct Info {
    bool exact = false;     // `lits` is the complete (finite) language of the node
    LitSet lits;
    bool has_pre = false, has_suf = false;
    LitSet pre, suf;        // every match starts with a member of pre / ends with a member of suf
    LitSet best;            // best required-factor set seen inside (may be empty)
};

std::string utf8(uint32_t c) {
    std::string s;
    if (c < 0x80) s.push_back((char)c);
    else if (c < 0x800) { s.push_back((char)(0xC0 | (c >> 6))); s.push_back((char)(0x80 | (c & 0x3F))); }
    else if (c < 0x10000) { s.push_back((char)(0xE0 | (c >> 12))); s.push_back((char)(0x80 | ((c >> 6) & 0x3F))); s.push_back((char)(0x80 | (c & 0x3F))); }
    else { s.push_back((char)(0xF0 | (c >> 18))); s.push_back((char)(0x80 | ((c >> 12) & 0x3F))); s.push_back((char)(0x80 | ((c >> 6) & 0x3F))); s.push_back((char)(0x80 | (c & 0x3F))); }
    return s;
}
inline bool alpha(uint32_t c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z'); }

// usefulness of a factor set: long members are selective, many members cost filter entries
double utility(const LitSet& v) {
    if (v.empty()) return -1e9;
    size_t m = SIZE_MAX;
    for (const auto& l : v) m = std::min(m, l.bytes.size());
    if (m == 0) return -1e9;
    return (double)std::min<size_t>(m, 12) - 2.0 * std::log2((double)v.size());
}
size_t min_len_of(const LitSet& v) {
    size_t m = SIZE_MAX;
    for (const auto& l : v) m = std::min(m, l.bytes.size());
    return v.empty() ? 0 : m;
}

void dedupe(LitSet& v) {
    std::sort(v.begin(), v.end(), [](const Literal& a, const Literal& b) { return a.bytes != b.bytes ? a.bytes < b.bytes : a.icase < b.icase; });
    v.erase(std::unique(v.begin(), v.end(), [](const Literal& a, const Literal& b) { return a.bytes == b.bytes && a.icase == b.icase; }), v.end());
}

bool cross(const LitSet& a, const LitSet& b, LitSet& out) {
    if (a.size() * b.size() > MAX_SET) return false;
    out.clear();
    for (const auto& x : a)
        for (const auto& y : b) {
            if (x.bytes.size() + y.bytes.size() > MAX_LEN) return false;
            Literal l;
            l.icase = x.icase || y.icase;
            l.bytes = x.bytes + y.bytes;
            auto letters = [](const std::string& b) { for (unsigned char ch : b) if ((ch | 0x20) >= 'a' && (ch | 0x20) <= 'z') return true; return false; };
            l.lossy = x.lossy || y.lossy || (l.icase && ((!x.icase && letters(x.bytes)) || (!y.icase && letters(y.bytes))));
            if (l.icase) for (auto& ch : l.bytes) if (ch >= 'A' && ch <= 'Z') ch = (char)(ch + 32);
            out.push_back(std::move(l));
        }
    dedupe(out);
    return true;
}

struct Picker {
    size_t min_len;
    LitSet best;
    void consider(const LitSet& cand) {
        if (cand.empty() || min_len_of(cand) < 1) return;
        const bool ok_c = min_len_of(cand) >= min_len, ok_b = !best.empty() && min_len_of(best) >= min_len;
        if (best.empty() || (ok_c && !ok_b) || (ok_c == ok_b && utility(cand) > utility(best))) best = cand;
    }
};

const LitSet EMPTY_WORD{Literal{"", false}};

Info walk(const Ast& a, size_t min_len) {
    Info r;
    Picker pk{min_len, {}};
    auto finish = [&]() {
        if (r.exact) { r.has_pre = r.has_suf = true; r.pre = r.suf = r.lits; pk.consider(r.lits); }
        if (r.has_pre) pk.consider(r.pre);
        if (r.has_suf) pk.consider(r.suf);
        r.best = pk.best;
        return r;
    };
    switch (a.kind) {
        case AstKind::Empty: case AstKind::Assert: case AstKind::Look:
            r.exact = true; r.lits = EMPTY_WORD; r.has_pre = r.has_suf = true; r.pre = r.suf = EMPTY_WORD; return r;
        case AstKind::Backref: return r;
        case AstKind::Set: {
            size_t count = 0;
            for (const auto& x : a.set) { count += (size_t)(x.hi - x.lo) + 1; if (count > MAX_CLASS + 1) break; }
            if (count == 0 || count > MAX_CLASS + 1) return r;
            std::vector<uint32_t> cps;
            for (const auto& x : a.set) for (uint32_t c = x.lo; c <= x.hi; ++c) cps.push_back(c);
            r.exact = true;
            // ASCII case pairs collapse to one case-insensitive member
            std::vector<char> used(cps.size(), 0);
            for (size_t i = 0; i < cps.size(); ++i) {
                if (used[i]) continue;
                bool paired = false;
                if (alpha(cps[i]))
                    for (size_t j = i + 1; j < cps.size(); ++j)
                        if (!used[j] && (cps[i] ^ 0x20) == cps[j]) { used[j] = 1; paired = true; break; }
                if (paired) r.lits.push_back(Literal{std::string(1, (char)(cps[i] | 0x20)), true});
                else r.lits.push_back(Literal{utf8(cps[i]), false});
            }
            if (r.lits.size() > MAX_CLASS) { r.exact = false; r.lits.clear(); return r; }
            return finish();
        }
        case AstKind::Group: case AstKind::Atomic: return walk(*a.kids[0], min_len);
        case AstKind::Alt: {
            bool all_exact = true, all_best = true, all_pre = true, all_suf = true;
            LitSet uni, bestu, preu, sufu;
            for (const auto& k : a.kids) {
                Info ki = walk(*k, min_len);
                if (ki.exact) uni.insert(uni.end(), ki.lits.begin(), ki.lits.end()); else all_exact = false;
                if (ki.best.empty() || min_len_of(ki.best) < min_len) all_best = false; else bestu.insert(bestu.end(), ki.best.begin(), ki.best.end());
                if (ki.has_pre) preu.insert(preu.end(), ki.pre.begin(), ki.pre.end()); else all_pre = false;
                if (ki.has_suf) sufu.insert(sufu.end(), ki.suf.begin(), ki.suf.end()); else all_suf = false;
            }
            dedupe(uni); dedupe(bestu); dedupe(preu); dedupe(sufu);
            if (all_exact && uni.size() <= MAX_SET) { r.exact = true; r.lits = uni; }
            if (all_pre && preu.size() <= MAX_SET) { r.has_pre = true; r.pre = preu; }
            if (all_suf && sufu.size() <= MAX_SET) { r.has_suf = true; r.suf = sufu; }
            if (all_best && bestu.size() <= MAX_SET) pk.consider(bestu);
            if (r.exact) pk.consider(r.lits);
            r.best = pk.best;
            return r;
        }
        case AstKind::Cat: {
            // `acc`: every match of the kids seen so far ends with a member of acc (it may carry the
            // required suffix of an inexact kid in); `fresh`: the exact run since the last inexact kid
            LitSet acc = EMPTY_WORD, fresh = EMPTY_WORD;
            bool whole = true;      // acc is still the exact language of the prefix
            for (const auto& k : a.kids) {
                Info ki = walk(*k, min_len);
                pk.consider(ki.best);
                if (ki.exact) {
                    LitSet nx;
                    if (cross(acc, ki.lits, nx)) { acc.swap(nx); }
                    else {
                        pk.consider(acc);
                        if (whole) { r.has_pre = true; r.pre = acc; }
                        whole = false;
                        acc = ki.lits;
                    }
                    if (cross(fresh, ki.lits, nx)) fresh.swap(nx);
                    else { pk.consider(fresh); fresh = ki.lits; }
                } else {
                    if (ki.has_pre) { LitSet nx; if (cross(acc, ki.pre, nx)) pk.consider(nx); if (cross(fresh, ki.pre, nx)) pk.consider(nx); }
                    pk.consider(acc);
                    pk.consider(fresh);
                    if (whole) {   // everything before this kid was exact: it is a required prefix
                        LitSet nx;
                        r.has_pre = true;
                        if (ki.has_pre && cross(acc, ki.pre, nx)) r.pre = nx; else r.pre = acc;
                    }
                    whole = false;
                    acc = ki.has_suf ? ki.suf : EMPTY_WORD;
                    fresh = EMPTY_WORD;
                }
            }
            pk.consider(acc);
            pk.consider(fresh);
            if (whole) { r.exact = true; r.lits = acc; }
            r.has_suf = true; r.suf = acc;
            if (r.exact) { r.has_pre = true; r.pre = acc; }
            r.best = pk.best;
            return r;
        }
        case AstKind::Repeat: {
            Info ki = walk(*a.kids[0], min_len);
            if (a.min == 0) {
                if (a.max == 1 && ki.exact && ki.lits.size() < MAX_SET) {
                    r.exact = true; r.lits = ki.lits; r.lits.push_back(Literal{"", false}); dedupe(r.lits);
                }
                return r;     // may match nothing: requires nothing
            }
            pk.consider(ki.best);
            if (ki.exact) {
                pk.consider(ki.lits);
                // fixed small counts are exact; otherwise the match starts and ends with min(min,2) copies
                const int copies = std::min(a.min, a.min == a.max ? 4 : 2);
                LitSet acc = EMPTY_WORD;
                bool ok = true;
                for (int i = 0; i < copies && ok; ++i) { LitSet nx; ok = cross(acc, ki.lits, nx); if (ok) acc.swap(nx); }
                if (ok && a.min == a.max && copies == a.min) { r.exact = true; r.lits = acc; }
                else {
                    r.has_pre = r.has_suf = true;
                    r.pre = r.suf = ok ? acc : ki.lits;
                }
            } else {
                if (ki.has_pre) { r.has_pre = true; r.pre = ki.pre; }
                if (ki.has_suf) { r.has_suf = true; r.suf = ki.suf; }
            }
            return finish();
        }
    }
    return r;
}

}  // namespace

namespace {

const Ast* strip_groups(const Ast* a) {
    while (a->kind == AstKind::Group) a = a->kids[0].get();
    return a;
}
bool is_word_set(const CpSet& s) {
    return s.size() == 4 && s[0].lo == '0' && s[0].hi == '9' && s[1].lo == 'A' && s[1].hi == 'Z' && s[2].lo == '_' && s[2].hi == '_' &&
           s[3].lo == 'a' && s[3].hi == 'z';
}
inline bool word_byte_l(unsigned char b) { return (b >= '0' && b <= '9') || (b >= 'A' && b <= 'Z') || b == '_' || (b >= 'a' && b <= 'z'); }

// walk() treats zero-width nodes as the empty string, which is sound for "every match contains this
// literal" but not for deciding a match: a branch that hides an assertion, a look-around or a
// back-reference anywhere inside (e.g. `word(?: \A)`) is not a plain literal.
bool has_zero_width_or_backref(const Ast& a) {
    if (a.kind == AstKind::Assert || a.kind == AstKind::Look || a.kind == AstKind::Backref) return true;
    for (const auto& k : a.kids) if (has_zero_width_or_backref(*k)) return true;
    return false;
}

bool local_branch(const Ast* br, std::vector<LocalLiteral>& out, size_t min_len) {
    br = strip_groups(br);
    std::vector<const Ast*> kids;
    if (br->kind == AstKind::Cat) for (const auto& k : br->kids) kids.push_back(k.get());
    else kids.push_back(br);
    size_t lo = 0, hi = kids.size();
    bool left_b = false, right_b = false, skipped_w = false;
    if (lo < hi && kids[lo]->kind == AstKind::Assert && kids[lo]->assert_kind == AssertKind::WordB) { left_b = true; ++lo; }
    if (lo < hi && kids[lo]->kind == AstKind::Repeat && kids[lo]->min == 0 && kids[lo]->max < 0 && kids[lo]->mode != 2 &&
        kids[lo]->kids[0]->kind == AstKind::Set && is_word_set(kids[lo]->kids[0]->set)) { skipped_w = true; ++lo; }
    if (lo < hi && kids[hi - 1]->kind == AstKind::Assert && kids[hi - 1]->assert_kind == AssertKind::WordB) { right_b = true; --hi; }
    if (lo >= hi) return false;
    LitSet acc = EMPTY_WORD;
    for (size_t i = lo; i < hi; ++i) {
        if (has_zero_width_or_backref(*kids[i])) return false;
        Info ki = walk(*kids[i], min_len);
        if (!ki.exact) return false;
        LitSet nx;
        if (!cross(acc, ki.lits, nx)) return false;
        acc.swap(nx);
    }
    if (acc.empty() || acc.size() > 64) return false;
    for (const auto& l : acc) {
        if (l.bytes.size() < min_len || l.lossy) return false;
        const bool first_word = word_byte_l((unsigned char)l.bytes[0]);
        LocalLiteral ll;
        ll.lit = l;
        ll.right_boundary = right_b;
        if (skipped_w) {
            // "\w*LIT" and "\b\w*LIT" hold at every occurrence whose first byte is a word byte
            // (the word run that contains it starts at a boundary); otherwise keep the automaton
            if (!first_word) return false;
            ll.left_boundary = false;
        } else ll.left_boundary = left_b;
        out.push_back(std::move(ll));
    }
    return true;
}

}  // namespace

bool local_verify_form(const Ast& root, std::vector<LocalLiteral>& out, size_t min_len) {
    out.clear();
    const Ast* r = strip_groups(&root);
    if (r->kind == AstKind::Alt) {
        for (const auto& k : r->kids) if (!local_branch(k.get(), out, min_len)) { out.clear(); return false; }
        return !out.empty();
    }
    if (!local_branch(r, out, min_len)) { out.clear(); return false; }
    return true;
}

std::vector<Literal> required_literals(const Ast& root, size_t min_len) {
    Info i = walk(root, min_len);
    Picker pk{min_len, i.best};
    if (i.exact) pk.consider(i.lits);
    if (i.has_pre) pk.consider(i.pre);
    if (i.has_suf) pk.consider(i.suf);
    if (pk.best.empty() || min_len_of(pk.best) < min_len) return {};
    for (const auto& l : pk.best) if (l.bytes.empty()) return {};
    return pk.best;
}

}  // namespace lpx
