// Required-literal extraction for the device prefilter: a set L of byte strings such that every
// match of the regex contains some member of L. Lines holding no member of L cannot satisfy
// Matcher.find() (AnalysisService.java:95), so only lines where the gram filter fires are handed
// to the DFA / backtracking verifiers. Lookaround bodies are ignored (sound: they consume nothing).
#include <algorithm>

#include "rx.h"

namespace lpx {
namespace {

constexpr size_t MAX_SET = 24;     // alternatives kept in one literal set
constexpr size_t MAX_LEN = 48;     // bytes kept per literal

struct Info {
    bool exact = false;            // `lits` is the complete (finite) language of the node
    std::vector<Literal> lits;     // exact language, or empty
    std::vector<Literal> best;     // best required-factor set seen inside (may be empty)
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

// quality of a candidate set: (min length, -count); unusable if any member is shorter than min_len
struct Score { long min_len; long neg_count; };
Score score(const std::vector<Literal>& v) {
    if (v.empty()) return {0, 0};
    size_t m = SIZE_MAX;
    for (const auto& l : v) m = std::min(m, l.bytes.size());
    return {(long)m, -(long)v.size()};
}
bool better(const std::vector<Literal>& a, const std::vector<Literal>& b) {
    Score x = score(a), y = score(b);
    // prefer longer minimum length, but a much smaller set wins ties
    if (x.min_len != y.min_len) return x.min_len > y.min_len;
    return x.neg_count > y.neg_count;
}

void dedupe(std::vector<Literal>& v) {
    std::sort(v.begin(), v.end(), [](const Literal& a, const Literal& b) { return a.bytes != b.bytes ? a.bytes < b.bytes : a.icase < b.icase; });
    v.erase(std::unique(v.begin(), v.end(), [](const Literal& a, const Literal& b) { return a.bytes == b.bytes && a.icase == b.icase; }), v.end());
}

bool cross(const std::vector<Literal>& a, const std::vector<Literal>& b, std::vector<Literal>& out) {
    if (a.size() * b.size() > MAX_SET) return false;
    out.clear();
    for (const auto& x : a)
        for (const auto& y : b) {
            if (x.bytes.size() + y.bytes.size() > MAX_LEN) return false;
            Literal l;
            l.icase = x.icase || y.icase;
            l.bytes = x.bytes + y.bytes;
            if (l.icase) for (auto& ch : l.bytes) if (ch >= 'A' && ch <= 'Z') ch = (char)(ch + 32);
            out.push_back(std::move(l));
        }
    dedupe(out);
    return true;
}

void consider(std::vector<Literal>& best, const std::vector<Literal>& cand) {
    if (cand.empty()) return;
    for (const auto& l : cand) if (l.bytes.empty()) return;  // contains "", requires nothing
    if (best.empty() || better(cand, best)) best = cand;
}

Info walk(const Ast& a) {
    Info r;
    switch (a.kind) {
        case AstKind::Empty: case AstKind::Assert: case AstKind::Look:
            r.exact = true; r.lits = {Literal{"", false}}; return r;
        case AstKind::Backref: return r;
        case AstKind::Set: {
            // single code point, an ASCII case pair, or a handful of code points
            size_t count = 0;
            for (const auto& x : a.set) { count += (size_t)(x.hi - x.lo) + 1; if (count > 4) break; }
            if (count == 0 || count > 4) return r;
            std::vector<uint32_t> cps;
            for (const auto& x : a.set) for (uint32_t c = x.lo; c <= x.hi; ++c) cps.push_back(c);
            if (cps.size() == 2 && alpha(cps[0]) && (cps[0] ^ 0x20) == cps[1]) {
                r.exact = true; r.lits = {Literal{std::string(1, (char)(cps[0] | 0x20)), true}}; r.best = r.lits; return r;
            }
            r.exact = true;
            for (uint32_t c : cps) r.lits.push_back(Literal{utf8(c), false});
            r.best = r.lits;
            return r;
        }
        case AstKind::Group: case AstKind::Atomic: return walk(*a.kids[0]);
        case AstKind::Alt: {
            bool all_exact = true, all_best = true;
            std::vector<Literal> uni, bestu;
            for (const auto& k : a.kids) {
                Info ki = walk(*k);
                if (ki.exact) uni.insert(uni.end(), ki.lits.begin(), ki.lits.end()); else all_exact = false;
                if (ki.best.empty()) all_best = false; else bestu.insert(bestu.end(), ki.best.begin(), ki.best.end());
            }
            dedupe(uni);
            dedupe(bestu);
            if (all_exact && uni.size() <= MAX_SET) { r.exact = true; r.lits = uni; }
            if (all_best && bestu.size() <= MAX_SET) r.best = bestu;
            if (r.exact) consider(r.best, r.lits);
            return r;
        }
        case AstKind::Cat: {
            std::vector<Literal> acc{Literal{"", false}};
            bool whole = true;
            for (const auto& k : a.kids) {
                Info ki = walk(*k);
                consider(r.best, ki.best);
                if (ki.exact) {
                    std::vector<Literal> nx;
                    if (cross(acc, ki.lits, nx)) { acc.swap(nx); continue; }
                    // too large: bank what we have, restart from this kid
                    consider(r.best, acc);
                    whole = false;
                    acc = ki.lits;
                    if (acc.size() > MAX_SET) acc = {Literal{"", false}};
                    continue;
                }
                consider(r.best, acc);
                whole = false;
                acc = {Literal{"", false}};
            }
            consider(r.best, acc);
            if (whole) { r.exact = true; r.lits = acc; }
            return r;
        }
        case AstKind::Repeat: {
            Info ki = walk(*a.kids[0]);
            if (a.min == 0) {
                if (a.max == 1 && ki.exact && ki.lits.size() < MAX_SET) {
                    r.exact = true; r.lits = ki.lits; r.lits.push_back(Literal{"", false}); dedupe(r.lits);
                }
                return r;
            }
            // at least one iteration happens
            r.best = ki.best;
            if (ki.exact) {
                consider(r.best, ki.lits);
                if (a.min == a.max && a.min <= 4) {
                    std::vector<Literal> acc{Literal{"", false}};
                    bool ok = true;
                    for (int i = 0; i < a.min && ok; ++i) { std::vector<Literal> nx; ok = cross(acc, ki.lits, nx); if (ok) acc.swap(nx); }
                    if (ok) { r.exact = true; r.lits = acc; consider(r.best, acc); }
                }
            }
            return r;
        }
    }
    return r;
}

}  // namespace

std::vector<Literal> required_literals(const Ast& root, size_t min_len) {
    Info i = walk(root);
    std::vector<Literal> best = i.best;
    if (i.exact) consider(best, i.lits);
    if (best.empty()) return {};
    for (const auto& l : best) if (l.bytes.size() < min_len) return {};
    return best;
}

}  // namespace lpx
