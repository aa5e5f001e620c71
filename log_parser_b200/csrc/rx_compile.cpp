// AST -> UTF-8 byte NFA -> DFA tables for the device walker (existence of a match on one line;
// what Matcher.find() returns at AnalysisService.java:95, ScoringService.java:281,300,330 and
// ContextAnalysisService.java:64-79).
//
// Zero-width assertions are resolved during subset construction from one bit of left context
// kept in the DFA state (previous byte is an ASCII word byte / at line start) and the symbol
// about to be consumed (RE2-style). `$`/\Z need "next code point is the final line terminator":
// the walker announces that position with the virtual symbol SYM_TPOS; end of line is SYM_EOT.
#include <algorithm>
#include <cstring>
#include <map>
#include <unordered_map>

#include "rx.h"

namespace lpx {
namespace {

struct NEdge { uint8_t lo, hi; int to; };
struct NAssert { AssertKind k; int to; };
struct NState {
    std::vector<NEdge> edges;
    std::vector<int> eps;
    std::vector<NAssert> asserts;
};

struct Nfa {
    std::vector<NState> st;
    int accept = -1;
    bool uses_eol = false;
    uint32_t limit;
    explicit Nfa(uint32_t lim) : limit(lim) {}
    int add() {
        if (st.size() >= limit) throw std::length_error("nfa too large");
        st.emplace_back();
        return (int)st.size() - 1;
    }
};
struct Frag { int in, out; };

inline bool word_byte(int b) { return (b >= '0' && b <= '9') || (b >= 'A' && b <= 'Z') || b == '_' || (b >= 'a' && b <= 'z'); }

// ---- code point range -> UTF-8 byte-range sequences
struct Seq { int n; uint8_t lo[4], hi[4]; };

int enc(uint32_t c, uint8_t* o) {
    if (c < 0x80) { o[0] = (uint8_t)c; return 1; }
    if (c < 0x800) { o[0] = 0xC0 | (c >> 6); o[1] = 0x80 | (c & 0x3F); return 2; }
    if (c < 0x10000) { o[0] = 0xE0 | (c >> 12); o[1] = 0x80 | ((c >> 6) & 0x3F); o[2] = 0x80 | (c & 0x3F); return 3; }
    o[0] = 0xF0 | (c >> 18); o[1] = 0x80 | ((c >> 12) & 0x3F); o[2] = 0x80 | ((c >> 6) & 0x3F); o[3] = 0x80 | (c & 0x3F);
    return 4;
}
void split_range(uint32_t lo, uint32_t hi, std::vector<Seq>& out) {
    if (lo > hi) return;
    if (lo <= 0xDFFF && hi >= 0xD800) {  // surrogates are not scalar values
        if (lo < 0xD800) split_range(lo, 0xD7FF, out);
        if (hi > 0xDFFF) split_range(0xE000, hi, out);
        return;
    }
    static const uint32_t cut[3] = {0x7F, 0x7FF, 0xFFFF};
    for (uint32_t c : cut)
        if (lo <= c && hi > c) { split_range(lo, c, out); split_range(c + 1, hi, out); return; }
    if (hi < 0x80) { Seq s; s.n = 1; s.lo[0] = (uint8_t)lo; s.hi[0] = (uint8_t)hi; out.push_back(s); return; }
    for (int i = 1; i < 4; ++i) {
        uint32_t m = (1u << (6 * i)) - 1;
        if ((lo & ~m) != (hi & ~m)) {
            if ((lo & m) != 0) { split_range(lo, lo | m, out); split_range((lo | m) + 1, hi, out); return; }
            if ((hi & m) != m) { split_range(lo, (hi & ~m) - 1, out); split_range(hi & ~m, hi, out); return; }
        }
    }
    uint8_t a[4], b[4];
    int n = enc(lo, a);
    enc(hi, b);
    Seq s; s.n = n;
    for (int i = 0; i < n; ++i) { s.lo[i] = a[i]; s.hi[i] = b[i]; }
    out.push_back(s);
}

struct Builder {
    Nfa& n;
    bool rev;     // build the automaton of the reversed language (walked from a match's end towards the line start)
    explicit Builder(Nfa& nfa, bool reversed = false) : n(nfa), rev(reversed) {}

    Frag set_frag(const CpSet& set) {
        int in = n.add(), out = n.add();
        std::vector<Seq> seqs;
        for (const auto& r : set) split_range(r.lo, std::min<uint32_t>(r.hi, 0x10FFFF), seqs);
        if (rev) {      // bytes of each code point in reverse order; plain chains
            for (const auto& s : seqs) {
                int cur = in;
                for (int i = s.n - 1; i >= 0; --i) {
                    const int to = i == 0 ? out : n.add();
                    n.st[cur].edges.push_back({s.lo[i], s.hi[i], to});
                    cur = to;
                }
            }
            return {in, out};
        }
        // share continuation tails keyed by (remaining ranges) to keep big classes small
        std::map<std::vector<uint8_t>, int> tails;
        for (const auto& s : seqs) {
            int cur = in;
            for (int i = 0; i < s.n; ++i) {
                int to;
                if (i == s.n - 1) to = out;
                else {
                    std::vector<uint8_t> key;
                    for (int k = i + 1; k < s.n; ++k) { key.push_back(s.lo[k]); key.push_back(s.hi[k]); }
                    key.push_back((uint8_t)(s.n - i));
                    auto it = tails.find(key);
                    if (it != tails.end()) { to = it->second; n.st[cur].edges.push_back({s.lo[i], s.hi[i], to}); goto next_seq; }
                    to = n.add();
                    tails[key] = to;
                }
                n.st[cur].edges.push_back({s.lo[i], s.hi[i], to});
                cur = to;
            }
        next_seq:;
        }
        return {in, out};
    }

    Frag empty_frag() { int s = n.add(); return {s, s}; }

    Frag repeat(const Ast& body, int lo, int hi) {
        Frag acc = empty_frag();
        auto append = [&](Frag f) { n.st[acc.out].eps.push_back(f.in); acc.out = f.out; };
        for (int i = 0; i < lo; ++i) append(build(body));
        if (hi < 0) {  // X*
            int hub = n.add();
            n.st[acc.out].eps.push_back(hub);
            Frag f = build(body);
            n.st[hub].eps.push_back(f.in);
            n.st[f.out].eps.push_back(hub);
            acc.out = hub;
        } else if (hi > lo) {  // (X(X(...)?)?)?
            int end = n.add();
            n.st[acc.out].eps.push_back(end);
            int cur = acc.out;
            for (int i = lo; i < hi; ++i) {
                Frag f = build(body);
                n.st[cur].eps.push_back(f.in);
                n.st[f.out].eps.push_back(end);
                cur = f.out;
            }
            acc.out = end;
        }
        return acc;
    }

    Frag build(const Ast& a) {
        switch (a.kind) {
            case AstKind::Empty: return empty_frag();
            case AstKind::Set: return set_frag(a.set);
            case AstKind::Group: return build(*a.kids[0]);
            case AstKind::Cat: {
                Frag acc = empty_frag();
                for (size_t i = 0; i < a.kids.size(); ++i) {
                    Frag f = build(*a.kids[rev ? a.kids.size() - 1 - i : i]);
                    n.st[acc.out].eps.push_back(f.in);
                    acc.out = f.out;
                }
                return acc;
            }
            case AstKind::Alt: {
                int in = n.add(), out = n.add();
                for (const auto& k : a.kids) { Frag f = build(*k); n.st[in].eps.push_back(f.in); n.st[f.out].eps.push_back(out); }
                return {in, out};
            }
            case AstKind::Repeat: return repeat(*a.kids[0], a.min, a.max);
            case AstKind::Assert: {
                if (a.assert_kind == AssertKind::BolMulti || a.assert_kind == AssertKind::EolMulti)
                    throw std::logic_error("multiline anchors are not DFA-representable here");
                if (a.assert_kind == AssertKind::Eol) n.uses_eol = true;
                int in = n.add(), out = n.add();
                AssertKind k = a.assert_kind;
                if (rev) k = k == AssertKind::Begin ? AssertKind::End : k == AssertKind::End ? AssertKind::Begin : k;   // \b, \B are symmetric
                n.st[in].asserts.push_back({k, out});
                return {in, out};
            }
            default: throw std::logic_error("non-regular node in DFA build");
        }
    }
};

// Unanchored search: a prefix loop in front of the pattern that skips whole UTF-8 code points, so that a
// match (and the zero-width assertions at its start) is only ever tried at a code point boundary -- the
// positions java.util.regex knows. A byte-wise loop would let `\B` or an empty match succeed between the
// two bytes of "é". Malformed sequences are skipped byte-wise.
void add_search_loop(Nfa& nfa, int start) {
    const int c1 = nfa.add(), c2 = nfa.add(), c3 = nfa.add();      // 1, 2, 3 continuation bytes still to come
    auto lead_edges = [&](int from) {
        nfa.st[from].edges.push_back({0x00, 0x7F, start});
        nfa.st[from].edges.push_back({0xC0, 0xC1, start});
        nfa.st[from].edges.push_back({0xC2, 0xDF, c1});
        nfa.st[from].edges.push_back({0xE0, 0xEF, c2});
        nfa.st[from].edges.push_back({0xF0, 0xF4, c3});
        nfa.st[from].edges.push_back({0xF5, 0xFF, start});
    };
    lead_edges(start);
    nfa.st[start].edges.push_back({0x80, 0xBF, start});             // stray continuation byte
    lead_edges(c1); lead_edges(c2); lead_edges(c3);                 // truncated sequence: resynchronise
    nfa.st[c1].edges.push_back({0x80, 0xBF, start});
    nfa.st[c2].edges.push_back({0x80, 0xBF, c1});
    nfa.st[c3].edges.push_back({0x80, 0xBF, c2});
}

struct Ctx {  // left context carried in a DFA state
    bool prev_word = false, at_begin = false, at_tpos = false;
    int bits() const { return (prev_word ? 1 : 0) | (at_begin ? 2 : 0) | (at_tpos ? 4 : 0); }
};

struct Subset {
    const Nfa& n;
    std::vector<int> mark;
    int stamp = 0;
    explicit Subset(const Nfa& nfa) : n(nfa), mark(nfa.st.size(), 0) {}

    // epsilon closure under the assertions that hold here; returns true if accept is reached
    bool closure(const std::vector<int>& kernel, const Ctx& c, bool next_is_eot, bool next_word, std::vector<int>& out) {
        ++stamp;
        out.clear();
        std::vector<int> stack(kernel.begin(), kernel.end());
        for (int s : kernel) mark[s] = stamp;
        bool acc = false;
        const bool wordb = c.prev_word != next_word;
        while (!stack.empty()) {
            int s = stack.back();
            stack.pop_back();
            out.push_back(s);
            if (s == n.accept) acc = true;
            const NState& st = n.st[s];
            for (int t : st.eps) if (mark[t] != stamp) { mark[t] = stamp; stack.push_back(t); }
            for (const auto& a : st.asserts) {
                bool ok = false;
                switch (a.k) {
                    case AssertKind::Begin: ok = c.at_begin; break;
                    case AssertKind::End: ok = next_is_eot; break;
                    case AssertKind::Eol: ok = next_is_eot || c.at_tpos; break;
                    case AssertKind::WordB: ok = wordb; break;
                    case AssertKind::NotWordB: ok = !wordb; break;
                    default: break;
                }
                if (ok && mark[a.to] != stamp) { mark[a.to] = stamp; stack.push_back(a.to); }
            }
        }
        return acc;
    }
};

}  // namespace

namespace {
// mode 0: unanchored search from the line start. mode 1: a match must START at the walk's first byte;
// mode 2: the reversed language, a match must END where the (backward) walk starts. Modes 1/2 get three
// start states for the context on the other side of the starting point: line edge, non-word byte, word byte.
bool build_dfa_impl(const Ast& root, Dfa& out, const DfaLimits& lim, int mode) {
    out = Dfa();
    Nfa nfa(lim.max_nfa);
    int start;
    try {
        Builder b(nfa, mode == 2);
        start = nfa.add();
        if (mode == 0) add_search_loop(nfa, start);
        Frag f = b.build(root);
        nfa.st[start].eps.push_back(f.in);
        nfa.accept = f.out;
    } catch (const std::length_error&) {
        return false;
    }
    if (mode == 2 && nfa.uses_eol) return false;      // `$` looks at what follows the match: not walkable backwards

    // byte classes from edge boundaries (+ word-byte boundaries for \b)
    bool cut[257];
    memset(cut, 0, sizeof cut);
    cut[0] = cut[256] = true;
    for (const auto& s : nfa.st)
        for (const auto& e : s.edges) { cut[e.lo] = true; cut[(int)e.hi + 1] = true; }
    for (int b = 0; b < 256; ++b)
        if (b > 0 && word_byte(b) != word_byte(b - 1)) cut[b] = true;
    std::vector<int> rep;  // representative byte per class
    uint8_t cls_of[256];
    for (int b = 0; b < 256; ++b) {
        if (cut[b]) rep.push_back(b);
        cls_of[b] = (uint8_t)(rep.size() - 1);
    }
    const int nb = (int)rep.size();
    const int SYM_TPOS = nb, SYM_EOT = nb + 1, nsym = nb + 2;

    Subset sub(nfa);
    std::map<std::pair<std::vector<int>, int>, int> ids;
    std::vector<std::pair<std::vector<int>, Ctx>> states;
    std::vector<std::vector<int>> trans;  // -1 DEAD, -2 MATCH, else state id
    auto intern = [&](std::vector<int> k, Ctx c) -> int {
        std::sort(k.begin(), k.end());
        k.erase(std::unique(k.begin(), k.end()), k.end());
        if (!nfa.uses_eol) c.at_tpos = false;
        auto key = std::make_pair(k, c.bits());
        auto it = ids.find(key);
        if (it != ids.end()) return it->second;
        int id = (int)states.size();
        ids.emplace(std::move(key), id);
        states.emplace_back(std::move(k), c);
        trans.emplace_back(nsym, -1);
        return id;
    };
    Ctx c0;
    c0.at_begin = true;
    intern({start}, c0);
    const int n_starts = mode == 0 ? 1 : 3;
    if (mode != 0) {
        Ctx c1, c2;
        c2.prev_word = true;
        const int s1 = intern({start}, c1), s2 = intern({start}, c2);
        if (s1 != 1 || s2 != 2) throw std::logic_error("anchored start states");
    }
    std::vector<int> clo, nxt;
    for (size_t si = 0; si < states.size(); ++si) {
        if (states.size() > lim.max_raw_states) return false;
        std::vector<int> kernel = states[si].first;
        Ctx c = states[si].second;
        for (int cl = 0; cl < nb; ++cl) {
            int b = rep[cl];
            if (sub.closure(kernel, c, false, word_byte(b), clo)) { trans[si][cl] = -2; continue; }
            nxt.clear();
            for (int s : clo)
                for (const auto& e : nfa.st[s].edges)
                    if (b >= e.lo && b <= e.hi) nxt.push_back(e.to);
            if (nxt.empty()) { trans[si][cl] = -1; continue; }
            Ctx c2;
            c2.prev_word = word_byte(b);
            trans[si][cl] = intern(nxt, c2);
        }
        trans[si][SYM_EOT] = sub.closure(kernel, c, true, false, clo) ? -2 : -1;
        if (nfa.uses_eol && !c.at_tpos) { Ctx c2 = c; c2.at_tpos = true; trans[si][SYM_TPOS] = intern(kernel, c2); }
        else trans[si][SYM_TPOS] = (int)si;
    }
    const int ns = (int)states.size();

    // states that cannot reach MATCH become DEAD (gives anchored patterns an early exit)
    std::vector<char> live(ns, 0);
    bool changed = true;
    while (changed) {
        changed = false;
        for (int s = 0; s < ns; ++s) {
            if (live[s]) continue;
            for (int x = 0; x < nsym; ++x) {
                int t = trans[s][x];
                if (t == -2 || (t >= 0 && live[t])) { live[s] = 1; changed = true; break; }
            }
        }
    }
    // Moore minimisation over {DEAD, MATCH, live states}
    std::vector<int> part(ns, 0);  // block id of each live state; dead states are -1
    for (int s = 0; s < ns; ++s) part[s] = live[s] ? 0 : -1;
    int nblocks = 1;
    for (;;) {
        std::map<std::vector<int>, int> sig2b;
        std::vector<int> np(ns, -1);
        for (int s = 0; s < ns; ++s) {
            if (!live[s]) continue;
            std::vector<int> sig;
            sig.reserve(nsym + 1);
            sig.push_back(part[s]);
            for (int x = 0; x < nsym; ++x) {
                int t = trans[s][x];
                sig.push_back(t == -2 ? -2 : (t < 0 || !live[t]) ? -1 : part[t]);
            }
            auto it = sig2b.find(sig);
            if (it == sig2b.end()) it = sig2b.emplace(std::move(sig), (int)sig2b.size()).first;
            np[s] = it->second;
        }
        int nb2 = (int)sig2b.size();
        part.swap(np);
        if (nb2 == nblocks) break;
        nblocks = nb2;
    }
    bool any_start_live = false;
    for (int k = 0; k < n_starts; ++k) any_start_live |= live[k] != 0;
    if (!any_start_live) {  // pattern can never match
        out.n_classes = 3;
        memset(out.classmap, 0, 256);
        out.trans.assign(2 * 3, DFA_DEAD);
        out.trans[1 * 3 + 0] = out.trans[1 * 3 + 1] = out.trans[1 * 3 + 2] = DFA_MATCH;
        out.start = DFA_DEAD;
        out.start_ctx[0] = out.start_ctx[1] = out.start_ctx[2] = DFA_DEAD;
        return true;
    }
    // block table
    std::vector<std::vector<int>> bt(nblocks, std::vector<int>(nsym, 0));
    std::vector<char> seen(nblocks, 0);
    for (int s = 0; s < ns; ++s) {
        if (!live[s] || seen[part[s]]) continue;
        seen[part[s]] = 1;
        for (int x = 0; x < nsym; ++x) {
            int t = trans[s][x];
            bt[part[s]][x] = t == -2 ? DFA_MATCH : (t < 0 || !live[t]) ? DFA_DEAD : part[t] + 2;
        }
    }
    // merge byte classes with identical columns
    std::map<std::vector<int>, int> col2c;
    std::vector<int> cmap(nb);
    for (int cl = 0; cl < nb; ++cl) {
        std::vector<int> col(nblocks);
        for (int b = 0; b < nblocks; ++b) col[b] = bt[b][cl];
        auto it = col2c.find(col);
        if (it == col2c.end()) it = col2c.emplace(std::move(col), (int)col2c.size()).first;
        cmap[cl] = it->second;
    }
    const int nb2 = (int)col2c.size();
    const int nsym2 = nb2 + 2;
    if (nblocks + 2 > 65535 || (uint32_t)nblocks + 2 > lim.max_states) return false;
    out.n_classes = (uint16_t)nsym2;
    for (int b = 0; b < 256; ++b) out.classmap[b] = (uint8_t)cmap[cls_of[b]];
    out.trans.assign((size_t)(nblocks + 2) * nsym2, DFA_DEAD);
    for (int x = 0; x < nsym2; ++x) out.trans[(size_t)DFA_MATCH * nsym2 + x] = DFA_MATCH;
    for (int b = 0; b < nblocks; ++b) {
        uint16_t* row = &out.trans[(size_t)(b + 2) * nsym2];
        for (int cl = 0; cl < nb; ++cl) row[cmap[cl]] = (uint16_t)bt[b][cl];
        row[nb2] = (uint16_t)bt[b][SYM_TPOS];
        row[nb2 + 1] = (uint16_t)bt[b][SYM_EOT];
    }
    out.start = live[0] ? (uint16_t)(part[0] + 2) : DFA_DEAD;
    for (int k = 0; k < 3; ++k) out.start_ctx[k] = k < n_starts ? (live[k] ? (uint16_t)(part[k] + 2) : DFA_DEAD) : out.start;
    return true;
}
}  // namespace

bool build_dfa(const Ast& root, Dfa& out, const DfaLimits& lim) { return build_dfa_impl(root, out, lim, 0); }
bool build_dfa_anchored(const Ast& root, bool reverse, Dfa& out, const DfaLimits& lim) { return build_dfa_impl(root, out, lim, reverse ? 2 : 1); }

bool build_multi_dfa(const std::vector<const Ast*>& roots, MultiDfa& out, uint32_t max_states, uint32_t max_nfa) {
    out = MultiDfa();
    const int K = (int)roots.size();
    if (K == 0 || K > 32) return false;
    Nfa nfa(max_nfa);
    std::vector<int> starts(K), accepts(K), owner;
    try {
        Builder b(nfa);
        for (int k = 0; k < K; ++k) {
            const size_t first = nfa.st.size();
            starts[k] = nfa.add();
            add_search_loop(nfa, starts[k]);
            Frag f = b.build(*roots[k]);
            nfa.st[starts[k]].eps.push_back(f.in);
            accepts[k] = f.out;
            owner.resize(nfa.st.size(), k);
            (void)first;
        }
    } catch (const std::length_error&) {
        return false;
    }
    std::vector<int> accept_of(nfa.st.size(), -1);
    for (int k = 0; k < K; ++k) accept_of[accepts[k]] = k;

    bool cut[257];
    memset(cut, 0, sizeof cut);
    cut[0] = cut[256] = true;
    for (const auto& s : nfa.st)
        for (const auto& e : s.edges) { cut[e.lo] = true; cut[(int)e.hi + 1] = true; }
    for (int b = 1; b < 256; ++b)
        if (word_byte(b) != word_byte(b - 1)) cut[b] = true;
    std::vector<int> rep;
    uint8_t cls_of[256];
    for (int b = 0; b < 256; ++b) {
        if (cut[b]) rep.push_back(b);
        cls_of[b] = (uint8_t)(rep.size() - 1);
    }
    const int nb = (int)rep.size();
    const int SYM_TPOS = nb, SYM_EOT = nb + 1, nsym = nb + 2;

    Subset sub(nfa);
    struct Key {
        std::vector<int> k; int ctx; uint32_t mask;
        bool operator<(const Key& o) const { return mask != o.mask ? mask < o.mask : ctx != o.ctx ? ctx < o.ctx : k < o.k; }
    };
    std::map<Key, int> ids;
    struct St { std::vector<int> k; Ctx c; uint32_t mask; };
    std::vector<St> states;
    std::vector<std::vector<int>> trans;
    auto intern = [&](std::vector<int> k, Ctx c, uint32_t mask) -> int {
        std::sort(k.begin(), k.end());
        k.erase(std::unique(k.begin(), k.end()), k.end());
        if (!nfa.uses_eol) c.at_tpos = false;
        if (k.empty()) c = Ctx();
        Key key{k, c.bits(), mask};
        auto it = ids.find(key);
        if (it != ids.end()) return it->second;
        int id = (int)states.size();
        ids.emplace(std::move(key), id);
        states.push_back(St{std::move(k), c, mask});
        trans.emplace_back(nsym, 0);
        return id;
    };
    Ctx c0;
    c0.at_begin = true;
    intern(std::vector<int>(starts.begin(), starts.end()), c0, 0);
    std::vector<int> clo, nxt;
    auto settle = [&](uint32_t mask, const std::vector<int>& closure_set, uint32_t& newmask) {
        newmask = mask;
        for (int s : closure_set) if (accept_of[s] >= 0) newmask |= 1u << accept_of[s];
    };
    for (size_t si = 0; si < states.size(); ++si) {
        if (states.size() > max_states * 8u + 64u) return false;
        const std::vector<int> kernel = states[si].k;
        const Ctx c = states[si].c;
        const uint32_t mask = states[si].mask;
        for (int cl = 0; cl < nb; ++cl) {
            const int b = rep[cl];
            sub.closure(kernel, c, false, word_byte(b), clo);
            uint32_t nm;
            settle(mask, clo, nm);
            nxt.clear();
            for (int s : clo) {
                if ((nm >> owner[s]) & 1u) continue;      // that regex is decided: drop its threads
                for (const auto& e : nfa.st[s].edges)
                    if (b >= e.lo && b <= e.hi) nxt.push_back(e.to);
            }
            Ctx c2;
            c2.prev_word = word_byte(b);
            trans[si][cl] = intern(nxt, c2, nm);
        }
        {
            sub.closure(kernel, c, true, false, clo);
            uint32_t nm;
            settle(mask, clo, nm);
            trans[si][SYM_EOT] = intern({}, Ctx(), nm);
        }
        if (nfa.uses_eol && !c.at_tpos && !kernel.empty()) { Ctx c2 = c; c2.at_tpos = true; trans[si][SYM_TPOS] = intern(kernel, c2, mask); }
        else trans[si][SYM_TPOS] = (int)si;
    }
    const int ns = (int)states.size();
    // Moore minimisation; the observable of a state is its mask
    std::vector<int> part(ns);
    {
        std::map<uint32_t, int> m2b;
        for (int s = 0; s < ns; ++s) {
            auto it = m2b.find(states[s].mask);
            if (it == m2b.end()) it = m2b.emplace(states[s].mask, (int)m2b.size()).first;
            part[s] = it->second;
        }
    }
    int nblocks = 0;
    for (int s = 0; s < ns; ++s) nblocks = std::max(nblocks, part[s] + 1);
    for (;;) {
        std::map<std::vector<int>, int> sig2b;
        std::vector<int> np(ns);
        for (int s = 0; s < ns; ++s) {
            std::vector<int> sig;
            sig.reserve(nsym + 1);
            sig.push_back(part[s]);
            for (int x = 0; x < nsym; ++x) sig.push_back(part[trans[s][x]]);
            auto it = sig2b.find(sig);
            if (it == sig2b.end()) it = sig2b.emplace(std::move(sig), (int)sig2b.size()).first;
            np[s] = it->second;
        }
        const int nb2 = (int)sig2b.size();
        part.swap(np);
        if (nb2 == nblocks) break;
        nblocks = nb2;
    }
    if ((uint32_t)nblocks > max_states || nblocks > 65535) return false;
    std::vector<std::vector<int>> bt(nblocks, std::vector<int>(nsym, 0));
    out.state_mask.assign(nblocks, 0);
    for (int s = 0; s < ns; ++s) {
        out.state_mask[part[s]] = states[s].mask;
        for (int x = 0; x < nsym; ++x) bt[part[s]][x] = part[trans[s][x]];
    }
    std::map<std::vector<int>, int> col2c;
    std::vector<int> cmap(nb);
    for (int cl = 0; cl < nb; ++cl) {
        std::vector<int> col(nblocks);
        for (int b = 0; b < nblocks; ++b) col[b] = bt[b][cl];
        auto it = col2c.find(col);
        if (it == col2c.end()) it = col2c.emplace(std::move(col), (int)col2c.size()).first;
        cmap[cl] = it->second;
    }
    const int nb2 = (int)col2c.size(), nsym2 = nb2 + 2;
    out.n_classes = (uint16_t)nsym2;
    for (int b = 0; b < 256; ++b) out.classmap[b] = (uint8_t)cmap[cls_of[b]];
    out.trans.assign((size_t)nblocks * nsym2, 0);
    out.state_final.assign(nblocks, 0);
    for (int b = 0; b < nblocks; ++b) {
        uint16_t* row = &out.trans[(size_t)b * nsym2];
        for (int cl = 0; cl < nb; ++cl) row[cmap[cl]] = (uint16_t)bt[b][cl];
        row[nb2] = (uint16_t)bt[b][SYM_TPOS];
        row[nb2 + 1] = (uint16_t)bt[b][SYM_EOT];
        bool fin = true;
        for (int x = 0; x < nsym2 && fin; ++x) fin = row[x] == b;
        out.state_final[b] = fin ? 1 : 0;
    }
    out.start = (uint16_t)part[0];
    return true;
}

uint32_t MultiDfa::run(const uint8_t* s, size_t n) const {
    if (!n_classes) return 0;
    long tpos = -1;
    if (n >= 1 && s[n - 1] == '\r') tpos = (long)n - 1;
    else if (n >= 2 && s[n - 2] == 0xC2 && s[n - 1] == 0x85) tpos = (long)n - 2;
    else if (n >= 3 && s[n - 3] == 0xE2 && s[n - 2] == 0x80 && (s[n - 1] == 0xA8 || s[n - 1] == 0xA9)) tpos = (long)n - 3;
    const unsigned nc = n_classes;
    unsigned st = start;
    for (size_t i = 0; i < n && !state_final[st]; ++i) {
        if ((long)i == tpos) st = trans[st * nc + nc - 2];
        st = trans[st * nc + classmap[s[i]]];
    }
    st = trans[st * nc + nc - 1];
    return state_mask[st];
}

// Same walk as the device verifier (kernels.cu: dfa_walk).
bool Dfa::match(const uint8_t* s, size_t n) const {
    if (!n_classes) return false;
    long tpos = -1;
    if (n >= 1 && s[n - 1] == '\r') tpos = (long)n - 1;
    else if (n >= 2 && s[n - 2] == 0xC2 && s[n - 1] == 0x85) tpos = (long)n - 2;
    else if (n >= 3 && s[n - 3] == 0xE2 && s[n - 2] == 0x80 && (s[n - 1] == 0xA8 || s[n - 1] == 0xA9)) tpos = (long)n - 3;
    const unsigned nc = n_classes;
    unsigned st = start;
    for (size_t i = 0; i < n && st > DFA_MATCH; ++i) {
        if ((long)i == tpos) st = trans[st * nc + nc - 2];
        st = trans[st * nc + classmap[s[i]]];
    }
    if (st > DFA_MATCH) st = trans[st * nc + nc - 1];
    return st == DFA_MATCH;
}

thread_local unsigned long long g_anchor_steps = 0;      // tuning aid: transitions taken by match_from / match_ending_at
static inline bool word_byte_pub(int b) { return (b >= '0' && b <= '9') || (b >= 'A' && b <= 'Z') || b == '_' || (b >= 'a' && b <= 'z'); }

bool Dfa::match_from(const uint8_t* s, size_t n, size_t pos) const {
    if (!n_classes || pos > n) return false;
    long tpos = -1;
    if (n >= 1 && s[n - 1] == '\r') tpos = (long)n - 1;
    else if (n >= 2 && s[n - 2] == 0xC2 && s[n - 1] == 0x85) tpos = (long)n - 2;
    else if (n >= 3 && s[n - 3] == 0xE2 && s[n - 2] == 0x80 && (s[n - 1] == 0xA8 || s[n - 1] == 0xA9)) tpos = (long)n - 3;
    const unsigned nc = n_classes;
    unsigned st = start_ctx[pos == 0 ? 0 : word_byte_pub(s[pos - 1]) ? 2 : 1];
    for (size_t i = pos; i < n && st > DFA_MATCH; ++i) {
        if ((long)i == tpos) st = trans[st * nc + nc - 2];
        st = trans[st * nc + classmap[s[i]]];
        ++g_anchor_steps;
    }
    if (st > DFA_MATCH) st = trans[st * nc + nc - 1];
    return st == DFA_MATCH;
}
bool Dfa::match_ending_at(const uint8_t* s, size_t n, size_t pos) const {
    if (!n_classes || pos > n) return false;
    const unsigned nc = n_classes;
    unsigned st = start_ctx[pos == n ? 0 : word_byte_pub(s[pos]) ? 2 : 1];
    for (size_t i = pos; i > 0 && st > DFA_MATCH; --i) { st = trans[st * nc + classmap[s[i - 1]]]; ++g_anchor_steps; }
    if (st > DFA_MATCH) st = trans[st * nc + nc - 1];
    return st == DFA_MATCH;
}

bool literals_cover(const Dfa& d, bool reverse, const std::vector<Literal>& lits) {
    if (lits.empty() || !d.n_classes) return false;
    // trie of the literals (reversed for the backward automaton); case-insensitive letters get both cases
    struct Node { std::map<int, int> kid; bool end = false; };
    std::vector<Node> trie(1);
    for (const auto& l : lits) {
        if (l.bytes.empty()) return false;
        std::vector<int> cur{0};
        for (size_t i = 0; i < l.bytes.size(); ++i) {
            const unsigned char ch = (unsigned char)l.bytes[reverse ? l.bytes.size() - 1 - i : i];
            std::vector<int> alts{ch};
            if (l.icase && ch >= 'a' && ch <= 'z') alts.push_back(ch - 32);
            std::vector<int> nxt;
            for (int node : cur)
                for (int b : alts) {
                    auto it = trie[node].kid.find(b);
                    int to;
                    if (it == trie[node].kid.end()) { to = (int)trie.size(); trie[node].kid[b] = to; trie.emplace_back(); }
                    else to = it->second;
                    nxt.push_back(to);
                }
            cur.swap(nxt);
            if (trie.size() > 200000) return false;
        }
        for (int node : cur) trie[node].end = true;
    }
    const unsigned nc = d.n_classes;
    std::map<std::pair<unsigned, int>, char> seen;
    // every way the automaton can stay alive from (state, trie node) must run through a complete literal
    // before a match can be reported
    std::vector<std::pair<unsigned, int>> stack;
    for (int k = 0; k < 3; ++k) if (d.start_ctx[k] > DFA_MATCH) stack.push_back({d.start_ctx[k], 0});
    for (int k = 0; k < 3; ++k) if (d.start_ctx[k] == DFA_MATCH) return false;
    size_t budget = 2000000;
    while (!stack.empty()) {
        auto cur = stack.back();
        stack.pop_back();
        if (seen.count(cur)) continue;
        seen[cur] = 1;
        const unsigned st = cur.first;
        const Node& nd = trie[cur.second];
        if (nd.end) continue;
        if (d.trans[st * nc + nc - 1] == DFA_MATCH) return false;             // match at end of line before a full literal
        const unsigned tp = d.trans[st * nc + nc - 2];                           // the `$` position announcement consumes no text
        if (tp == DFA_MATCH) return false;
        if (tp > DFA_MATCH && tp != st) stack.push_back({tp, cur.second});
        for (int b = 0; b < 256; ++b) {
            const unsigned t = d.trans[st * nc + d.classmap[b]];
            if (t == DFA_DEAD) continue;
            if (t == DFA_MATCH) return false;                                    // a match completed before this byte
            auto it = nd.kid.find(b);
            if (it == nd.kid.end()) return false;                                // alive on a byte no literal continues with
            stack.push_back({t, it->second});
            if (--budget == 0) return false;
        }
    }
    return true;
}


int choose_anchor(const Ast& root, const std::vector<Literal>& lits, Dfa& out, const DfaLimits& lim) {
    if (lits.empty()) return 0;
    for (int mode = 1; mode <= 2; ++mode) {
        Dfa d;
        if (!build_dfa_anchored(root, mode == 2, d, lim)) continue;
        if (literals_cover(d, mode == 2, lits)) { out = std::move(d); return mode; }
    }
    return 0;
}

}  // namespace lpx
