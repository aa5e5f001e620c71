// Host-compiler introspection for the CPU-only test tier, built into its own library
// (liblogparser_b200_debug.so, host code only; the product library does not contain it)
// (`pytest -m "not gpu"`): lets the tests diff the exact tables the device walks against the
// oracle without a GPU. Never called by lp_analyze*; it is not a CPU fallback of the product path.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>

#include "rx.h"
#include "plan.h"
#include "vm.h"

namespace lpx { extern thread_local unsigned long long g_anchor_steps; }
using namespace lpx;

namespace {
struct DebugRx {
    ParsedRegex parsed;
    bool has_dfa = false;
    Dfa dfa;
    bool has_vm = false;
    VmProgram vm;
    std::vector<Literal> lits;
    int anchor = 0;          // 0 none, 1 forward, 2 backward
    Dfa adfa;
    std::vector<LocalLiteral> local;      // non-empty: the regex is decided by boundary tests at its literal occurrences
    VmProgram span_vm;       // every regex has a bytecode program: spans come from the backtracking engine
};
void put_err(char* err, size_t n, const std::string& m) {
    if (!err || !n) return;
    size_t k = std::min(n - 1, m.size());
    memcpy(err, m.data(), k);
    err[k] = 0;
}
}  // namespace

extern "C" {

// kind: 0 ok, 1 PatternSyntaxException, 2 unsupported construct, 3 internal
void* lp_debug_rx_compile(const char* regex, int flags, int force_vm, char* err, size_t errlen, int* kind) {
    if (kind) *kind = 0;
    try {
        DebugRx* d = new DebugRx();
        try {
            d->parsed = parse_java_regex(regex, flags);
            const size_t min_lit = (force_vm >> 8) ? (size_t)(force_vm >> 8) : 4;   // tests lower it to exercise short literals
            force_vm &= 1;
            if (d->parsed.regular && !force_vm) d->has_dfa = build_dfa(*d->parsed.root, d->dfa);
            if (!d->has_dfa) { build_vm(d->parsed, d->vm); d->has_vm = true; }
            build_vm(d->parsed, d->span_vm);
            if (d->has_dfa && !local_verify_form(*d->parsed.root, d->local, min_lit)) d->local.clear();
            d->lits = required_literals(*d->parsed.root, min_lit);
            if (d->has_dfa) d->anchor = choose_anchor(*d->parsed.root, d->lits, d->adfa);
        } catch (...) { delete d; throw; }
        return d;
    } catch (const RegexSyntaxError& e) {
        if (kind) *kind = 1;
        put_err(err, errlen, std::string(e.what()) + " near index " + std::to_string(e.index));
    } catch (const RegexUnsupported& e) {
        if (kind) *kind = 2;
        put_err(err, errlen, std::string("unsupported: ") + e.what());
    } catch (const std::exception& e) {
        if (kind) *kind = 3;
        put_err(err, errlen, e.what());
    }
    return nullptr;
}
void lp_debug_rx_free(void* h) { delete static_cast<DebugRx*>(h); }
int lp_debug_rx_steps(void* h, const uint8_t* s, size_t n) { return lpx::vm_steps_host(static_cast<DebugRx*>(h)->span_vm, s, n); }
int lp_debug_rx_is_dfa(void* h) { return static_cast<DebugRx*>(h)->has_dfa ? 1 : 0; }
uint32_t lp_debug_rx_states(void* h) {
    DebugRx* d = static_cast<DebugRx*>(h);
    return d->has_dfa ? d->dfa.n_states() : (uint32_t)d->vm.code.size();
}
// Walks the same tables with the same loop as the device verifier. Returns 1/0, or <0 for a
// backtracking-budget overflow (the device reports the same condition as an error).
int lp_debug_rx_match(void* h, const uint8_t* s, size_t n) {
    DebugRx* d = static_cast<DebugRx*>(h);
    if (d->has_dfa) return d->dfa.match(s, n) ? 1 : 0;
    return vm_match_host(d->vm, s, n);
}
unsigned long long lp_debug_anchor_steps(int reset) { unsigned long long v = lpx::g_anchor_steps; if (reset) lpx::g_anchor_steps = 0; return v; }
// Byte offsets of the first find() (what k_spans computes); returns 1/0, <0 on a budget overflow.
int lp_debug_rx_span(void* h, const uint8_t* s, size_t n, uint32_t* start, uint32_t* end) {
    uint32_t span[2] = {0, 0};
    const int r = vm_span_host(static_cast<DebugRx*>(h)->span_vm, s, n, span);
    if (r > 0) { *start = span[0]; *end = span[1]; }
    return r;
}
// What k_resolve's boundary class decides: some occurrence of one of the literals whose neighbours satisfy
// the \\b conditions. -2 when the regex is not of that form.
int lp_debug_rx_match_local(void* h, const uint8_t* s, size_t n) {
    DebugRx* d = static_cast<DebugRx*>(h);
    if (d->local.empty()) return -2;
    auto word = [](uint8_t b) { return (b >= '0' && b <= '9') || (b >= 'A' && b <= 'Z') || b == '_' || (b >= 'a' && b <= 'z'); };
    for (const auto& ll : d->local) {
        const size_t m = ll.lit.bytes.size();
        for (size_t p = 0; p + m <= n; ++p) {
            bool ok = true;
            for (size_t j = 0; j < m && ok; ++j) {
                uint8_t b = s[p + j];
                if (ll.lit.icase && b >= 'A' && b <= 'Z') b += 32;
                ok = b == (uint8_t)ll.lit.bytes[j];
            }
            if (!ok) continue;
            if (ll.left_boundary && ((p > 0 && word(s[p - 1])) == word(s[p]))) continue;
            if (ll.right_boundary && (word(s[p + m - 1]) == (p + m < n && word(s[p + m])))) continue;
            return 1;
        }
    }
    return 0;
}
int lp_debug_rx_anchor(void* h) { return static_cast<DebugRx*>(h)->anchor; }
// What k_confirm decides for an anchored regex: some occurrence of one of its literals at which the
// anchored automaton accepts. -2 when the regex is not verified that way.
int lp_debug_rx_match_anchored(void* h, const uint8_t* s, size_t n) {
    DebugRx* d = static_cast<DebugRx*>(h);
    if (!d->anchor) return -2;
    int hit = 0;       // every occurrence is walked, as on the device
    for (const auto& l : d->lits) {
        const size_t m = l.bytes.size();
        for (size_t p = 0; p + m <= n; ++p) {
            bool ok = true;
            for (size_t j = 0; j < m && ok; ++j) {
                uint8_t b = s[p + j];
                if (l.icase && b >= 'A' && b <= 'Z') b += 32;
                ok = b == (uint8_t)l.bytes[j];
            }
            if (!ok) continue;
            if (d->anchor == 1 ? d->adfa.match_from(s, n, p) : d->adfa.match_ending_at(s, n, p + m)) hit = 1;
        }
    }
    return hit;
}
size_t lp_debug_rx_literals(void* h, char* buf, size_t cap) {
    DebugRx* d = static_cast<DebugRx*>(h);
    std::string out;
    for (const auto& l : d->lits) {
        out += l.icase ? "i:" : "s:";
        out += l.bytes;
        out.push_back('\n');
    }
    if (buf && cap) {
        size_t k = std::min(cap - 1, out.size());
        memcpy(buf, out.data(), k);
        buf[k] = 0;
    }
    return out.size();
}

// Builds one scan group (product DFA) from `n` regexes and walks `text` through its serialised
// tables exactly as k_dfa_scan does. Returns the hit mask (bit k = regex k), or 0xFFFFFFFF when the
// group cannot be built.
// The first-byte skip set of the scan group built from `n` regexes (k_dfa_scan does not walk a non-empty line whose
// first byte is in it). Returns 0, or -1 when the group cannot be built.
int lp_debug_group_skip(const char* const* regexes, const int* flags, size_t n, uint32_t* out8) {
    try {
        std::vector<ParsedRegex> pr;
        std::vector<const Ast*> roots;
        for (size_t i = 0; i < n; ++i) pr.push_back(parse_java_regex(regexes[i], flags ? flags[i] : 0));
        for (auto& p : pr) { if (!p.regular) return -1; roots.push_back(p.root.get()); }
        MultiDfa d;
        if (!build_multi_dfa(roots, d, 8192)) return -1;
        std::vector<uint32_t> rx(n), blob;
        for (size_t i = 0; i < n; ++i) rx[i] = (uint32_t)i + 4;
        serialise_group(rx, d, blob);
        group_skip_set(blob, out8);
        return 0;
    } catch (const std::exception&) {
        return -1;
    }
}
uint32_t lp_debug_group_match(const char* const* regexes, const int* flags, size_t n, const uint8_t* text, size_t len) {
    try {
        std::vector<ParsedRegex> pr;
        std::vector<const Ast*> roots;
        for (size_t i = 0; i < n; ++i) pr.push_back(parse_java_regex(regexes[i], flags ? flags[i] : 0));
        for (auto& p : pr) { if (!p.regular) return 0xFFFFFFFFu; roots.push_back(p.root.get()); }
        MultiDfa d;
        if (!build_multi_dfa(roots, d, 8192)) return 0xFFFFFFFFu;
        std::vector<uint32_t> rx(n), blob;
        for (size_t i = 0; i < n; ++i) rx[i] = (uint32_t)i + 4;
        serialise_group(rx, d, blob);
        const uint32_t a = run_group_blob(blob, text, len), b = d.run(text, len);
        return a == b ? a : 0xFFFFFFFEu;
    } catch (const std::exception&) {
        return 0xFFFFFFFFu;
    }
}

// CPU-only introspection of the plan lp_load_patterns would build (tests / tuning; no GPU needed):
// one text line per literal: "<gram hex>\t<icase>\t<literal>", then "#..." summary lines.
size_t lp_debug_plan_dump(const lp_pattern_desc* pats, size_t n, const uint8_t* sample, size_t sample_len, char* buf, size_t cap) {
    std::string out;
    try {
        lpx::Plan pl;
        std::vector<lpx::PlanPatternIn> in = lpx::plan_input_from_desc(pats, n, nullptr);
        lpx::build_plan(in, pl, nullptr);
        if (sample && sample_len) lpx::select_grams(pl, sample, sample_len);
        for (size_t l = 0; l < pl.lit_off.size(); ++l) {
            char g[16];
            const uint8_t* b = &pl.lit_bytes[pl.lit_off[l] + pl.lit_gram_off[l]];
            snprintf(g, sizeof g, "%02x%02x%02x%02x", b[0], b[1], b[2], b[3]);
            out += g; out += "\t"; out += pl.lit_icase[l] ? "1" : "0";
            out += ":";
            for (uint32_t j = pl.lit_rx_begin[l]; j < pl.lit_rx_begin[l + 1]; ++j) { out += std::to_string(pl.lit_rx[j]) + "/" + std::to_string((int)pl.lit_rx_mode[j]) + ","; }
            out += "\t";
            out.append(reinterpret_cast<const char*>(&pl.lit_bytes[pl.lit_off[l]]), pl.lit_len[l]);
            out += "\n";
        }
        out += "#regexes " + std::to_string(pl.n_rx) + " vm " + std::to_string(pl.n_vm) + " literals " + std::to_string(pl.lit_off.size()) +
               " groups " + std::to_string(pl.groups.size()) + " always_dfa " + std::to_string(pl.always_list.size()) +
               " always_vm " + std::to_string(pl.always_vm_list.size()) + "\n";
        for (const auto& g : pl.groups) {
            out += "#group states " + std::to_string(g.blob[0]) + " classes " + std::to_string(g.blob[1]) + " bytes " + std::to_string(g.blob.size() * 4) + " rx";
            for (uint32_t r : g.rx) out += " [" + pl.rx_src[r] + "]";
            out += "\n";
        }
        for (uint32_t r : pl.always_list) out += "#always_dfa [" + pl.rx_src[r] + "]\n";
        for (uint32_t r : pl.always_vm_list) out += "#always_vm [" + pl.rx_src[r] + "]\n";
    } catch (const std::exception& e) {
        out = std::string("!error ") + e.what() + "\n";
    }
    if (buf && cap) { size_t k = std::min(cap - 1, out.size()); memcpy(buf, out.data(), k); buf[k] = 0; }
    return out.size();
}


}  // extern "C"
