// Builds the device plan from the flattened pattern set (see plan.h).
#include "plan.h"

#include <algorithm>
#include <cstring>
#include <map>
#include <stdexcept>

namespace lpx {
namespace {

// Rough rarity of a byte in pod logs: rarer bytes make better filter grams.
int rarity(uint8_t b) {
    if (b >= 0x80) return 6;
    if (b >= 'A' && b <= 'Z') b = (uint8_t)(b + 32);
    if (b >= 'a' && b <= 'z') {
        static const char common[] = "etaoinsrlc";
        return strchr(common, (char)b) ? 1 : (strchr("dmupghf", (char)b) ? 2 : 3);
    }
    if (b >= '0' && b <= '9') return 2;
    if (b == ' ' || b == '.' || b == ':' || b == '-' || b == '_' || b == '/') return 1;
    return 4;
}

// every alternative starts with \A / non-MULTILINE ^: its automaton dies within a few bytes on most lines
bool begins_anchored(const Ast& a) {
    switch (a.kind) {
        case AstKind::Assert: return a.assert_kind == AssertKind::Begin;
        case AstKind::Group: return begins_anchored(*a.kids[0]);
        case AstKind::Cat: return !a.kids.empty() && begins_anchored(*a.kids[0]);
        case AstKind::Alt: { for (const auto& k : a.kids) if (!begins_anchored(*k)) return false; return !a.kids.empty(); }
        default: return false;
    }
}

bool blank(const std::string& s) {
    for (unsigned char c : s) if (c > ' ') return false;
    return true;
}

}  // namespace

void select_grams(Plan& pl, const uint8_t* sample, size_t sample_len);

std::vector<PlanPatternIn> plan_input_from_desc(const lp_pattern_desc* pats, size_t n, int64_t* bad_pattern) {
    std::vector<PlanPatternIn> in(n);
    for (size_t i = 0; i < n; ++i) {
        const lp_pattern_desc& d = pats[i];
        if (!d.primary_regex) { if (bad_pattern) *bad_pattern = (int64_t)i; throw std::invalid_argument("primary_regex is NULL"); }
        in[i].has_id = d.id != nullptr;
        if (d.id) in[i].id = d.id;
        in[i].primary = d.primary_regex;
        in[i].confidence = d.confidence; in[i].severity = d.severity_multiplier;
        in[i].has_context = d.has_context; in[i].before = d.lines_before; in[i].after = d.lines_after;
        for (uint32_t s = 0; s < d.n_secondaries; ++s) {
            if (!d.secondaries[s].regex) { if (bad_pattern) *bad_pattern = (int64_t)i; throw std::invalid_argument("secondary regex is NULL"); }
            in[i].secs.push_back({d.secondaries[s].regex, d.secondaries[s].weight, d.secondaries[s].proximity_window});
        }
        for (uint32_t s = 0; s < d.n_sequences; ++s) {
            PlanPatternIn::Seq q;
            q.bonus = d.sequences[s].bonus_multiplier;
            for (uint32_t e = 0; e < d.sequences[s].n_events; ++e) {
                if (!d.sequences[s].event_regex[e]) { if (bad_pattern) *bad_pattern = (int64_t)i; throw std::invalid_argument("sequence event regex is NULL"); }
                q.events.push_back(d.sequences[s].event_regex[e]);
            }
            in[i].seqs.push_back(std::move(q));
        }
    }
    return in;
}


void serialise_group(const std::vector<uint32_t>& rx, const MultiDfa& d, std::vector<uint32_t>& blob) {
    const uint32_t ns = d.n_states(), nc = d.n_classes, nbyte = nc - 2;
    std::vector<uint32_t> perm(ns), inv(ns);      // non-final states first
    uint32_t w = 0;
    for (uint32_t s2 = 0; s2 < ns; ++s2) if (!d.state_final[s2]) { perm[s2] = w; inv[w++] = s2; }
    const uint32_t first_final = w;
    for (uint32_t s2 = 0; s2 < ns; ++s2) if (d.state_final[s2]) { perm[s2] = w; inv[w++] = s2; }
    auto step = [&](uint32_t st, uint8_t byte) { return (uint32_t)d.trans[st * nc + d.classmap[byte]]; };
    // Quiescent pair: the state the automaton idles in after a boring non-word byte (qa) and
    // after a boring word byte (qb). On "boring" byte classes both idle states move to qa/qb
    // by word-ness alone, so the walker needs no table lookup for them.
    uint32_t qa = 0xFFFFFFFFu, qb = 0xFFFFFFFFu;
    std::vector<int> boring(nbyte, 0);      // 0 interesting, 1 boring non-word, 2 boring word
    {
        uint32_t a0 = d.start;
        for (int it = 0; it < 4; ++it) a0 = step(a0, 0x01);
        if (step(a0, 0x01) == a0 && !d.state_final[a0]) {
            for (const char* wb = "zqjxkvy_0"; *wb; ++wb) {
                uint32_t b0 = step(a0, (uint8_t)*wb);
                if (step(b0, (uint8_t)*wb) == b0 && step(b0, 0x01) == a0 && !d.state_final[b0]) { qa = a0; qb = b0; break; }
            }
        }
        if (qa != 0xFFFFFFFFu) {
            for (uint32_t c = 0; c < nbyte; ++c) {
                const uint32_t ta = d.trans[qa * nc + c], tb = d.trans[qb * nc + c];
                if (ta == tb && ta == qa) boring[c] = 1;
                else if (ta == tb && ta == qb) boring[c] = 2;
            }
        }
    }
    // classes renumbered: boring non-word, boring word, the rest
    std::vector<uint32_t> cperm(nc);
    uint32_t cw = 0;
    for (uint32_t c = 0; c < nbyte; ++c) if (boring[c] == 1) cperm[c] = cw++;
    const uint32_t nb0 = cw;
    for (uint32_t c = 0; c < nbyte; ++c) if (boring[c] == 2) cperm[c] = cw++;
    const uint32_t nb1 = cw;
    for (uint32_t c = 0; c < nbyte; ++c) if (boring[c] == 0) cperm[c] = cw++;
    cperm[nbyte] = nbyte; cperm[nbyte + 1] = nbyte + 1;
    blob.assign(5 + 32 + 64 + ns + (ns * nc + 1) / 2, 0);
    blob[0] = ns; blob[1] = nc; blob[2] = perm[d.start]; blob[3] = first_final; blob[4] = (uint32_t)rx.size();
    for (size_t k = 0; k < rx.size(); ++k) blob[5 + k] = rx[k];
    blob[33] = qa == 0xFFFFFFFFu ? 0xFFFFFFFFu : perm[qa];
    blob[34] = qb == 0xFFFFFFFFu ? 0xFFFFFFFFu : perm[qb];
    blob[35] = nb0; blob[36] = nb1;
    uint8_t* cmo = reinterpret_cast<uint8_t*>(&blob[37]);
    for (int b2 = 0; b2 < 256; ++b2) cmo[b2] = (uint8_t)cperm[d.classmap[b2]];
    for (uint32_t s2 = 0; s2 < ns; ++s2) blob[101 + s2] = d.state_mask[inv[s2]];
    uint16_t* t = reinterpret_cast<uint16_t*>(&blob[101 + ns]);
    for (uint32_t s2 = 0; s2 < ns; ++s2)
        for (uint32_t x = 0; x < nc; ++x) t[s2 * nc + cperm[x]] = (uint16_t)perm[d.trans[inv[s2] * nc + x]];
}

// Host mirror of the device walker in k_dfa_scan (same blob, same fast path); CPU-only tests use it.
uint32_t run_group_blob(const std::vector<uint32_t>& blob, const uint8_t* s, size_t n) {
    const uint32_t ns = blob[0], nc = blob[1], ff = blob[3];
    const uint8_t* cm = reinterpret_cast<const uint8_t*>(&blob[37]);
    const uint32_t* smask = &blob[101];
    const uint16_t* T = reinterpret_cast<const uint16_t*>(&blob[101 + ns]);
    long tpos = -1;
    if (n >= 1 && s[n - 1] == '\r') tpos = (long)n - 1;
    else if (n >= 2 && s[n - 2] == 0xC2 && s[n - 1] == 0x85) tpos = (long)n - 2;
    else if (n >= 3 && s[n - 3] == 0xE2 && s[n - 2] == 0x80 && (s[n - 1] == 0xA8 || s[n - 1] == 0xA9)) tpos = (long)n - 3;
    uint32_t st = blob[2];
    for (size_t i = 0; i < n && st < ff; ++i) {
        if ((long)i == tpos) st = T[st * nc + nc - 2];
        if (st >= ff) break;
        st = T[st * nc + cm[s[i]]];
    }
    return smask[T[st * nc + nc - 1]];
}

// Lines that the group's automaton rejects on their first byte (begin-anchored patterns: the stack-trace regex of CAS:31
// dies on anything but white space or 'a') need not be walked, nor even read beyond that byte. The bytes that can start a
// final line terminator (\r, U+0085, U+2028/9: the walker feeds a virtual symbol in front of them) are never skipped.
void group_skip_set(const std::vector<uint32_t>& blob, uint32_t out[8]) {
    const uint32_t ns = blob[0], nc = blob[1], ff = blob[3];
    const uint8_t* cm = reinterpret_cast<const uint8_t*>(&blob[37]);
    const uint32_t* smask = &blob[101];
    const uint16_t* T = reinterpret_cast<const uint16_t*>(&blob[101 + ns]);
    for (int w = 0; w < 8; ++w) out[w] = 0;
    if (blob[2] >= ff) return;          // decided before the first byte: leave it to the walker
    for (uint32_t b = 0; b < 256; ++b) {
        if (b == '\r' || b == 0xC2 || b == 0xE2) continue;
        const uint32_t st = T[blob[2] * nc + cm[b]];
        if (st >= ff && smask[T[st * nc + nc - 1]] == 0) out[b >> 5] |= 1u << (b & 31);
    }
}

void build_plan(const std::vector<PlanPatternIn>& pats, Plan& pl, int64_t* bad) {
    pl = Plan();
    if (bad) *bad = -1;
    std::map<std::pair<std::string, int>, uint32_t> ids;
    std::vector<int> rx_flags;
    std::vector<std::vector<Literal>> rx_lits;
    std::vector<ParsedRegex> parsed;
    std::vector<std::vector<uint8_t>> rx_lit_mode;     // parallel to rx_lits

    auto add_regex = [&](const std::string& src, int flags) -> uint32_t {
        auto key = std::make_pair(src, flags);
        auto it = ids.find(key);
        if (it != ids.end()) return it->second;
        ParsedRegex pr = parse_java_regex(src, flags);
        uint32_t id = pl.n_rx++;
        ids.emplace(key, id);
        pl.rx_src.push_back(src);
        rx_flags.push_back(flags);
        Dfa dfa;
        bool is_dfa = pr.regular && build_dfa(*pr.root, dfa);
        pl.rx_kind.push_back(is_dfa ? RXK_DFA : RXK_VM);
        pl.rx_dfa_off.push_back(0); pl.rx_dfa_len.push_back(0); pl.rx_ncls.push_back(0); pl.rx_start.push_back(0);
        pl.rx_vm_off.push_back(0); pl.rx_vm_len.push_back(0); pl.rx_vm_slots.push_back(0);
        pl.rx_vm_caps.push_back(0); pl.rx_vm_set_base.push_back(0);
        pl.rx_cmap.resize((size_t)pl.n_rx * 256, 0);
        pl.rx_vm_first.resize((size_t)pl.n_rx * 8, ~0u);
        pl.rx_adfa_off.push_back(0); pl.rx_ancls.push_back(0);
        pl.rx_astart.resize((size_t)pl.n_rx * 3, 0);
        pl.rx_acmap.resize((size_t)pl.n_rx * 256, 0);
        if (is_dfa) {
            pl.rx_dfa_off[id] = (uint32_t)pl.dfa_trans.size();
            pl.rx_dfa_len[id] = (uint32_t)dfa.trans.size();
            pl.rx_ncls[id] = dfa.n_classes;
            pl.rx_start[id] = dfa.start;
            pl.dfa_trans.insert(pl.dfa_trans.end(), dfa.trans.begin(), dfa.trans.end());
            memcpy(&pl.rx_cmap[(size_t)id * 256], dfa.classmap, 256);
        } else {
            VmProgram vp;
            build_vm(pr, vp);
            pl.rx_vm_off[id] = (uint32_t)pl.vm_code.size();
            pl.rx_vm_len[id] = (uint32_t)vp.code.size();
            pl.rx_vm_slots[id] = vp.n_slots;
            for (int w = 0; w < 8; ++w) pl.rx_vm_first[(size_t)id * 8 + w] = vp.first[w];
            pl.rx_vm_caps[id] = vp.n_caps;
            pl.rx_vm_set_base[id] = (uint32_t)pl.vm_set_off.size();
            uint32_t base = (uint32_t)pl.vm_set_ranges.size();
            for (uint32_t o : vp.set_off) pl.vm_set_off.push_back(o + base);
            pl.vm_set_ranges.insert(pl.vm_set_ranges.end(), vp.set_ranges.begin(), vp.set_ranges.end());
            pl.vm_code.insert(pl.vm_code.end(), vp.code.begin(), vp.code.end());
            ++pl.n_vm;
        }
        std::vector<Literal> lits;
        std::vector<uint8_t> modes;
        std::vector<LocalLiteral> local;
        if (is_dfa && local_verify_form(*pr.root, local, 4)) {
            for (const auto& l : local) {
                lits.push_back(l.lit);
                modes.push_back((uint8_t)(1 | (l.left_boundary ? 2 : 0) | (l.right_boundary ? 4 : 0)));
            }
        } else {
            lits = required_literals(*pr.root, 4);
            modes.assign(lits.size(), 0);
            // every match starts (or ends) with one of the literals: verified at the occurrence by an
            // anchored automaton, forwards from its first byte or backwards from its last
            Dfa ad;
            const int am = is_dfa ? choose_anchor(*pr.root, lits, ad) : 0;
            if (am) {
                modes.assign(lits.size(), (uint8_t)(am == 1 ? 8 : 16));
                pl.rx_adfa_off[id] = (uint32_t)pl.dfa_trans.size();
                pl.rx_ancls[id] = ad.n_classes;
                for (int k = 0; k < 3; ++k) pl.rx_astart[(size_t)id * 3 + k] = ad.start_ctx[k];
                pl.dfa_trans.insert(pl.dfa_trans.end(), ad.trans.begin(), ad.trans.end());
                memcpy(&pl.rx_acmap[(size_t)id * 256], ad.classmap, 256);
            }
        }
        pl.rx_always.push_back(lits.empty() ? 1 : 0);
        rx_lits.push_back(std::move(lits));
        rx_lit_mode.push_back(std::move(modes));
        parsed.push_back(std::move(pr));
        return id;
    };

    // ContextAnalysisService.java:27-34
    add_regex("\\b(ERROR|FATAL|CRITICAL|SEVERE)\\b", RXF_ICASE);
    add_regex("\\b(WARN|WARNING)\\b", RXF_ICASE);
    add_regex("^\\s*at\\s+[\\w\\.\\$]+\\(.*\\)\\s*$", 0);
    add_regex("\\b\\w*Exception\\b|\\b\\w*Error\\b", 0);

    pl.n_pat = (uint32_t)pats.size();
    std::map<std::string, int> slots;
    std::vector<std::vector<uint32_t>> by_primary;
    pl.sec_begin.push_back(0);
    pl.seq_begin.push_back(0);
    pl.seq_ev_begin.push_back(0);
    for (size_t p = 0; p < pats.size(); ++p) {
        const PlanPatternIn& in = pats[p];
        if (bad) *bad = (int64_t)p;
        if (in.has_context && (in.before < 0 || in.after < 0))
            throw std::invalid_argument("negative lines_before/lines_after (Arrays.copyOfRange would throw at AnalysisService.java:143-149)");
        uint32_t pr = add_regex(in.primary, 0);
        pl.pat_primary.push_back(pr);
        if (by_primary.size() < pl.n_rx) by_primary.resize(pl.n_rx);
        by_primary[pr].push_back((uint32_t)p);
        pl.pat_conf.push_back(in.confidence);
        pl.pat_sev.push_back(in.severity);
        pl.pat_has_ctx.push_back(in.has_context);
        pl.pat_before.push_back(in.before);
        pl.pat_after.push_back(in.after);
        if (!in.has_id || blank(in.id)) pl.pat_slot.push_back(-1);
        else {
            auto it = slots.find(in.id);
            if (it == slots.end()) { it = slots.emplace(in.id, (int)pl.slot_first_pat.size()).first; pl.slot_first_pat.push_back((uint32_t)p); }
            pl.pat_slot.push_back(it->second);
        }
        for (const auto& s : in.secs) {
            pl.sec_rx.push_back(add_regex(s.regex, 0));
            pl.sec_weight.push_back(s.weight);
            pl.sec_window.push_back(s.window);
        }
        pl.sec_begin.push_back((uint32_t)pl.sec_rx.size());
        for (const auto& q : in.seqs) {
            for (const auto& e : q.events) pl.ev_rx.push_back(add_regex(e, 0));
            pl.seq_ev_begin.push_back((uint32_t)pl.ev_rx.size());
            pl.seq_bonus.push_back(q.bonus);
        }
        pl.seq_begin.push_back((uint32_t)pl.seq_bonus.size());
    }
    if (bad) *bad = -1;
    {
        std::vector<std::vector<uint32_t>> members(pl.slot_first_pat.size());
        for (uint32_t p2 = 0; p2 < pl.n_pat; ++p2) if (pl.pat_slot[p2] >= 0) members[pl.pat_slot[p2]].push_back(p2);
        pl.slot_begin.assign(1, 0);
        for (auto& v : members) { pl.slot_pats.insert(pl.slot_pats.end(), v.begin(), v.end()); pl.slot_begin.push_back((uint32_t)pl.slot_pats.size()); }
    }
    by_primary.resize(pl.n_rx);
    // Match spans (LP_OPT_SPANS) come from the backtracking engine, whose priorities are java.util.regex's:
    // every primary regex gets a bytecode program, also those whose existence is decided by a DFA.
    for (uint32_t r : pl.pat_primary) {
        if (pl.rx_vm_len[r]) continue;
        VmProgram vp;
        build_vm(parsed[r], vp);
        pl.rx_vm_off[r] = (uint32_t)pl.vm_code.size();
        pl.rx_vm_len[r] = (uint32_t)vp.code.size();
        pl.rx_vm_slots[r] = vp.n_slots;
        for (int w = 0; w < 8; ++w) pl.rx_vm_first[(size_t)r * 8 + w] = vp.first[w];
        pl.rx_vm_caps[r] = vp.n_caps;
        pl.rx_vm_set_base[r] = (uint32_t)pl.vm_set_off.size();
        const uint32_t base = (uint32_t)pl.vm_set_ranges.size();
        for (uint32_t o : vp.set_off) pl.vm_set_off.push_back(o + base);
        pl.vm_set_ranges.insert(pl.vm_set_ranges.end(), vp.set_ranges.begin(), vp.set_ranges.end());
        pl.vm_code.insert(pl.vm_code.end(), vp.code.begin(), vp.code.end());
    }
    pl.rxp_begin.assign(pl.n_rx + 1, 0);
    for (uint32_t r = 0; r < pl.n_rx; ++r) {
        pl.rxp_begin[r + 1] = pl.rxp_begin[r] + (uint32_t)by_primary[r].size();
        pl.max_fanout = std::max<uint32_t>(pl.max_fanout, (uint32_t)by_primary[r].size());
        for (uint32_t p : by_primary[r]) pl.rxp_pat.push_back(p);
    }
    // ---- scan groups: greedily pack always-on DFA regexes into product automata that fit shared memory
    pl.rx_in_group.assign(pl.n_rx, 0);
    pl.rx_record.assign(pl.n_rx, 0);
    for (uint32_t r = RX_FIRST_USER; r < pl.n_rx; ++r) pl.rx_record[r] = 1;
    for (uint32_t r : pl.pat_primary) pl.rx_record[r] = 1;
    for (uint32_t r : pl.sec_rx) pl.rx_record[r] = 1;
    for (uint32_t r : pl.ev_rx) pl.rx_record[r] = 1;
    {
        constexpr size_t GROUP_BYTES = 40 * 1024, TOTAL_BYTES = 100 * 1024;
        std::vector<uint32_t> cand;
        for (uint32_t r = 0; r < pl.n_rx; ++r)
            if (pl.rx_always[r] && pl.rx_kind[r] == RXK_DFA) cand.push_back(r);
        // unanchored automata first, begin-anchored ones last: groups made only of anchored patterns
        // reach a final state early and stop costing table lookups
        std::stable_sort(cand.begin(), cand.end(), [&](uint32_t x, uint32_t y) {
            return (int)begins_anchored(*parsed[x].root) < (int)begins_anchored(*parsed[y].root);
        });
        auto serialise = [&](const std::vector<uint32_t>& rx, const MultiDfa& d, std::vector<uint32_t>& blob) { serialise_group(rx, d, blob); };
        size_t total = 0;
        std::vector<uint32_t> cur;
        MultiDfa cur_dfa;
        auto flush = [&]() {
            if (cur.empty()) return;
            Plan::ScanGroup g;
            g.rx = cur;
            serialise(cur, cur_dfa, g.blob);
            group_skip_set(g.blob, g.skip1);
            total += g.blob.size() * 4;
            for (uint32_t r : cur) pl.rx_in_group[r] = 1;
            pl.groups.push_back(std::move(g));
            cur.clear();
        };
        auto bytes_of = [](const MultiDfa& d) { return (size_t)(101 + d.n_states()) * 4 + (size_t)d.n_states() * d.n_classes * 2; };
        for (uint32_t r : cand) {
            std::vector<uint32_t> trial = cur;
            trial.push_back(r);
            std::vector<const Ast*> roots;
            for (uint32_t q : trial) roots.push_back(parsed[q].root.get());
            MultiDfa d;
            if (trial.size() <= 28 && build_multi_dfa(roots, d, 4096) && bytes_of(d) <= GROUP_BYTES && total + bytes_of(d) <= TOTAL_BYTES) {
                cur = trial; cur_dfa = d;
                continue;
            }
            flush();
            roots.assign(1, parsed[r].root.get());
            MultiDfa single;
            if (build_multi_dfa(roots, single, 4096) && bytes_of(single) <= GROUP_BYTES && total + bytes_of(single) <= TOTAL_BYTES) {
                cur.assign(1, r); cur_dfa = single;
            }   // else: stays on the global-table path (k_verify_always)
        }
        flush();
        pl.ctx_in_groups = pl.rx_in_group[RX_ERR] && pl.rx_in_group[RX_WARN] && pl.rx_in_group[RX_STACK] && pl.rx_in_group[RX_EXC];
        pl.ctx_any_in_groups = pl.rx_in_group[RX_ERR] || pl.rx_in_group[RX_WARN] || pl.rx_in_group[RX_STACK] || pl.rx_in_group[RX_EXC];
        for (uint32_t r = 0; r < RX_FIRST_USER; ++r)
            if (!pl.rx_in_group[r]) pl.rx_record[r] = 1;   // its context flag is derived from its records
    }
    for (uint32_t r = 0; r < pl.n_rx; ++r)
        if (pl.rx_always[r] && !pl.rx_in_group[r]) (pl.rx_kind[r] == RXK_DFA ? pl.always_list : pl.always_vm_list).push_back(r);

    // ---- literal table: de-duplicate (bytes, icase) across regexes
    std::map<std::pair<std::string, bool>, uint32_t> lit_ids;
    std::vector<std::vector<uint32_t>> lit_rx;
    std::vector<std::vector<uint8_t>> lit_mode;
    for (uint32_t r = 0; r < pl.n_rx; ++r) {
        for (size_t li = 0; li < rx_lits[r].size(); ++li) {
            const Literal& l = rx_lits[r][li];
            auto key = std::make_pair(l.bytes, l.icase);
            auto it = lit_ids.find(key);
            if (it == lit_ids.end()) {
                uint32_t id = (uint32_t)pl.lit_off.size();
                it = lit_ids.emplace(key, id).first;
                pl.lit_off.push_back((uint32_t)pl.lit_bytes.size());
                pl.lit_len.push_back((uint32_t)l.bytes.size());
                pl.lit_icase.push_back(l.icase ? 1 : 0);
                pl.lit_bytes.insert(pl.lit_bytes.end(), l.bytes.begin(), l.bytes.end());
                for (unsigned char ch : l.bytes) if (ch == '\r' || ch == '\n') pl.lits_have_line_breaks = true;
                lit_rx.emplace_back();
                lit_mode.emplace_back();
                pl.lit_gram_off.push_back(0);
            }
            auto& v = lit_rx[it->second];
            if (std::find(v.begin(), v.end(), r) == v.end()) { v.push_back(r); lit_mode[it->second].push_back(rx_lit_mode[r][li]); }
        }
    }
    pl.lit_rx_begin.push_back(0);
    for (size_t l = 0; l < lit_rx.size(); ++l) {
        pl.lit_rx.insert(pl.lit_rx.end(), lit_rx[l].begin(), lit_rx[l].end());
        pl.lit_rx_mode.insert(pl.lit_rx_mode.end(), lit_mode[l].begin(), lit_mode[l].end());
        pl.lit_rx_begin.push_back((uint32_t)pl.lit_rx.size());
    }
    select_grams(pl, nullptr, 0);
}

// (Re)chooses one filter gram per literal and rebuilds the Bloom bitmap and the exact gram table.
// Without a sample, the window that the fewest literals of the set share wins (a window common to
// many patterns, like the tail of "...Exception", is also frequent in the logs they are written
// for). With a sample of real log text, the window seen least often in the sample wins. The choice
// only changes how many candidates the prefilter emits, never the results.
void select_grams(Plan& pl, const uint8_t* sample, size_t sample_len) {
    auto folded_at = [&](uint32_t l, uint32_t o) {
        const uint8_t* b = &pl.lit_bytes[pl.lit_off[l] + o];
        return gram_fold((uint32_t)b[0] | ((uint32_t)b[1] << 8) | ((uint32_t)b[2] << 16) | ((uint32_t)b[3] << 24));
    };
    const size_t n_lit = pl.lit_off.size();
    constexpr uint32_t CB = 22;
    std::vector<uint32_t> seen_cnt;
    if (sample && sample_len >= 4) {
        seen_cnt.assign(1u << CB, 0);
        uint32_t w = (uint32_t)sample[0] | ((uint32_t)sample[1] << 8) | ((uint32_t)sample[2] << 16);
        for (size_t i = 3; i < sample_len; ++i) {
            w = (w & 0x00FFFFFFu) | ((uint32_t)sample[i] << 24);
            ++seen_cnt[(gram_fold(w) * GRAM_MUL) >> (32 - CB)];
            w >>= 8;
        }
    }
    std::map<uint32_t, uint32_t> df;
    for (uint32_t l = 0; l < n_lit; ++l) {
        std::vector<uint32_t> seen;
        for (uint32_t o = 0; o + 4 <= pl.lit_len[l]; ++o) {
            uint32_t f = folded_at(l, o);
            if (std::find(seen.begin(), seen.end(), f) == seen.end()) { seen.push_back(f); ++df[f]; }
        }
    }
    for (uint32_t l = 0; l < n_lit; ++l) {
        double best = -1e300; uint32_t best_off = 0;
        for (uint32_t o = 0; o + 4 <= pl.lit_len[l]; ++o) {
            const uint8_t* b = &pl.lit_bytes[pl.lit_off[l] + o];
            int rare = 0;
            for (int k = 0; k < 4; ++k) rare += rarity(b[k]);
            const bool distinct = b[0] != b[1] && b[1] != b[2] && b[2] != b[3];
            const uint32_t f = folded_at(l, o);
            double sc = rare * 2 + (distinct ? 1 : 0);
            if (!seen_cnt.empty()) sc -= 1000.0 * seen_cnt[(f * GRAM_MUL) >> (32 - CB)];
            else sc -= 6.0 * (df[f] - 1);
            if (sc > best) { best = sc; best_off = o; }
        }
        pl.lit_gram_off[l] = best_off;
    }
    pl.bitmap.assign((1u << GRAM_BITS) / 32, 0);
    pl.gram_lits.clear();
    std::map<uint32_t, std::vector<uint32_t>> gram2lits;
    for (uint32_t l = 0; l < n_lit; ++l) {
        const uint32_t f = folded_at(l, pl.lit_gram_off[l]);
        gram2lits[f].push_back(l);
        const uint32_t h = gram_hash1(f);
        pl.bitmap[h >> 5] |= 1u << (h & 31);
        const uint32_t h2 = (f * GRAM_MUL2) >> (32 - GRAM_BITS);       // second Bloom probe
        pl.bitmap[h2 >> 5] |= 1u << (h2 & 31);
    }
    uint32_t bits = 4;
    while ((1u << bits) < gram2lits.size() * 2 + 1) ++bits;
    pl.table_bits = bits;
    pl.table_key.assign(1u << bits, 0);
    pl.table_begin.assign(1u << bits, 0);
    pl.table_count.assign(1u << bits, 0);
    for (auto& kv : gram2lits) {
        uint32_t slot = gram_hash2(kv.first, bits);
        while (pl.table_key[slot] != 0) slot = (slot + 1) & ((1u << bits) - 1);
        pl.table_key[slot] = kv.first;
        pl.table_begin[slot] = (uint32_t)pl.gram_lits.size();
        pl.table_count[slot] = (uint32_t)kv.second.size();
        pl.gram_lits.insert(pl.gram_lits.end(), kv.second.begin(), kv.second.end());
    }
}

}  // namespace lpx
