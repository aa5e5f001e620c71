// Builds the device plan from the flattened pattern set (see plan.h).
#include "plan.h"

#include <algorithm>
#include <cstring>
#include <map>

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

bool blank(const std::string& s) {
    for (unsigned char c : s) if (c > ' ') return false;
    return true;
}

}  // namespace

void build_plan(const std::vector<PlanPatternIn>& pats, Plan& pl, int64_t* bad) {
    pl = Plan();
    if (bad) *bad = -1;
    std::map<std::pair<std::string, int>, uint32_t> ids;
    std::vector<int> rx_flags;
    std::vector<std::vector<Literal>> rx_lits;

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
        pl.rx_dfa_off.push_back(0); pl.rx_ncls.push_back(0); pl.rx_start.push_back(0);
        pl.rx_vm_off.push_back(0); pl.rx_vm_len.push_back(0); pl.rx_vm_slots.push_back(0);
        pl.rx_vm_caps.push_back(0); pl.rx_vm_set_base.push_back(0);
        pl.rx_cmap.resize((size_t)pl.n_rx * 256, 0);
        if (is_dfa) {
            pl.rx_dfa_off[id] = (uint32_t)pl.dfa_trans.size();
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
            pl.rx_vm_caps[id] = vp.n_caps;
            pl.rx_vm_set_base[id] = (uint32_t)pl.vm_set_off.size();
            uint32_t base = (uint32_t)pl.vm_set_ranges.size();
            for (uint32_t o : vp.set_off) pl.vm_set_off.push_back(o + base);
            pl.vm_set_ranges.insert(pl.vm_set_ranges.end(), vp.set_ranges.begin(), vp.set_ranges.end());
            pl.vm_code.insert(pl.vm_code.end(), vp.code.begin(), vp.code.end());
            ++pl.n_vm;
        }
        std::vector<Literal> lits = required_literals(*pr.root, 4);
        pl.rx_always.push_back(lits.empty() ? 1 : 0);
        rx_lits.push_back(std::move(lits));
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
    by_primary.resize(pl.n_rx);
    pl.rxp_begin.assign(pl.n_rx + 1, 0);
    for (uint32_t r = 0; r < pl.n_rx; ++r) {
        pl.rxp_begin[r + 1] = pl.rxp_begin[r] + (uint32_t)by_primary[r].size();
        pl.max_fanout = std::max<uint32_t>(pl.max_fanout, (uint32_t)by_primary[r].size());
        for (uint32_t p : by_primary[r]) pl.rxp_pat.push_back(p);
    }
    for (uint32_t r = 0; r < pl.n_rx; ++r)
        if (pl.rx_always[r]) (pl.rx_kind[r] == RXK_DFA ? pl.always_list : pl.always_vm_list).push_back(r);

    // ---- literal table: de-duplicate (bytes, icase) across regexes
    std::map<std::pair<std::string, bool>, uint32_t> lit_ids;
    std::vector<std::vector<uint32_t>> lit_rx;
    for (uint32_t r = 0; r < pl.n_rx; ++r) {
        for (const Literal& l : rx_lits[r]) {
            auto key = std::make_pair(l.bytes, l.icase);
            auto it = lit_ids.find(key);
            if (it == lit_ids.end()) {
                uint32_t id = (uint32_t)pl.lit_off.size();
                it = lit_ids.emplace(key, id).first;
                pl.lit_off.push_back((uint32_t)pl.lit_bytes.size());
                pl.lit_len.push_back((uint32_t)l.bytes.size());
                pl.lit_icase.push_back(l.icase ? 1 : 0);
                pl.lit_bytes.insert(pl.lit_bytes.end(), l.bytes.begin(), l.bytes.end());
                lit_rx.emplace_back();
                // choose the rarest 4-byte window of the literal as its filter gram
                int best = -1; uint32_t best_off = 0;
                for (uint32_t o = 0; o + 4 <= l.bytes.size(); ++o) {
                    int sc = 0;
                    for (int k = 0; k < 4; ++k) sc += rarity((uint8_t)l.bytes[o + k]);
                    bool distinct = l.bytes[o] != l.bytes[o + 1] && l.bytes[o + 1] != l.bytes[o + 2] && l.bytes[o + 2] != l.bytes[o + 3];
                    sc = sc * 2 + (distinct ? 1 : 0);
                    if (sc > best) { best = sc; best_off = o; }
                }
                pl.lit_gram_off.push_back(best_off);
            }
            auto& v = lit_rx[it->second];
            if (std::find(v.begin(), v.end(), r) == v.end()) v.push_back(r);
        }
    }
    pl.lit_rx_begin.push_back(0);
    for (auto& v : lit_rx) {
        pl.lit_rx.insert(pl.lit_rx.end(), v.begin(), v.end());
        pl.lit_rx_begin.push_back((uint32_t)pl.lit_rx.size());
    }

    // ---- gram bitmap + exact gram table
    pl.bitmap.assign((1u << GRAM_BITS) / 32, 0);
    std::map<uint32_t, std::vector<uint32_t>> gram2lits;
    for (uint32_t l = 0; l < pl.lit_off.size(); ++l) {
        const uint8_t* b = &pl.lit_bytes[pl.lit_off[l] + pl.lit_gram_off[l]];
        uint32_t w = (uint32_t)b[0] | ((uint32_t)b[1] << 8) | ((uint32_t)b[2] << 16) | ((uint32_t)b[3] << 24);
        uint32_t f = gram_fold(w);
        gram2lits[f].push_back(l);
        uint32_t h = gram_hash1(f);
        pl.bitmap[h >> 5] |= 1u << (h & 31);
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
