// Device kernels of the B200 hot path (sm_100a). HBM-bound byte/integer work: no tensor cores.
//
//   k_filter_count / k_scan_tiles / k_lines / k_log_meta   String.split("\\r?\\n")  AnalysisService.java:53
//                                                          + literal-gram prefilter for :89-95
//   k_confirm / k_resolve / k_verify / k_verify_vm /       Matcher.find() per (line, regex)  AS:93-95,
//   k_dfa_scan / k_verify_always / k_verify_vm_always      SS:281,300,330, CAS:64-79
//   k_emit_events / k_tile_sort / k_tile_rank              event list in discovery order AS:100-109, ranking
//   k_score                                                ScoringService.java:63-347, CAS:46-117, FTS:64-93
//   k_range_tables / k_prior_prefix                        multi-GPU composition (no reference counterpart)
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace lpk {

constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITERS = 4;
constexpr uint32_t TILE_BYTES = SCAN_THREADS * 16 * SCAN_ITERS;   // 32 KiB of log per tile
constexpr uint32_t GRAM_BITS = 19;
constexpr uint32_t GRAM_WORDS = (1u << GRAM_BITS) / 32;
constexpr uint32_t GRAM_MUL = 0x9E3779B1u;
constexpr uint32_t GRAM_MUL2 = 0x85EBCA6Bu;

struct DevPlan {
    // key layouts of this call: record = regex << gb | line, event = line << pb | pattern
    uint32_t gb, pb;
    // regex universe
    uint32_t n_rx;
    const uint8_t* rx_kind;
    const uint32_t* rx_dfa_off;
    const uint32_t* rx_dfa_len;       // entries (u16) of the regex's whole-line automaton in dfa_trans
    const uint16_t* rx_ncls;
    const uint16_t* rx_start;
    const uint16_t* dfa_trans;
    const uint8_t* rx_cmap;
    // anchored automata (verification at a literal occurrence): {offset into dfa_trans, n_classes,
    // start(line edge) | start(after non-word) << 16, start(after word)} and a 256-byte class map per regex
    const uint4* anchor_desc;
    const uint8_t* rx_acmap;
    const uint32_t* always_list;      // DFA regexes without a usable literal
    uint32_t n_always;
    const uint32_t* always_vm_list;   // bytecode regexes without a usable literal
    uint32_t n_always_vm;
    // literals / gram filter
    // Tables packed so that k_confirm, whose lanes all read different addresses, needs one load each:
    //   gram_slots[2^table_bits]  {folded gram (0 = empty), first literal, literal count, -}
    //   lit_desc[]                literals grouped by gram: {offset into lit_bytes8, len | gram offset << 8 |
    //                             icase << 16 | first byte is a word byte << 17 | last << 18, first entry of
    //                             lit_rx_packed, entry count}
    //   lit_bytes8                literal bytes, each literal starting on an 8-byte boundary
    //   lit_rx_packed[]           regex id | verification mode << 24
    const uint32_t* bitmap;
    uint32_t table_bits;
    const uint4* gram_slots;
    const uint4* lit_desc;
    const uint8_t* lit_bytes8;
    const uint32_t* lit_rx_packed;
    // patterns
    uint32_t n_pat;
    const uint32_t* pat_primary;
    const double* pat_conf;
    const double* pat_sev;
    const int32_t* pat_has_ctx;
    const int32_t* pat_before;
    const int32_t* pat_after;
    const int32_t* pat_slot;
    const uint32_t* slot_first_pat;
    const uint32_t* slot_begin;
    const uint32_t* slot_pats;
    const uint32_t* sec_begin;
    const uint32_t* sec_rx;
    const double* sec_weight;
    const int32_t* sec_window;
    const uint32_t* seq_begin;
    const uint32_t* seq_ev_begin;
    const uint32_t* ev_rx;
    const double* seq_bonus;
    const uint32_t* rxp_begin;
    const uint32_t* rxp_pat;
    // backtracking programs
    const uint32_t* rx_vm_off;
    const uint32_t* rx_vm_len;
    const uint32_t* rx_vm_slots;
    const uint32_t* rx_vm_caps;
    const uint32_t* rx_vm_first;      // 8 words per regex
    const uint32_t* rx_vm_set_base;
    const void* vm_code;
    const uint32_t* vm_set_off;
    const uint32_t* vm_set_ranges;
};

struct ScoreCfg {
    double early, max_early, penalty, max_ctx, freq_thr, max_pen;
    int32_t max_window, hours;
    const double* decay_lut;   // exp(-d/decay) for d = 0..max_window
};

// counters block in device memory (one 64-byte struct, zeroed per call)
struct Counters {
    unsigned long long total_nl;
    unsigned int n_cand;
    unsigned int n_tasks;
    unsigned int n_recs;
    unsigned int n_events;
    unsigned int overflow;     // bit0 cand, bit1 tasks, bit2 recs, bit3 events, bit4 backtrack budget, bit5 vm tasks, bit6 line arrays, bit7 item queues
    unsigned int n_unique;     // written by the unique pass (includes one sentinel group when padding was sorted)
    unsigned int n_vm_tasks;
    unsigned int n_lines_total;    // newlines + logs (G), published by k_scan_tiles
    unsigned int n_items[4];       // (literal occurrence, regex) pairs per verification class, k_confirm -> k_resolve
    unsigned int n_budget;         // (line, regex) evaluations the backtracking verifier gave up on (step / stack budget)
    unsigned int vm_pool_next;     // deep backtracking stacks handed out so far in this call
    unsigned int n_scan;           // lines the always-on scan has to walk (k_expand's list)
};

}  // namespace lpk
