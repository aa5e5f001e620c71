// Device kernels of the B200 hot path (sm_100a). HBM-bound byte/integer work: no tensor cores.
//
//   k_count_nl / k_scan_tiles / k_lines_filter   String.split("\\r?\\n")  AnalysisService.java:53
//                                                + multi-literal gram prefilter for :89-95
//   k_confirm / k_verify / k_verify_always       Matcher.find() per (line, regex)  AS:93-95,
//                                                SS:281,300,330, CAS:64-79
//   k_score                                      ScoringService.java:63-347, CAS:46-117, FTS:64-93
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
    const uint16_t* rx_ncls;
    const uint16_t* rx_start;
    const uint16_t* dfa_trans;
    const uint8_t* rx_cmap;
    const uint32_t* always_list;      // DFA regexes without a usable literal
    uint32_t n_always;
    const uint32_t* always_vm_list;   // bytecode regexes without a usable literal
    uint32_t n_always_vm;
    // literals / gram filter
    const uint32_t* bitmap;
    uint32_t table_bits;
    const uint32_t* table_key;
    const uint32_t* table_begin;
    const uint32_t* table_count;
    const uint32_t* gram_lits;
    const uint8_t* lit_bytes;
    const uint32_t* lit_off;
    const uint32_t* lit_len;
    const uint32_t* lit_gram_off;
    const uint8_t* lit_icase;
    const uint32_t* lit_rx_begin;
    const uint32_t* lit_rx;
    const uint8_t* lit_rx_mode;
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
    unsigned int overflow;     // bit0 cand, bit1 tasks, bit2 recs, bit3 events, bit4 backtrack budget, bit5 vm tasks, bit6 line arrays
    unsigned int n_unique;     // written by the unique pass (includes one sentinel group when padding was sorted)
    unsigned int n_vm_tasks;
    unsigned int n_lines_total;    // newlines + logs (G), published by k_scan_tiles
    unsigned int pad[4];
};

}  // namespace lpk
