// Host-side "plan": everything lp_load_patterns derives from the flattened pattern set before it
// is uploaded — the de-duplicated regex universe with its DFA tables / bytecode, the literal
// prefilter (gram bitmap + exact gram table) and the per-pattern scoring parameters.
// Replaces AnalysisService.java:56-86 (compile) and the object graph ScoringService walks
// (SS:163-178, 201-214, 232-258).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/logparser_b200.h"
#include "rx.h"
#include "vm.h"

namespace lpx {

constexpr uint32_t RX_ERR = 0, RX_WARN = 1, RX_STACK = 2, RX_EXC = 3, RX_FIRST_USER = 4;
constexpr uint32_t GRAM_BITS = 19;             // 2^19-bit (64 KiB) first-level bitmap in shared memory
constexpr uint32_t GRAM_MUL = 0x9E3779B1u;
constexpr uint32_t GRAM_MUL2 = 0x85EBCA6Bu;

enum RxKind : uint8_t { RXK_DFA = 0, RXK_VM = 1 };

struct PlanPatternIn {            // mirror of lp_pattern_desc with owned strings
    std::string id; bool has_id = false;
    std::string primary;
    double confidence = 0, severity = 1;
    int has_context = 0, before = 0, after = 0;
    struct Sec { std::string regex; double weight; int window; };
    struct Seq { std::vector<std::string> events; double bonus; };
    std::vector<Sec> secs;
    std::vector<Seq> seqs;
};

struct Plan {
    // ---- regex universe (ids 0..3 are the context regexes of CAS:27-34)
    uint32_t n_rx = 0;
    std::vector<std::string> rx_src;
    std::vector<uint8_t> rx_kind;         // RxKind
    std::vector<uint8_t> rx_always;       // 1: no usable literal -> every line is a candidate
    std::vector<uint32_t> rx_dfa_off;     // offset into dfa_trans (u16 units)
    std::vector<uint32_t> rx_dfa_len;     // entries of the whole-line automaton
    std::vector<uint16_t> rx_ncls;
    std::vector<uint16_t> rx_start;
    std::vector<uint16_t> dfa_trans;
    std::vector<uint8_t> rx_cmap;         // 256 per regex
    std::vector<uint32_t> rx_vm_first;    // 8 per regex: bytes a match of the bytecode program can begin with
    // anchored automaton (rx_compile.cpp: build_dfa_anchored) of regexes decided at their literal occurrences
    std::vector<uint32_t> rx_adfa_off;    // offset into dfa_trans
    std::vector<uint16_t> rx_ancls;
    std::vector<uint16_t> rx_astart;      // 3 per regex: walk starts at the line edge / next to a non-word byte / next to a word byte
    std::vector<uint8_t> rx_acmap;        // 256 per regex
    std::vector<uint32_t> always_list, always_vm_list;
    // bytecode programs (RXK_VM)
    std::vector<uint32_t> rx_vm_off, rx_vm_len, rx_vm_slots, rx_vm_caps, rx_vm_set_base;
    std::vector<VmInsn> vm_code;
    std::vector<uint32_t> vm_set_off, vm_set_ranges;

    // ---- always-on scan groups: product DFAs walked over every line from shared memory
    // (context regexes ERR/WARN/EXC/STACK + every DFA regex without a usable literal)
    struct ScanGroup {
        std::vector<uint32_t> rx;        // regex ids, bit k of the group's mask <-> rx[k]
        std::vector<uint32_t> blob;      // serialised tables (see k_dfa_scan)
        uint32_t skip1[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // bit b: a non-empty line whose first byte is b is decided, with no hit, after that byte
    };
    std::vector<ScanGroup> groups;
    std::vector<uint8_t> rx_in_group;    // 1: decided by a scan group
    std::vector<uint8_t> rx_record;      // 1: matches must be recorded as (regex,line) (0: context flag only)
    bool ctx_in_groups = false;          // all four context regexes are decided by scan groups
    bool ctx_any_in_groups = false;      // at least one is: k_dfa_scan writes the flag byte of every line

    // ---- literals and the gram filter
    std::vector<uint8_t> lit_bytes;
    std::vector<uint32_t> lit_off, lit_len, lit_gram_off;
    std::vector<uint8_t> lit_icase;
    bool lits_have_line_breaks = false;               // some literal holds '\r' or '\n' (k_confirm then looks lines up eagerly)
    std::vector<uint32_t> lit_rx_begin, lit_rx;       // CSR literal -> regex ids
    std::vector<uint8_t> lit_rx_mode;                 // per CSR entry: 0 walk the line; 1 | left\b<<1 | right\b<<2: boundary test at the occurrence; 8 / 16: anchored automaton forwards / backwards from the occurrence
    std::vector<uint32_t> bitmap;                     // 2^GRAM_BITS bits
    uint32_t table_bits = 0;
    std::vector<uint32_t> table_key;                  // folded gram, 0 = empty (0x20202020-folded grams are never 0)
    std::vector<uint32_t> table_begin, table_count;   // into gram_lits
    std::vector<uint32_t> gram_lits;

    // ---- patterns
    uint32_t n_pat = 0;
    std::vector<uint32_t> pat_primary;                // regex id
    std::vector<double> pat_conf, pat_sev;
    std::vector<int32_t> pat_has_ctx, pat_before, pat_after;
    std::vector<int32_t> pat_slot;                    // frequency slot or -1 (FTS keys by id string)
    std::vector<uint32_t> slot_first_pat;
    std::vector<uint32_t> slot_begin, slot_pats;      // CSR frequency slot -> patterns sharing that id
    std::vector<uint32_t> sec_begin, sec_rx; std::vector<double> sec_weight; std::vector<int32_t> sec_window;
    std::vector<uint32_t> seq_begin, seq_ev_begin, ev_rx; std::vector<double> seq_bonus;
    std::vector<uint32_t> rxp_begin, rxp_pat;         // CSR primary regex -> patterns (list order)
    uint32_t max_fanout = 1;

    uint32_t n_vm = 0;
};

// lp_pattern_desc[] of the C ABI -> owned strings; throws std::invalid_argument on a NULL regex (*bad = its pattern).
std::vector<PlanPatternIn> plan_input_from_desc(const lp_pattern_desc* pats, size_t n, int64_t* bad);
// Throws RegexSyntaxError / RegexUnsupported; *bad receives the flat pattern index.
void build_plan(const std::vector<PlanPatternIn>& pats, Plan& out, int64_t* bad);
void serialise_group(const std::vector<uint32_t>& rx, const MultiDfa& d, std::vector<uint32_t>& blob);
uint32_t run_group_blob(const std::vector<uint32_t>& blob, const uint8_t* s, size_t n);
// bit b set: from the start state, byte b leads to a final state whose hit mask is empty (group_first_byte_dead)
void group_skip_set(const std::vector<uint32_t>& blob, uint32_t out[8]);
// Re-picks the filter grams (optionally against a sample of real log text) and rebuilds the filter tables.
void select_grams(Plan& pl, const uint8_t* sample, size_t sample_len);

inline uint32_t gram_fold(uint32_t w) { return w | 0x20202020u; }
inline uint32_t gram_hash1(uint32_t folded) { return (folded * GRAM_MUL) >> (32 - GRAM_BITS); }
inline uint32_t gram_hash2(uint32_t folded, uint32_t bits) { return (folded * GRAM_MUL2) >> (32 - bits); }

}  // namespace lpx
