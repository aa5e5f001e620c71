// liblogparser_b200.so: kernels + host orchestration + the C ABI of include/logparser_b200.h.
// See kernels.cuh for the kernel -> reference map. No CPU fallback exists in this file.
#include <cub/cub.cuh>
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <chrono>
#include <cstring>
#include <dlfcn.h>
#include <functional>
#include <limits>
#include <memory>
#include <mutex>
#include <thread>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/logparser_b200.h"
#include "kernels.cuh"
#include "plan.h"
#include "vm_exec.h"

using namespace lpk;

// ============================================================================================
//                                         device code
// ============================================================================================
namespace {
#include "k_scan.cuh"
#include "k_fscan.cuh"
#include "k_verify.cuh"
#include "k_score.cuh"
#include "k_sort.cuh"
}  // namespace

// ============================================================================================
//                                         host code
// ============================================================================================
namespace {

thread_local std::string g_last_error;

struct CudaFail : std::runtime_error {
    explicit CudaFail(const std::string& m) : std::runtime_error(m) {}
};
struct LimitFail : std::runtime_error {
    explicit LimitFail(const std::string& m) : std::runtime_error(m) {}
};
#define CK(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess)                                                                            \
            throw CudaFail(std::string(#call) + ": " + cudaGetErrorString(e_) + " (" __FILE__ ":" + std::to_string(__LINE__) + ")"); \
    } while (0)

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    void ensure(size_t n) {
        if (n <= cap) return;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = std::max<size_t>(n + n / 4, 256);
        CK(cudaMalloc(&p, want));
        cap = want;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return static_cast<T*>(p); }
};

struct PinnedBuf {      // grow-only page-locked host staging: device->host copies into it are truly asynchronous
    uint8_t* p = nullptr;
    size_t cap = 0;
    void ensure(size_t n) {
        if (n <= cap) return;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        size_t want = std::max<size_t>(n + n / 2, 1 << 16);
        CK(cudaMallocHost(reinterpret_cast<void**>(&p), want));
        cap = want;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

template <class T>
const T* upload(std::vector<DevBuf>& keep, const std::vector<T>& v) {
    DevBuf b;
    b.ensure(std::max<size_t>(v.size() * sizeof(T), 16));
    if (!v.empty()) CK(cudaMemcpy(b.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    keep.push_back(b);
    return static_cast<const T*>(b.p);
}

}  // namespace

// Page-locked buffers travel between a context and the results it hands out: the device writes a
// call's outputs into one, the lp_result then owns it (no copy) and gives it back when it is freed.
struct StagePool {
    std::mutex mu;
    std::vector<PinnedBuf> free_list;
    bool closed = false;
    PinnedBuf acquire() {
        std::lock_guard<std::mutex> lk(mu);
        if (free_list.empty()) return PinnedBuf();
        PinnedBuf b = free_list.back();
        free_list.pop_back();
        return b;
    }
    void give_back(PinnedBuf b) {
        std::lock_guard<std::mutex> lk(mu);
        if (closed || free_list.size() >= 4) b.release(); else free_list.push_back(b);
    }
    void close() {
        std::lock_guard<std::mutex> lk(mu);
        closed = true;
        for (auto& b : free_list) b.release();
        free_list.clear();
    }
};

struct lp_result {
    PinnedBuf hold;                           // event columns live here (pinned, written by the device)
    std::shared_ptr<StagePool> pool;
    const uint32_t *line = nullptr, *pattern = nullptr, *rank = nullptr, *span_start = nullptr, *span_end = nullptr;
    const double *score = nullptr, *factors = nullptr;
    const uint64_t* freq_delta = nullptr;
    uint64_t n_patterns = 0;
    std::vector<uint64_t> log_n_lines, log_event_begin;
    std::vector<uint32_t> line_start, line_end, zero_rank;
    std::vector<unsigned long long> own;      // a result put together from several devices keeps its columns here
    uint64_t n_logs = 0, n_events = 0, total_lines = 0, budget_exceeded = 0;
    bool has_factors = false, has_offsets = false;
    ~lp_result() { if (hold.p) { if (pool) pool->give_back(hold); else hold.release(); } }
};

struct lp_ctx {
    int device = 0;
    lp_scoring_cfg cfg{};
    cudaStream_t stream = nullptr;
    std::mutex mu;
    int n_sm = 148;

    lpx::Plan plan;
    bool loaded = false;
    std::vector<DevBuf> plan_bufs;
    std::vector<const uint32_t*> d_group_blob;
    std::vector<uint32_t> group_words;
    const uint8_t* d_rx_record = nullptr;
    uint32_t n_group_record_rx = 0;
    DevBuf f_bitmap, f_key, f_lits, f_bytes;   // prefilter tables (re-uploaded by calibration)
    DevBuf line_skip;                          // per line: the always-on scan need not walk it (written by k_expand)
    DevBuf scan_list;                           // lines the always-on scan walks (k_expand)
    DevBuf f_bits, f_gkeys, fs_lists;           // fused scan: first-stage bit table, exact gram keys, per-span newline counts and lists
    const uint8_t* d_byte_table = nullptr;
    int scan_mode = 0;                         // 0: fused single-pass scan (k_fscan); 1: k_filter_count + k_scan_tiles + k_lines (LP_SCAN_MODE=legacy)
    uint32_t anchor_budget = ANCHOR_BUDGET;     // LP_ANCHOR_BUDGET
    uint32_t vm_blocks_per_sm = 12;            // k_verify_vm: blocks of 4 warps per SM (40 registers: 48 resident warps); the more warps, the fewer evaluations each one serialises (LP_VM_BLOCKS)
    uint32_t f_mul = fs::MUL1;                 // first-stage hash multiplier of k_fscan (pick_multiplier)
    int fscan_variant = 0;                     // first-stage lookup of k_fscan: 0 IMAD.HI hash, 1 IMAD+SHF hash, 2 byte table (LP_FSCAN_VARIANT)
    std::vector<uint32_t> lit_off8;   // literal -> offset in the 8-byte aligned device copy of the literal bytes
    bool calibrated = false;
    int recalibrations = 0;
    double cal_rate = -1.0;                     // candidates per line on the first call after a calibration (recalibration trigger)
    // what one worker learned from its first log serves the others: the owner context keeps the calibrated filter
    struct SharedCal { std::mutex mu; uint64_t version = 0; std::vector<uint32_t> lit_gram_off, bitmap, table_key, table_begin, table_count, gram_lits; uint32_t table_bits = 0, f_mul = 0; } shared_cal;
    uint64_t cal_seen = 0;                      // version of the owner's shared calibration this worker has adopted
    DevPlan dp{};
    DevBuf d_lut;
    ScoreCfg sc{};

    // scratch (grow-only)
    DevBuf text, log_off, tile_counts, counters, line_start, line_end, first_line, n_lines;
    DevBuf items, ss_keys, ss_idx;
    size_t est_items = 0;
    int occ_confirm = 0, occ_resolve = 0;
    DevBuf cand, tasks, group_hist, vm_tasks, vm_tasks_alt, vm_stack, recs, recs_alt, cub_tmp, rx_begin, ctx_flags;
    DevBuf ev_key, ev_key_alt, ev_log, slot_key, slot_key_alt, ev_idx, ev_idx_alt, rank_in_slot, freq_delta, prior;
    DevBuf out_arena, sort_key, sort_key_alt, order, order_alt, log_sorted, log_sorted_alt;
    DevBuf log_event_begin, n_unique;
    const uint8_t* host_src = nullptr;      // set by lp_analyze*: the log still has to be copied in (run_scan does it piecewise)
    cudaStream_t copy_stream = nullptr;
    std::vector<cudaEvent_t> chunk_ev;
    PinnedBuf stage;
    std::shared_ptr<StagePool> pool = std::make_shared<StagePool>();
    size_t est_lines = 0, est_cand = 0, est_tasks = 0, est_vm_tasks = 0, est_recs = 0, est_events = 0;
    int retries = 0;

    // scan-phase state kept for the two-phase batch API
    struct Pending {
        bool valid = false;
        const uint8_t* d_text = nullptr;
        uint64_t len = 0;
        uint32_t n_logs = 0, G = 0, n_unique = 0, m = 0;
        int options = 0;
        bool range = false;                 // line-range shard: events only for the lines between the halos
        uint32_t halo_before = 0, halo_after = 0;
    } pend;
    DevBuf seq_tab_off, seq_tab, seq_T, prior_eff, any_flag, dist_mine, dist_all, range_dev;
    void* nccl_comm = nullptr;              // ncclComm_t of lp_comm_init (NCCL is loaded at run time, see NcclApi)
    std::function<void()> after_unique;     // run_scan calls it once the unique records exist (the collective starts its exchange there)
    cudaEvent_t dist_ev[2] = {nullptr, nullptr};
    uint32_t comm_rank = 0, comm_world = 1;
    std::vector<uint32_t> h_seq_tab_off;    // per sequence, offset of its chain table in a shard summary
    uint32_t n_seq = 0;

    cudaEvent_t ev[10]{};
    lp_timings tm{};
    uint64_t launches = 0;

    // Further workers of this context: same pattern set and configuration, own stream, scratch and estimates. A stateless
    // request (lp_analyze, lp_analyze_device, lp_analyze_batch*) runs on whichever worker is free, so that requests from
    // several threads overlap on the GPU (one's host->device copy under another's kernels). The two-phase, collective and
    // line-range calls keep state between calls and stay on the context itself.
    std::vector<lp_ctx*> extra;
    // a multi-device context (lp_create(cfg, -1) / lp_create_multi): the contexts of the other devices; lp_analyze_batch
    // cuts a batch into one block of logs per device
    std::vector<lp_ctx*> shards;
    lp_ctx* owner = nullptr;                    // the context this worker belongs to (itself for the context)
    std::atomic<uint32_t> next_worker{0};
    std::atomic<lp_ctx*> last_used{nullptr};
};

namespace {

template <class T>
const T* upload_into(DevBuf& b, const std::vector<T>& v) {
    b.ensure(std::max<size_t>(v.size() * sizeof(T), 16));
    if (!v.empty()) CK(cudaMemcpy(b.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return static_cast<const T*>(b.p);
}
void upload_filter(lp_ctx* c) {
    const lpx::Plan& pl = c->plan;
    DevPlan& d = c->dp;
    d.bitmap = upload_into(c->f_bitmap, pl.bitmap);
    {   // first-stage table of k_filter_count: one byte per 16-bit hash of a folded gram
        std::vector<uint8_t> bt(65536, 0);
        for (uint32_t key : pl.table_key) if (key) bt[(key * GRAM_MUL) >> 16] = 1;
        c->d_byte_table = upload_into(c->f_bytes, bt);
    }
    if (c->fscan_variant != 2) {   // k_fscan: one bit per gram at (row, rotation - 1) of the per-lane table; every lane's copy is the same
        std::vector<uint32_t> bits(fs::ROWS, 0);
        for (uint32_t key : pl.table_key) if (key) {
            const uint32_t H = fs::host_hash(c->fscan_variant, key, c->f_mul);
            bits[fs::host_row(H)] |= 1u << ((H + 31u) & 31u);
        }
        upload_into(c->f_bits, bits);
    }
    upload_into(c->f_gkeys, pl.table_key);
    d.table_bits = pl.table_bits;
    // literal descriptors grouped by gram slot (a literal has exactly one gram)
    auto is_word = [](unsigned char b) { return (b >= '0' && b <= '9') || (b >= 'A' && b <= 'Z') || b == '_' || (b >= 'a' && b <= 'z'); };
    std::vector<uint4> slots(pl.table_key.size()), descs;
    descs.reserve(pl.lit_off.size());
    for (size_t sl = 0; sl < pl.table_key.size(); ++sl) {
        slots[sl] = make_uint4(pl.table_key[sl], (uint32_t)descs.size(), pl.table_key[sl] ? pl.table_count[sl] : 0u, 0u);
        if (!pl.table_key[sl]) continue;
        for (uint32_t q = 0; q < pl.table_count[sl]; ++q) {
            const uint32_t l = pl.gram_lits[pl.table_begin[sl] + q];
            const uint32_t len = pl.lit_len[l];
            const unsigned char first = pl.lit_bytes[pl.lit_off[l]], last = pl.lit_bytes[pl.lit_off[l] + len - 1];
            if (len > 255 || pl.lit_gram_off[l] > 255) throw std::logic_error("literal too long for the packed descriptor");
            descs.push_back(make_uint4(c->lit_off8[l], len | (pl.lit_gram_off[l] << 8) | (pl.lit_icase[l] ? 1u << 16 : 0u) |
                                                       (is_word(first) ? 1u << 17 : 0u) | (is_word(last) ? 1u << 18 : 0u),
                                       pl.lit_rx_begin[l], pl.lit_rx_begin[l + 1] - pl.lit_rx_begin[l]));
        }
    }
    if (descs.empty()) descs.push_back(make_uint4(0, 0, 0, 0));
    d.gram_slots = upload_into(c->f_key, slots);
    d.lit_desc = upload_into(c->f_lits, descs);
}
// The first-stage table of k_fscan holds one bit per gram at a place given by one multiplicative hash, so a window
// of the log that is no gram passes it when its hash lands on a set bit. The 4-byte windows of log text are few and
// heavily repeated; among 256 odd multipliers the one whose set bits are hit by the least sampled text wins (the
// mass of false first-stage hits varies several-fold between multipliers). Results never depend on the choice: every
// first-stage hit is checked against the exact gram keys.
void pick_multiplier(lp_ctx* c, const std::vector<uint8_t>& sample) {
    c->f_mul = fs::MUL1;
    if (c->fscan_variant == 2 || sample.size() < 4096) return;
    const size_t n = std::min<size_t>(sample.size(), (size_t)1 << 20) - 3;
    std::vector<uint32_t> grams(n);
    for (size_t i = 0; i < n; ++i)
        grams[i] = ((uint32_t)sample[i] | ((uint32_t)sample[i + 1] << 8) | ((uint32_t)sample[i + 2] << 16) | ((uint32_t)sample[i + 3] << 24)) | 0x20202020u;
    std::sort(grams.begin(), grams.end());
    std::vector<uint32_t> keys;
    for (uint32_t k : c->plan.table_key) if (k) keys.push_back(k);
    std::sort(keys.begin(), keys.end());
    std::vector<std::pair<uint32_t, uint32_t>> text;      // distinct windows of the sample that are no gram, with their counts
    for (size_t i = 0; i < n;) {
        size_t j = i;
        while (j < n && grams[j] == grams[i]) ++j;
        if (!std::binary_search(keys.begin(), keys.end(), grams[i])) text.emplace_back(grams[i], (uint32_t)(j - i));
        i = j;
    }
    uint64_t best = ~0ull;
    std::vector<uint32_t> bits(fs::ROWS);
    uint32_t m = fs::MUL1;
    for (int trial = 0; trial < 256; ++trial, m = (m * 0x2C1B3C6Du + 0x297A2D39u) | 1u) {
        std::fill(bits.begin(), bits.end(), 0u);
        for (uint32_t k : keys) { const uint32_t H = fs::host_hash(c->fscan_variant, k, m); bits[fs::host_row(H)] |= 1u << ((H + 31u) & 31u); }
        uint64_t mass = 0;
        for (const auto& g : text) {
            const uint32_t H = fs::host_hash(c->fscan_variant, g.first, m);
            if ((bits[fs::host_row(H)] >> ((H + 31u) & 31u)) & 1u) mass += g.second;
        }
        if (mass < best) { best = mass; c->f_mul = m; }
    }
}
// One-time tuning of the literal prefilter against the first log this context sees: 4 x 1 MiB of
// its text decide which 4-byte window of each literal is the rarest. Results never depend on it.
void calibrate_filter(lp_ctx* c, const uint8_t* d_text, uint64_t len) {
    const uint64_t piece = 1u << 20;
    std::vector<uint8_t> sample;
    for (int k = 0; k < 4; ++k) {
        const uint64_t off = (len / 4) * k & ~15ull;
        if (off >= len) break;
        const uint64_t n = std::min<uint64_t>(piece, len - off);
        const size_t at = sample.size();
        sample.resize(at + n + 1);
        CK(cudaMemcpyAsync(sample.data() + at, d_text + off, n, cudaMemcpyDeviceToHost, c->stream));
        sample[at + n] = '\n';
        if (len <= piece) break;
    }
    CK(cudaStreamSynchronize(c->stream));
    lpx::select_grams(c->plan, sample.data(), sample.size());
    pick_multiplier(c, sample);
    upload_filter(c);
    c->calibrated = true;
}
void upload_plan(lp_ctx* c) {
    for (auto& b : c->plan_bufs) b.release();
    c->plan_bufs.clear();
    const lpx::Plan& pl = c->plan;
    DevPlan& d = c->dp;
    auto& K = c->plan_bufs;
    d.n_rx = pl.n_rx;
    d.rx_kind = upload(K, pl.rx_kind);
    d.rx_dfa_off = upload(K, pl.rx_dfa_off);
    d.rx_dfa_len = upload(K, pl.rx_dfa_len);
    d.rx_ncls = upload(K, pl.rx_ncls);
    d.rx_start = upload(K, pl.rx_start);
    d.dfa_trans = upload(K, pl.dfa_trans);
    d.rx_cmap = upload(K, pl.rx_cmap);
    {
        std::vector<uint4> ad(std::max<uint32_t>(pl.n_rx, 1));
        for (uint32_t r = 0; r < pl.n_rx; ++r)
            ad[r] = make_uint4(pl.rx_adfa_off[r], pl.rx_ancls[r], (uint32_t)pl.rx_astart[r * 3] | ((uint32_t)pl.rx_astart[r * 3 + 1] << 16),
                               pl.rx_astart[r * 3 + 2]);
        d.anchor_desc = upload(K, ad);
    }
    d.rx_acmap = upload(K, pl.rx_acmap);
    d.always_list = upload(K, pl.always_list);
    d.n_always = (uint32_t)pl.always_list.size();
    d.always_vm_list = upload(K, pl.always_vm_list);
    d.n_always_vm = (uint32_t)pl.always_vm_list.size();
    {
        std::vector<uint8_t> b8;
        c->lit_off8.assign(pl.lit_off.size(), 0);
        for (size_t l = 0; l < pl.lit_off.size(); ++l) {
            c->lit_off8[l] = (uint32_t)b8.size();
            b8.insert(b8.end(), pl.lit_bytes.begin() + pl.lit_off[l], pl.lit_bytes.begin() + pl.lit_off[l] + pl.lit_len[l]);
            b8.resize((b8.size() + 7) & ~size_t(7), 0);
        }
        b8.resize(b8.size() + 8, 0);
        d.lit_bytes8 = upload(K, b8);
        std::vector<uint32_t> packed(pl.lit_rx.size());
        if (pl.lit_rx.size() >= (1u << 24)) throw std::length_error("more than 2^24 (literal, regex) pairs");
        auto is_word = [](unsigned char b) { return (b >= '0' && b <= '9') || (b >= 'A' && b <= 'Z') || b == '_' || (b >= 'a' && b <= 'z'); };
        for (size_t l = 0; l < pl.lit_off.size(); ++l) {
            const uint32_t edge = (is_word(pl.lit_bytes[pl.lit_off[l]]) ? 32u : 0u) | (is_word(pl.lit_bytes[pl.lit_off[l] + pl.lit_len[l] - 1]) ? 64u : 0u);
            for (uint32_t j = pl.lit_rx_begin[l]; j < pl.lit_rx_begin[l + 1]; ++j) {
                if (pl.lit_rx[j] >= (1u << 24)) throw std::length_error("more than 2^24 distinct regexes");
                packed[j] = pl.lit_rx[j] | (((uint32_t)pl.lit_rx_mode[j] | edge) << 24);
            }
        }
        if (packed.empty()) packed.push_back(0);
        d.lit_rx_packed = upload(K, packed);
    }
    upload_filter(c);
    d.n_pat = pl.n_pat;
    d.pat_primary = upload(K, pl.pat_primary);
    d.pat_conf = upload(K, pl.pat_conf);
    d.pat_sev = upload(K, pl.pat_sev);
    d.pat_has_ctx = upload(K, pl.pat_has_ctx);
    d.pat_before = upload(K, pl.pat_before);
    d.pat_after = upload(K, pl.pat_after);
    d.pat_slot = upload(K, pl.pat_slot);
    d.slot_first_pat = upload(K, pl.slot_first_pat);
    d.slot_begin = upload(K, pl.slot_begin);
    d.slot_pats = upload(K, pl.slot_pats);
    d.sec_begin = upload(K, pl.sec_begin);
    d.sec_rx = upload(K, pl.sec_rx);
    d.sec_weight = upload(K, pl.sec_weight);
    d.sec_window = upload(K, pl.sec_window);
    d.seq_begin = upload(K, pl.seq_begin);
    d.seq_ev_begin = upload(K, pl.seq_ev_begin);
    d.ev_rx = upload(K, pl.ev_rx);
    d.seq_bonus = upload(K, pl.seq_bonus);
    d.rxp_begin = upload(K, pl.rxp_begin);
    d.rxp_pat = upload(K, pl.rxp_pat);
    d.rx_vm_off = upload(K, pl.rx_vm_off);
    d.rx_vm_len = upload(K, pl.rx_vm_len);
    d.rx_vm_slots = upload(K, pl.rx_vm_slots);
    d.rx_vm_caps = upload(K, pl.rx_vm_caps);
    d.rx_vm_first = upload(K, pl.rx_vm_first);
    d.rx_vm_set_base = upload(K, pl.rx_vm_set_base);
    d.vm_code = upload(K, pl.vm_code);
    d.vm_set_off = upload(K, pl.vm_set_off);
    d.vm_set_ranges = upload(K, pl.vm_set_ranges);
    c->d_group_blob.clear();
    c->group_words.clear();
    c->n_group_record_rx = 0;
    for (const auto& g : pl.groups) {
        c->d_group_blob.push_back(upload(K, g.blob));
        c->group_words.push_back((uint32_t)g.blob.size());
        for (uint32_t r : g.rx) c->n_group_record_rx += pl.rx_record[r];
    }
    c->d_rx_record = upload(K, pl.rx_record);
}

inline uint32_t blocks_for(uint64_t n, uint32_t threads) { return (uint32_t)std::max<uint64_t>(1, (n + threads - 1) / threads); }

template <class K>
void sort_keys(lp_ctx* c, K* in, K* out, uint32_t n, int end_bit, int begin_bit = 0) {
    size_t tmp = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, tmp, in, out, (int)n, begin_bit, end_bit, c->stream);
    c->cub_tmp.ensure(tmp);
    CK(cub::DeviceRadixSort::SortKeys(c->cub_tmp.p, tmp, in, out, (int)n, begin_bit, end_bit, c->stream));
    c->launches += 3;
}
template <class K, class V>
void sort_pairs(lp_ctx* c, K* kin, K* kout, V* vin, V* vout, uint32_t n, int end_bit) {
    size_t tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp, kin, kout, vin, vout, (int)n, 0, end_bit, c->stream);
    c->cub_tmp.ensure(tmp);
    CK(cub::DeviceRadixSort::SortPairs(c->cub_tmp.p, tmp, kin, kout, vin, vout, (int)n, 0, end_bit, c->stream));
    c->launches += 3;
}
// n <= SS_MAX_TILES * SS_TILE keys: tile sort + counting rank (see k_tile_sort). Exactly one of out_sorted / out_rank is used.
bool small_sort(lp_ctx* c, const unsigned long long* in, uint32_t n, int end_bit, unsigned long long* out_sorted, uint32_t* out_rank,
                const unsigned int* n_valid) {
    const uint32_t n_tiles = (n + SS_TILE - 1) / SS_TILE;
    if (n == 0 || n_tiles > SS_MAX_TILES) return false;
    if (n <= TR_MAX) {       // a small request: ranks by counting, one launch
        k_tiny_rank<<<(n + TR_ELEMS - 1) / TR_ELEMS, TR_THREADS, 0, c->stream>>>(in, n, out_sorted, out_rank, n_valid);
        c->launches += 1;
        return true;
    }
    c->ss_keys.ensure((size_t)n_tiles * SS_TILE * 8);
    c->ss_idx.ensure((size_t)n_tiles * SS_TILE * 4);
    k_tile_sort<<<n_tiles, SS_THREADS, sizeof(SsSorter::TempStorage), c->stream>>>(in, n, end_bit, c->ss_keys.as<unsigned long long>(), c->ss_idx.as<uint32_t>());
    k_tile_rank<<<blocks_for(n_tiles * SS_TILE, 128), 128, 0, c->stream>>>(c->ss_keys.as<unsigned long long>(), c->ss_idx.as<uint32_t>(), n_tiles, n,
                                                                          out_sorted, out_rank, n_valid);
    c->launches += 2;
    return true;
}
int bits_for(uint64_t v) { int b = 1; while (b < 64 && (v >> b)) ++b; return b; }

// ---- optimistic capacities -------------------------------------------------------------------
// Every queue and output of the pipeline is sized from what earlier calls on this context needed
// (first call: from the log size), kernels take their live counts from device-side counters, and
// padding slots hold all-ones sentinels that sort last. The host therefore synchronises once per
// call; a call whose real counts exceed an estimate is detected at that point and simply re-run
// with larger buffers (first calls on a context, or a log unlike its predecessors).
struct Retry {};

void bump(size_t& est, uint64_t observed) { est = std::max<size_t>(est, (size_t)(observed + observed / 4 + 1024)); }
// After a call that fitted: the estimate follows what the log needed, dropping by at most a quarter per call. Sorts and
// scans run over the estimated capacities, so a first guess that was several times too large (or one huge log) must not
// tax every later call; the slow decay keeps a workload that alternates between log sizes from re-running every time.
void settle(size_t& est, uint64_t observed) {
    const size_t need = (size_t)(observed + observed / 4 + 1024);
    est = std::max(need, est - est / 4);
}

void check_estimates(lp_ctx* c, const Counters& h, uint32_t n_logs, bool events_too, int options) {
    bool retry = false;
    const uint64_t G = (uint64_t)h.total_nl + n_logs;
    if (G >= 0xFFFFFFF0ull) throw LimitFail("too many lines");
    if ((h.overflow & 16u) && (options & LP_OPT_STRICT_BUDGET)) throw std::overflow_error("backtracking verifier exceeded its step/stack budget");
    if (G + 1 > c->est_lines) { bump(c->est_lines, G + 1); retry = true; }
    if (h.n_cand > c->est_cand) { bump(c->est_cand, h.n_cand); retry = true; }
    if (h.n_tasks > c->est_tasks) { bump(c->est_tasks, h.n_tasks); retry = true; }
    if (h.n_vm_tasks > c->est_vm_tasks) { bump(c->est_vm_tasks, h.n_vm_tasks); retry = true; }
    const uint32_t max_items = std::max(std::max(h.n_items[0], h.n_items[1]), std::max(h.n_items[2], h.n_items[3]));
    if (max_items > c->est_items) { bump(c->est_items, max_items); retry = true; }
    if (h.n_recs > c->est_recs) { bump(c->est_recs, h.n_recs); retry = true; }
    if (events_too && h.n_events > c->est_events) { bump(c->est_events, h.n_events); retry = true; }
    if (!retry && (h.overflow & ~16u)) {   // a flag without a matching count: grow everything once
        c->est_cand *= 2; c->est_tasks *= 2; c->est_recs *= 2; c->est_events *= 2; c->est_vm_tasks *= 2; c->est_items *= 2;
        retry = true;
    }
    if (retry) {
        if (++c->retries > 12) throw LimitFail("queues keep overflowing");
        throw Retry();
    }
    c->retries = 0;
    // remember what this log needed (with headroom) for the next call
    settle(c->est_lines, G + 1); settle(c->est_cand, h.n_cand); settle(c->est_tasks, h.n_tasks);
    settle(c->est_vm_tasks, h.n_vm_tasks); settle(c->est_recs, h.n_recs); settle(c->est_items, max_items);
    if (events_too) settle(c->est_events, h.n_events);
}

EstCaps est_caps(const lp_ctx* c) {
    return EstCaps{c->est_lines, c->est_cand, c->est_tasks, c->est_vm_tasks, c->est_recs, c->est_items, c->est_events};
}

// ---- phase 1: split + prefilter + verify + sorted unique records + events in discovery order.
// Launches only; the caller synchronises.
struct RangeShard { uint32_t halo_before, halo_after; };

void run_scan(lp_ctx* c, const uint8_t* d_text, uint64_t len, const uint64_t* h_off, uint32_t n_logs, int options,
              const RangeShard* rg = nullptr) {
    if (!c->loaded) throw std::invalid_argument("lp_load_patterns has not been called");
    if (len >= 0xFFFFFFF0ull) throw LimitFail("log/batch larger than 4 GiB - 16 bytes; shard it (see DESIGN.md)");
    if (n_logs == 0) throw std::invalid_argument("n_logs == 0");
    if ((reinterpret_cast<uintptr_t>(d_text) & 15) != 0) throw std::invalid_argument("device log pointer must be 16-byte aligned");
    for (uint32_t k = 0; k < n_logs; ++k) if (h_off[k] > h_off[k + 1]) throw std::invalid_argument("offsets must be non-decreasing");
    if (h_off[0] != 0 || h_off[n_logs] != len) throw std::invalid_argument("offsets[0] must be 0 and offsets[n_logs] the total size");
    cudaStream_t st = c->stream;
    c->owner->last_used.store(c);               // lp_get_timings reports the worker that ran last
    c->pend = lp_ctx::Pending();
    c->launches = 0;
    c->tm = lp_timings();

    const bool use_filter = !c->plan.lit_off.empty();
    if (use_filter && len >= 4096) {
        lp_ctx::SharedCal& sh = c->owner->shared_cal;
        std::unique_lock<std::mutex> lk(sh.mu);
        if (sh.version > c->cal_seen && !(c->calibrated == false && c->recalibrations > 0)) {
            // another worker of this context has already tuned the filter against a log: adopt its choice
            lpx::Plan& pl = c->plan;
            pl.lit_gram_off = sh.lit_gram_off; pl.bitmap = sh.bitmap; pl.table_key = sh.table_key; pl.table_begin = sh.table_begin;
            pl.table_count = sh.table_count; pl.gram_lits = sh.gram_lits; pl.table_bits = sh.table_bits;
            c->f_mul = sh.f_mul;
            c->cal_seen = sh.version;
            lk.unlock();
            upload_filter(c);
            c->calibrated = true;
            c->cal_rate = -1.0;
        } else if (!c->calibrated) {
            lk.unlock();
            calibrate_filter(c, d_text, len);
            c->cal_rate = -1.0;
            lk.lock();
            const lpx::Plan& pl = c->plan;
            sh.lit_gram_off = pl.lit_gram_off; sh.bitmap = pl.bitmap; sh.table_key = pl.table_key; sh.table_begin = pl.table_begin;
            sh.table_count = pl.table_count; sh.gram_lits = pl.gram_lits; sh.table_bits = pl.table_bits; sh.f_mul = c->f_mul;
            c->cal_seen = ++sh.version;
        }
    }
    const uint32_t n_always = (uint32_t)c->plan.always_list.size();
    const uint32_t n_always_vm = (uint32_t)c->plan.always_vm_list.size();
    const bool has_vm = c->plan.n_vm > 0;
    // one lane per (line, regex) evaluation; the top of its backtracking stack is in local memory, deeper entries come
    // from a pool (VM_POOL_STACKS x 16 KB, allocated only when such patterns exist)
    const uint32_t vm_grid = (uint32_t)c->n_sm * c->vm_blocks_per_sm, vm_block = 128;

    // first-call estimates (later calls reuse what earlier logs needed; an under-estimate costs one re-run)
    if (c->est_lines == 0) {
        c->est_lines = (size_t)(len / 64 + n_logs + 4096);
        c->est_cand = (size_t)(len / 256 + (1u << 16));
        c->est_tasks = (size_t)(len / 256 + (1u << 16));
        c->est_items = (size_t)(len / 256 + (1u << 16));
        c->est_vm_tasks = (size_t)(len / 1024 + (1u << 14));
        c->est_recs = (size_t)(len / 256 + (1u << 16));
        c->est_events = (size_t)(len / 2048 + (1u << 14));
    }
    c->est_lines = std::max<size_t>(c->est_lines, (size_t)n_logs + 16);
    const size_t line_cap = c->est_lines;
    const size_t rec_cap = c->est_recs;
    if (rec_cap >= 0x7FFFFFFFull || line_cap >= 0xFFFFFFF0ull) throw LimitFail("record capacity exceeds 2^31; shard the batch");
    const size_t ev_cap = c->est_events;
    c->dp.gb = (uint32_t)bits_for(line_cap);          // record key = regex << gb | line
    c->dp.pb = (uint32_t)bits_for(std::max<uint32_t>(c->dp.n_pat, 1));   // event key = line << pb | pattern

    const bool fused = c->scan_mode == 0;
    const uint32_t tile_bytes = fused ? fs::TILE : TILE_BYTES;
    const uint32_t n_tiles = (uint32_t)((len + 1 + tile_bytes - 1) / tile_bytes);   // covers position `len` too
    c->log_off.ensure((n_logs + 1) * sizeof(uint64_t));
    c->tile_counts.ensure((size_t)n_tiles * 4);
    c->counters.ensure(sizeof(Counters));
    c->line_start.ensure((line_cap + 1) * 4);
    c->line_end.ensure((line_cap + 1) * 4);
    c->first_line.ensure((size_t)(n_logs + 1) * 4);
    c->n_lines.ensure((size_t)n_logs * 4);
    c->ctx_flags.ensure((line_cap / 4 + 2) * 4);
    c->cand.ensure(c->est_cand * 4);
    c->tasks.ensure(c->est_tasks * 16);
    c->group_hist.ensure((size_t)(c->dp.n_rx + 1) * 4 * 4);      // histogram + cursors, for the DFA tasks and the bytecode tasks
    c->items.ensure(c->est_items * 8 * N_CLASSES);
    if (has_vm) { c->vm_tasks.ensure(c->est_vm_tasks * 8); c->vm_tasks_alt.ensure(c->est_vm_tasks * 8); }
    if (has_vm || !c->plan.always_vm_list.empty()) c->vm_stack.ensure((size_t)VM_POOL_STACKS * lpx::VM_STACK_WORDS * 4);
    c->recs.ensure(rec_cap * 8);
    c->recs_alt.ensure(rec_cap * 8);
    c->rx_begin.ensure((size_t)(c->dp.n_rx + 2) * 4);
    c->ev_key.ensure(ev_cap * 8 + 8);
    c->ev_key_alt.ensure(ev_cap * 8 + 8);
    Counters* ctr = c->counters.as<Counters>();

    // only the counters have to be clear before the first kernel; everything else is queued behind it so
    // that the GPU starts on the log as early as possible
    CK(cudaMemsetAsync(ctr, 0, sizeof(Counters), st));
    CK(cudaEventRecord(c->ev[0], st));
    const size_t smem = use_filter ? 65536 : 0;
    FscanArgs fz{};
    if (fused) {
        // per tile: 16 span counts, then per tile 16 lists of fs::WLIST entries
        c->fs_lists.ensure((size_t)n_tiles * 16 * 2 + (size_t)n_tiles * 16 * fs::WLIST * 2);
        fz.text = d_text; fz.len = len; fz.n_tiles = n_tiles;
        fz.warp_counts = c->fs_lists.as<uint16_t>();
        fz.tile_counts = c->tile_counts.as<uint32_t>();
        CK(cudaMemsetAsync(c->tile_counts.p, 0, (size_t)n_tiles * 4, st));
        fz.nl_lists = c->fs_lists.as<uint16_t>() + (size_t)n_tiles * 16;
        fz.mul = c->f_mul;
        { const uint32_t kk[8] = {1u, 1024u, 128u, 1u << 25, 1u << 26, 1u << 27, 1u << 28, 1u << 15}; for (int q = 0; q < 8; ++q) fz.k[q] = kk[q]; }
        fz.bit_table = c->fscan_variant == 2 ? reinterpret_cast<const uint32_t*>(c->d_byte_table) : c->f_bits.as<uint32_t>();
        fz.gram_keys = c->f_gkeys.as<uint32_t>(); fz.table_bits = c->dp.table_bits;
        fz.cand = c->cand.as<uint32_t>(); fz.cand_cap = (uint32_t)std::min<size_t>(c->est_cand, 0xFFFFFFFFu); fz.ctr = ctr;
        fz.use_filter = use_filter ? 1 : 0;
        // a line is left out of the always-on scan when ALL its groups are decided, without a hit, by its first byte
        for (int w = 0; w < 8; ++w) fz.skip1[w] = c->plan.groups.empty() ? 0u : 0xFFFFFFFFu;
        for (const auto& g : c->plan.groups) for (int w = 0; w < 8; ++w) fz.skip1[w] &= g.skip1[w];
        c->line_skip.ensure(line_cap + 16);
        c->scan_list.ensure((line_cap + 16) * 4);
    }
    auto launch_fscan = [&](uint32_t tile_begin, uint32_t tile_end) {
        fz.tile_begin = tile_begin; fz.tile_end = tile_end;
        // one CTA (32 warps, a pipeline each) per SM; fewer when the piece has less than one span per warp
        const uint32_t grid = std::max<uint32_t>(1, std::min<uint32_t>(((tile_end - tile_begin) * fs::SPANS + 31) / 32, (uint32_t)c->n_sm));
        switch (c->fscan_variant) {
            case 1: k_fscan<1><<<grid, fs::THREADS, fs::SMEM_BYTES, st>>>(fz); break;
            case 2: k_fscan<2><<<grid, fs::THREADS, fs::SMEM_BYTES, st>>>(fz); break;
            case 3: k_fscan<3><<<grid, fs::THREADS, fs::SMEM_BYTES, st>>>(fz); break;
            default: k_fscan<0><<<grid, fs::THREADS, fs::SMEM_BYTES, st>>>(fz); break;
        }
        CK(cudaGetLastError());
    };
    FilterArgs fa{};
    fa.text = d_text; fa.len = len; fa.n_tiles = n_tiles; fa.tile_counts = c->tile_counts.as<uint32_t>(); fa.bitmap = c->dp.bitmap; fa.byte_table = c->d_byte_table;
    fa.cand = c->cand.as<uint32_t>(); fa.cand_cap = (uint32_t)std::min<size_t>(c->est_cand, 0xFFFFFFFFu); fa.ctr = ctr;
    fa.use_filter = use_filter ? 1 : 0;
    CK(cudaEventRecord(c->ev[1], st));
    if (c->host_src) {
        // The log is still in (page-locked) host memory: it is copied in up to 8 pieces on a second
        // stream and each piece is scanned as soon as it has landed, so the copy hides the scan. A piece
        // carries 16 bytes of its successor for the windows that straddle the cut.
        const uint8_t* host = c->host_src;
        c->host_src = nullptr;                  // a re-run of this call finds the text on the device
        uint8_t* dst = const_cast<uint8_t*>(d_text);
        const uint64_t piece = std::max<uint64_t>(((len / 8 + TILE_BYTES) / TILE_BYTES) * TILE_BYTES, 256 * (uint64_t)TILE_BYTES);   // a multiple of both tile sizes
        CK(cudaEventRecord(c->ev[5], c->copy_stream));
        uint32_t k = 0;
        for (uint64_t off = 0; off < len || k == 0; off += piece, ++k) {
            const uint64_t n = std::min<uint64_t>(piece, len - off), n_copy = std::min<uint64_t>(n + 16, len - off);
            if (n_copy) CK(cudaMemcpyAsync(dst + off, host + off, n_copy, cudaMemcpyHostToDevice, c->copy_stream));
            if (k >= c->chunk_ev.size()) { cudaEvent_t e; CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); c->chunk_ev.push_back(e); }
            CK(cudaEventRecord(c->chunk_ev[k], c->copy_stream));
            CK(cudaStreamWaitEvent(st, c->chunk_ev[k], 0));
            fa.tile_begin = (uint32_t)(off / tile_bytes);
            fa.tile_end = off + n >= len ? n_tiles : (uint32_t)((off + n) / tile_bytes);
            if (fused) launch_fscan(fa.tile_begin, fa.tile_end);
            else k_filter_count<<<std::min<uint32_t>(fa.tile_end - fa.tile_begin, (uint32_t)c->n_sm * 3), SCAN_THREADS, smem, st>>>(fa);
            if (len == 0) break;
        }
        CK(cudaEventRecord(c->ev[8], c->copy_stream));
        c->launches += k - 1;
    } else {
        fa.tile_begin = 0; fa.tile_end = n_tiles;
        if (fused) launch_fscan(0, n_tiles);
        else k_filter_count<<<std::min<uint32_t>(n_tiles, (uint32_t)c->n_sm * 3), SCAN_THREADS, smem, st>>>(fa);
    }
    CK(cudaEventRecord(c->ev[2], st));
    CK(cudaMemcpyAsync(c->log_off.p, h_off, (n_logs + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(c->recs.p, 0xFF, rec_cap * 8, st));            // padding sentinels sort last
    CK(cudaMemsetAsync(c->ev_key_alt.p, 0xFF, ev_cap * 8, st));
    if (!c->plan.ctx_any_in_groups || fused) CK(cudaMemsetAsync(c->ctx_flags.p, 0, (line_cap / 4 + 2) * 4, st));   // else k_dfa_scan (walking every line) writes every byte
    k_scan_tiles<<<1, 1024, n_tiles <= 16384 ? (size_t)n_tiles * 4 : 0, st>>>(c->tile_counts.as<uint32_t>(), nullptr, n_tiles, n_logs, ctr);
    if (fused) {
        ExpandArgs ea{};
        ea.text = d_text; ea.len = len; ea.n_tiles = n_tiles; ea.tile_prefix = c->tile_counts.as<uint32_t>(); ea.warp_counts = c->fs_lists.as<uint16_t>(); ea.nl_lists = c->fs_lists.as<uint16_t>() + (size_t)n_tiles * 16;
        ea.log_off = c->log_off.as<uint64_t>(); ea.n_logs = n_logs;
        ea.line_start = c->line_start.as<uint32_t>(); ea.line_end = c->line_end.as<uint32_t>();
        ea.first_line = c->first_line.as<uint32_t>(); ea.line_cap = (uint32_t)line_cap; ea.ctr = ctr;
        ea.line_skip = c->plan.groups.empty() ? nullptr : c->line_skip.as<uint8_t>();
        ea.scan_list = c->plan.groups.empty() ? nullptr : c->scan_list.as<uint32_t>();
        for (int w = 0; w < 8; ++w) ea.skip1[w] = fz.skip1[w];
        k_expand<<<std::min<uint32_t>(n_tiles, (uint32_t)c->n_sm * 16), EX_THREADS, 0, st>>>(ea);
    } else {
    ScanArgs sa{};
    sa.text = d_text; sa.len = len; sa.n_tiles = n_tiles; sa.tile_prefix = c->tile_counts.as<uint32_t>();
    sa.log_off = c->log_off.as<uint64_t>(); sa.n_logs = n_logs;
    sa.line_start = c->line_start.as<uint32_t>(); sa.line_end = c->line_end.as<uint32_t>();
    sa.first_line = c->first_line.as<uint32_t>(); sa.line_cap = (uint32_t)line_cap; sa.bitmap = c->dp.bitmap;
    sa.cand = nullptr; sa.cand_cap = 0;
    sa.ctr = ctr; sa.use_filter = 0;      // offsets only: the prefilter ran in k_filter_count
    k_lines<<<std::min<uint32_t>(n_tiles, (uint32_t)c->n_sm * 4), LN_TILE_THREADS, 0, st>>>(sa);
    }
    c->launches += 2;
    k_log_meta<<<std::min<uint32_t>(n_logs, (uint32_t)c->n_sm * 8), LM_THREADS, 0, st>>>(c->first_line.as<uint32_t>(), c->line_start.as<uint32_t>(),
                                                         c->line_end.as<uint32_t>(), n_logs, (uint32_t)line_cap, rg ? 1u : 0u, c->n_lines.as<uint32_t>());
    CK(cudaEventRecord(c->ev[9], st));
    c->launches += 2;

    VerifyArgs va{};
    va.text = d_text; va.len = len; va.line_start = c->line_start.as<uint32_t>(); va.line_end = c->line_end.as<uint32_t>();
    va.tasks = c->tasks.as<uint4>(); va.task_cap = (uint32_t)std::min<size_t>(c->est_tasks, 0xFFFFFFFFu);
    va.recs = c->recs.as<unsigned long long>(); va.rec_cap = (uint32_t)rec_cap; va.ctr = ctr; va.n_lines_total = (uint32_t)line_cap;
    va.first_line = c->first_line.as<uint32_t>(); va.n_lines = c->n_lines.as<uint32_t>(); va.n_logs = n_logs;
    va.vm_tasks = c->vm_tasks.as<unsigned long long>(); va.vm_task_cap = (uint32_t)std::min<size_t>(c->est_vm_tasks, 0xFFFFFFFFu);
    va.vm_stack = c->vm_stack.as<uint32_t>(); va.vm_pool_n = VM_POOL_STACKS;
    if (use_filter) {
        ConfirmArgs ca{};
        ca.text = d_text; ca.len = len; ca.cand = c->cand.as<uint32_t>(); ca.cand_cap = fa.cand_cap;
        ca.items = c->items.as<unsigned long long>(); ca.item_cap = (uint32_t)std::min<size_t>(c->est_items, 0xFFFFFFFFu); ca.ctr = ctr;
        ResolveArgs ra{};
        ra.text = d_text; ra.len = len; ra.items = ca.items; ra.item_cap = ca.item_cap;
        ra.tile_prefix = c->tile_counts.as<uint32_t>(); ra.n_tiles = n_tiles; ra.tile_bytes = tile_bytes;
        ra.span_counts = fused ? c->fs_lists.as<uint16_t>() : nullptr; ra.log_off = c->log_off.as<uint64_t>(); ra.n_logs = n_logs;
        ra.line_start = c->line_start.as<uint32_t>(); ra.line_end = c->line_end.as<uint32_t>(); ra.n_lines_total = (uint32_t)line_cap;
        ra.tasks = c->tasks.as<uint4>(); ra.task_cap = va.task_cap; ra.ctr = ctr; ra.budget = c->anchor_budget;
        ra.vm_tasks = c->vm_tasks.as<unsigned long long>(); ra.vm_task_cap = va.vm_task_cap;
        ra.recs = c->recs.as<unsigned long long>(); ra.rec_cap = (uint32_t)rec_cap;
        // grids of exactly one resident wave (grid-stride kernels; a partial second wave would run alone at the end)
        if (!c->occ_confirm) {
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c->occ_confirm, k_confirm, 256, 0));
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c->occ_resolve, k_resolve, 256, 0));
            c->occ_confirm = std::max(1, c->occ_confirm); c->occ_resolve = std::max(1, c->occ_resolve);
            // (fewer blocks per SM measured slower for both, and k_resolve squeezed to 40 registers for 6 blocks much slower)
            if (const char* v = getenv("LP_CONFIRM_BLOCKS")) c->occ_confirm = std::max(1, std::min(c->occ_confirm, atoi(v)));
            if (const char* v = getenv("LP_RESOLVE_BLOCKS")) c->occ_resolve = std::max(1, std::min(c->occ_resolve, atoi(v)));
        }
        k_confirm<<<c->n_sm * c->occ_confirm, 256, 0, st>>>(ca, c->dp);
        k_resolve<<<c->n_sm * c->occ_resolve, 256, 0, st>>>(ra, c->dp);
        c->launches += 1;
        // bytecode tasks are grouped by regex before they are interpreted (see k_group_count)
        const uint32_t bins = c->dp.n_rx + 1;
        uint32_t* const gh = c->group_hist.as<uint32_t>();
        auto group_by_regex = [&](const unsigned long long* in, const unsigned int* n_ptr, uint32_t cap, uint32_t* hist, unsigned long long* out) {
            k_group_count<<<c->n_sm * 2, 256, 0, st>>>(in, n_ptr, cap, c->dp.n_rx, hist);
            k_group_scan<<<1, 1024, 0, st>>>(hist, bins, hist + bins);
            k_group_scatter<<<c->n_sm * 2, 256, 0, st>>>(in, n_ptr, cap, c->dp.n_rx, hist + bins, out);
            c->launches += 3;
        };
        if (has_vm) CK(cudaMemsetAsync(gh, 0, (size_t)bins * 4 * 4, st));
        k_verify<<<c->n_sm * 8, 256, 0, st>>>(va, c->dp);
        c->launches += 2;
        if (has_vm) {
            // the lanes of a warp then interpret the same program and stay together far longer than with tasks in arrival order
            group_by_regex(c->vm_tasks.as<unsigned long long>(), &ctr->n_vm_tasks, va.vm_task_cap, gh + 2 * bins, c->vm_tasks_alt.as<unsigned long long>());
            va.vm_tasks = c->vm_tasks_alt.as<unsigned long long>();
            k_verify_vm<<<vm_grid, vm_block, 0, st>>>(va, c->dp);
            c->launches += 1;
        }
    }
    if (n_always) { k_verify_always<<<std::min<uint32_t>(blocks_for(line_cap, 256), c->n_sm * 8), 256, 0, st>>>(va, c->dp); c->launches += 1; }
    if (n_always_vm) { k_verify_vm_always<<<vm_grid, vm_block, 0, st>>>(va, c->dp); c->launches += 1; }
    // always-on product-DFA scan (context regexes + literal-free DFA patterns)
    CK(cudaEventRecord(c->ev[6], st));
    for (size_t g0 = 0; g0 < c->d_group_blob.size(); g0 += DS_MAXG) {
        DfaScanArgs da{};
        da.text = d_text; da.len = len;
        da.line_start = c->line_start.as<uint32_t>(); da.line_end = c->line_end.as<uint32_t>(); da.n_lines_total = (uint32_t)line_cap;
        da.first_line = c->first_line.as<uint32_t>(); da.n_lines = c->n_lines.as<uint32_t>(); da.n_logs = n_logs;
        size_t tab_bytes = 0;
        da.n_groups = (uint32_t)std::min<size_t>(DS_MAXG, c->d_group_blob.size() - g0);
        for (int w = 0; w < 8; ++w) da.skip1[w] = 0xFFFFFFFFu;
        da.line_skip = fused ? c->line_skip.as<uint8_t>() : nullptr;      // the flags hold for every group (k_fscan used the intersection)
        da.scan_list = fused ? c->scan_list.as<uint32_t>() : nullptr;
        for (uint32_t q = 0; q < da.n_groups; ++q) {
            for (int w = 0; w < 8; ++w) da.skip1[w] &= c->plan.groups[g0 + q].skip1[w];
            da.blob[q] = c->d_group_blob[g0 + q]; da.blob_words[q] = c->group_words[g0 + q];
            tab_bytes += ((size_t)c->group_words[g0 + q] + 3) / 4 * 16;
        }
        da.rx_record = c->d_rx_record; da.ctx_flags = c->ctx_flags.as<uint8_t>();
        da.write_flags = c->plan.ctx_any_in_groups ? (g0 == 0 ? 1 : 2) : 0;
        da.recs = c->recs.as<unsigned long long>(); da.rec_cap = (uint32_t)rec_cap; da.gb = c->dp.gb; da.ctr = ctr;
        const uint32_t per_sm = (uint32_t)std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / std::max<size_t>(tab_bytes, 1)));
        const uint32_t grid = std::min<uint32_t>(blocks_for(line_cap, DS_THREADS), (uint32_t)c->n_sm * per_sm);
        switch (da.n_groups) {
            case 1: k_dfa_scan<1><<<grid, DS_THREADS, tab_bytes, st>>>(da); break;
            case 2: k_dfa_scan<2><<<grid, DS_THREADS, tab_bytes, st>>>(da); break;
            case 3: k_dfa_scan<3><<<grid, DS_THREADS, tab_bytes, st>>>(da); break;
            default: k_dfa_scan<4><<<grid, DS_THREADS, tab_bytes, st>>>(da); break;
        }
        c->launches += 1;
    }
    CK(cudaEventRecord(c->ev[7], st));
    CK(cudaEventRecord(c->ev[3], st));

    // ---- sort + unique the (regex, line) records; rec_cap slots are sorted, unused ones are sentinels
    {
        // (the all-ones padding has regex field 2^bits - 1 > n_rx - 1: it sorts last without an extra bit)
        const int rec_bits = std::min<int>(64, (int)c->dp.gb + bits_for(c->dp.n_rx));
        // small requests (C1: one 10k-line log; a shard of a log cut over 8 GPUs): tile sort + counting rank, two launches
        // instead of the device radix sort's six, which are pure launch and look-back latency at this size
        if (rec_cap > 8 * (size_t)SS_TILE ||
            !small_sort(c, c->recs.as<unsigned long long>(), (uint32_t)rec_cap, rec_bits, c->recs_alt.as<unsigned long long>(), nullptr, nullptr))
            sort_keys<unsigned long long>(c, c->recs.as<unsigned long long>(), c->recs_alt.as<unsigned long long>(), (uint32_t)rec_cap, rec_bits);
        size_t tmp = 0;
        cub::DeviceSelect::Unique(nullptr, tmp, c->recs_alt.as<unsigned long long>(), c->recs.as<unsigned long long>(),
                                  &ctr->n_unique, (int)rec_cap, st);
        c->cub_tmp.ensure(tmp);
        CK(cub::DeviceSelect::Unique(c->cub_tmp.p, tmp, c->recs_alt.as<unsigned long long>(), c->recs.as<unsigned long long>(),
                                     &ctr->n_unique, (int)rec_cap, st));
        c->launches += 2;
    }
    if (c->after_unique) c->after_unique();
    k_rx_bounds<<<blocks_for((uint64_t)(c->dp.n_rx + 1) * 32, 128), 128, 0, st>>>(c->recs.as<unsigned long long>(), ctr, c->dp.n_rx, c->dp.gb, c->rx_begin.as<uint32_t>());
    c->launches += 1;
    if (!c->plan.ctx_in_groups) {
        k_ctx_flags<<<c->n_sm * 4, 256, 0, st>>>(c->recs.as<unsigned long long>(), c->rx_begin.as<uint32_t>(), c->dp.gb, c->ctx_flags.as<uint32_t>());
        c->launches += 1;
    }
    // ---- events in discovery order (global line, pattern index)
    k_emit_events<<<c->n_sm * 4, 256, 0, st>>>(c->recs.as<unsigned long long>(), c->dp, c->ev_key_alt.as<unsigned long long>(), (uint32_t)ev_cap, ctr,
                                              rg ? rg->halo_before : 0u, rg ? rg->halo_after : 0u);
    if (!small_sort(c, c->ev_key_alt.as<unsigned long long>(), (uint32_t)ev_cap, std::min<int>(64, (int)c->dp.gb + (int)c->dp.pb + 1),
                    c->ev_key.as<unsigned long long>(), nullptr, nullptr))
        sort_keys<unsigned long long>(c, c->ev_key_alt.as<unsigned long long>(), c->ev_key.as<unsigned long long>(), (uint32_t)ev_cap,
                                      std::min<int>(64, (int)c->dp.gb + (int)c->dp.pb + 1));
    c->launches += 1;

    c->pend.valid = true;
    c->pend.d_text = d_text; c->pend.len = len; c->pend.n_logs = n_logs; c->pend.G = (uint32_t)line_cap;
    c->pend.m = (uint32_t)ev_cap; c->pend.options = options;
    if (rg) { c->pend.range = true; c->pend.halo_before = rg->halo_before; c->pend.halo_after = rg->halo_after; }
}

// Counts per pattern for the two-phase (multi-GPU) API; synchronises and validates the estimates.
void per_pattern_counts(lp_ctx* c, uint64_t* counts_out) {
    cudaStream_t st = c->stream;
    const uint32_t np = c->dp.n_pat;
    Counters* ctr = c->counters.as<Counters>();
    c->freq_delta.ensure((size_t)std::max<uint32_t>(np, 1) * 8);
    CK(cudaMemsetAsync(c->freq_delta.p, 0, (size_t)std::max<uint32_t>(np, 1) * 8, st));
    k_count_per_pattern<<<c->n_sm * 4, 256, 0, st>>>(c->recs.as<unsigned long long>(), ctr, c->dp, c->freq_delta.as<unsigned long long>(),
                                                    c->pend.range ? c->pend.halo_before : 0u, c->pend.range ? c->pend.halo_after : 0u);
    c->launches += 1;
    Counters h{};
    CK(cudaMemcpyAsync(&h, ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st));
    if (np) CK(cudaMemcpyAsync(counts_out, c->freq_delta.p, (size_t)np * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    check_estimates(c, h, c->pend.n_logs, true, c->pend.options);
}

// The one synchronisation of a call. The pipeline runs for well under a millisecond on typical logs, so the
// calling thread polls the stream for up to 1 ms (a blocking wait costs 10-20 us of wake-up latency) and falls
// back to a blocking wait when the work takes longer (a log that is still being copied from the host).
void wait_for_stream(cudaStream_t st) {
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
        const cudaError_t q = cudaStreamQuery(st);
        if (q == cudaSuccess) return;
        if (q != cudaErrorNotReady) CK(q);
        if (std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(1)) break;
    }
    CK(cudaStreamSynchronize(st));
}

// ---- phase 2: frequency ranks, scoring, ranking, result download (one synchronisation)
struct RangeGlobal {
    long long line_base, n_global; const long long* seq_T;        // seq_T: host array [n_seq]
    const long long* dev_range = nullptr;                         // or: {line_base, n_global, error} and seq_T composed on the device (k_range_compose)
    const long long* dev_seq_T = nullptr;
};

lp_result* run_score(lp_ctx* c, const uint64_t* prior_count, const RangeGlobal* rgl = nullptr,
                     const unsigned long long* prior_dev = nullptr, const unsigned long long* flag_dev = nullptr,
                     unsigned long long* flag_out = nullptr, cudaEvent_t wait_before_score = nullptr) {
    if (!c->pend.valid) throw std::invalid_argument("no scan is pending on this context");
    cudaStream_t st = c->stream;
    const uint32_t mc = c->pend.m, n_logs = c->pend.n_logs, line_cap = c->pend.G, np = c->dp.n_pat;
    const int options = c->pend.options;
    Counters* ctr = c->counters.as<Counters>();

    c->freq_delta.ensure((size_t)std::max<uint32_t>(np, 1) * 8);
    CK(cudaMemsetAsync(c->freq_delta.p, 0, (size_t)std::max<uint32_t>(np, 1) * 8, st));
    c->log_event_begin.ensure((size_t)(n_logs + 1) * 8);
    c->ev_log.ensure((size_t)mc * 4); c->slot_key.ensure((size_t)mc * 4); c->slot_key_alt.ensure((size_t)mc * 4);
    c->ev_idx.ensure((size_t)mc * 4); c->ev_idx_alt.ensure((size_t)mc * 4); c->rank_in_slot.ensure((size_t)mc * 4);
    // Everything the host reads back lives in ONE device arena with the layout of the page-locked buffer it is
    // copied to, so a call ends with a single device->host copy (small copies cost ~4 us each on the stream).
    const bool want_rank = !(options & LP_OPT_NO_RANK), want_factors = (options & LP_OPT_FACTORS) != 0, want_spans = (options & LP_OPT_SPANS) != 0;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 63) & ~size_t(63); return o; };
    const size_t o_ctr = take(sizeof(Counters)), o_line = take((size_t)mc * 4), o_pat = take((size_t)mc * 4), o_score = take((size_t)mc * 8),
                 o_rank = take((size_t)mc * 4), o_fac = take(want_factors ? (size_t)mc * 56 : 0),
                 o_delta = take((size_t)np * 8), o_nl = take((size_t)n_logs * 4), o_leb = take((size_t)(n_logs + 1) * 8),
                 o_ss = take(want_spans ? (size_t)mc * 4 : 0), o_se = take(want_spans ? (size_t)mc * 4 : 0), o_flag = take(32);
    c->out_arena.ensure(off);
    uint8_t* const arena = c->out_arena.as<uint8_t>();
    uint32_t* const d_rank = reinterpret_cast<uint32_t*>(arena + o_rank);
    c->sort_key.ensure((size_t)mc * 8); c->sort_key_alt.ensure((size_t)mc * 8);
    c->order.ensure((size_t)mc * 4); c->order_alt.ensure((size_t)mc * 4);
    c->log_sorted.ensure((size_t)mc * 4); c->log_sorted_alt.ensure((size_t)mc * 4);
    const unsigned long long* prior_d = nullptr;
    if (prior_count && np) {
        c->prior.ensure((size_t)np * 8);
        CK(cudaMemcpyAsync(c->prior.p, prior_count, (size_t)np * 8, cudaMemcpyHostToDevice, st));
        prior_d = c->prior.as<unsigned long long>();
    }
    if (prior_dev) prior_d = prior_dev;
    ScoreArgs sa{};
    sa.ev_key = c->ev_key.as<unsigned long long>(); sa.m_cap = mc; sa.ctr = ctr; sa.ev_log = c->ev_log.as<uint32_t>();
    sa.ev_idx = c->ev_idx.as<uint32_t>(); sa.freq_delta = c->freq_delta.as<unsigned long long>(); sa.n_logs = n_logs;
    sa.rank_in_slot = c->rank_in_slot.as<uint32_t>(); sa.prior = prior_d;
    sa.first_line = c->first_line.as<uint32_t>(); sa.n_lines = c->n_lines.as<uint32_t>();
    sa.keys = c->recs.as<unsigned long long>(); sa.rx_begin = c->rx_begin.as<uint32_t>();
    sa.ctx_flags = c->ctx_flags.as<uint8_t>();
    sa.out_line = reinterpret_cast<uint32_t*>(arena + o_line); sa.out_pat = reinterpret_cast<uint32_t*>(arena + o_pat);
    sa.out_score = reinterpret_cast<double*>(arena + o_score);
    sa.out_factors = want_factors ? reinterpret_cast<double*>(arena + o_fac) : nullptr;
    sa.sort_key = c->sort_key.as<unsigned long long>();
    if (rgl && rgl->dev_range) {
        sa.range_dev = rgl->dev_range; sa.n_global = 1; sa.core_lo = c->pend.halo_before;       // (n_global is read from range_dev by the kernel)
        if (c->n_seq) sa.seq_T = rgl->dev_seq_T;
    } else if (rgl) {
        sa.line_base = rgl->line_base; sa.n_global = rgl->n_global; sa.core_lo = c->pend.halo_before;
        if (c->n_seq) {
            c->seq_T.ensure((size_t)c->n_seq * 8);
            CK(cudaMemcpyAsync(c->seq_T.p, rgl->seq_T, (size_t)c->n_seq * 8, cudaMemcpyHostToDevice, st));
            sa.seq_T = c->seq_T.as<long long>();
        }
    }
    if (wait_before_score) CK(cudaStreamWaitEvent(st, wait_before_score, 0));     // the prior counts arrive on another stream
    // 8 lanes per event (cooperative searches: short dependent chains) while the events are few enough for one wave of
    // warps; with more events the kernel is bound by the number of warp-wide loads and a thread per event is faster
    // (C4: 4.6*10^4 events, score phase 0.35 ms with groups of 8, 0.29 ms with threads)
    if (mc <= 32768) k_score<8, 256><<<blocks_for(mc, 32), 256, 0, st>>>(sa, c->dp, c->sc);
    else k_score<1, 128><<<blocks_for(mc, 128), 128, 0, st>>>(sa, c->dp, c->sc);
    if (n_logs > 1 || mc > SS_MAX_TILES * (uint32_t)SS_TILE) {      // per-log event ranges: trivial for one log unless the large-sort ranking path needs them on the device
        k_log_event_begin<<<blocks_for(n_logs + 1, 128), 128, 0, st>>>(c->ev_key.as<unsigned long long>(), ctr, c->first_line.as<uint32_t>(),
                                                                       n_logs, c->dp.pb, c->log_event_begin.as<unsigned long long>());
        c->launches += 1;
    }
    c->launches += 1;
    if (!(options & LP_OPT_NO_RANK) && n_logs == 1 &&
        small_sort(c, c->sort_key.as<unsigned long long>(), mc, 64, nullptr, d_rank, &ctr->n_events)) {
        // one log, few events: ranks by tile sort + counting (descending score, ties keep discovery order)
    } else if (!(options & LP_OPT_NO_RANK)) {
        // stable sort by descending score (ties keep discovery order), then stably by log. The first step is the tile sort
        // + counting rank when the events fit it (a device-wide radix sort of a few ten thousand 64-bit keys is eight
        // passes of launch and look-back latency).
        if (small_sort(c, c->sort_key.as<unsigned long long>(), mc, 64, nullptr, c->order_alt.as<uint32_t>(), &ctr->n_events)) {
            k_invert_rank<<<blocks_for(mc, 256), 256, 0, st>>>(c->order_alt.as<uint32_t>(), ctr, mc, c->order.as<uint32_t>());
            c->launches += 1;
        } else
            sort_pairs<unsigned long long, uint32_t>(c, c->sort_key.as<unsigned long long>(), c->sort_key_alt.as<unsigned long long>(),
                                                     c->ev_idx.as<uint32_t>(), c->order.as<uint32_t>(), mc, 64);
        uint32_t* order = c->order.as<uint32_t>();
        if (n_logs > 1) {
            k_gather_u32<<<blocks_for(mc, 256), 256, 0, st>>>(c->ev_log.as<uint32_t>(), order, mc, c->log_sorted.as<uint32_t>());
            sort_pairs<uint32_t, uint32_t>(c, c->log_sorted.as<uint32_t>(), c->log_sorted_alt.as<uint32_t>(), order,
                                           c->order_alt.as<uint32_t>(), mc, bits_for(n_logs));
            order = c->order_alt.as<uint32_t>();
            c->launches += 1;
        }
        k_write_rank<<<blocks_for(mc, 256), 256, 0, st>>>(order, c->ev_log.as<uint32_t>(), c->log_event_begin.as<unsigned long long>(), ctr,
                                                          d_rank);
        c->launches += 1;
    }
    if (want_spans) {
        const uint32_t sg = (uint32_t)c->n_sm * 2, sb = 128;
        c->vm_stack.ensure((size_t)VM_POOL_STACKS * lpx::VM_STACK_WORDS * 4);
        k_spans<<<sg, sb, 0, st>>>(c->ev_key.as<unsigned long long>(), ctr, mc, c->dp, c->pend.d_text, c->line_start.as<uint32_t>(),
                                   c->line_end.as<uint32_t>(), c->vm_stack.as<uint32_t>(), VM_POOL_STACKS, ctr, reinterpret_cast<uint32_t*>(arena + o_ss),
                                   reinterpret_cast<uint32_t*>(arena + o_se));
        c->launches += 1;
    }
    CK(cudaEventRecord(c->ev[4], st));

    // ---- download into page-locked staging (asynchronous copies, one synchronisation), sizes from
    // the estimates; trimmed after the counters arrive
    std::unique_ptr<lp_result> res(new lp_result());
    res->n_logs = n_logs;
    // the small items (counters, per-pattern counts, per-log line counts and event ranges, status word) are
    // packed behind the event columns by one tiny kernel, then the arena goes to the host in one copy
    k_pack_tail<<<1, 256, 0, st>>>(ctr, c->freq_delta.as<unsigned long long>(), np, c->n_lines.as<uint32_t>(), n_logs,
                                   c->log_event_begin.as<unsigned long long>(), flag_dev, rgl ? rgl->dev_range : nullptr, reinterpret_cast<Counters*>(arena + o_ctr),
                                   reinterpret_cast<unsigned long long*>(arena + o_delta), reinterpret_cast<uint32_t*>(arena + o_nl),
                                   reinterpret_cast<unsigned long long*>(arena + o_leb), reinterpret_cast<unsigned long long*>(arena + o_flag));
    c->launches += 1;
    if (!c->stage.p) c->stage = c->pool->acquire();
    c->stage.ensure(off);
    uint8_t* hs = c->stage.p;
    CK(cudaMemcpyAsync(hs, arena, off, cudaMemcpyDeviceToHost, st));
    wait_for_stream(st);
    CK(cudaGetLastError());
    Counters h;
    memcpy(&h, hs + o_ctr, sizeof h);
    if (flag_dev && flag_out) memcpy(flag_out, hs + o_flag, 8);
    check_estimates(c, h, n_logs, true, options);       // throws Retry when a buffer was too small

    const uint32_t m = h.n_events;
    const uint32_t G = (uint32_t)(h.total_nl + n_logs);
    res->n_events = m;
    res->total_lines = G;
    res->budget_exceeded = h.n_budget;
    // the result takes the staging buffer over: its columns are read in place
    res->line = reinterpret_cast<uint32_t*>(hs + o_line);
    res->pattern = reinterpret_cast<uint32_t*>(hs + o_pat);
    res->score = reinterpret_cast<double*>(hs + o_score);
    if (want_rank) res->rank = reinterpret_cast<uint32_t*>(hs + o_rank);
    else { res->zero_rank.assign(m, 0); res->rank = res->zero_rank.data(); }
    if (want_factors) { res->has_factors = true; res->factors = reinterpret_cast<double*>(hs + o_fac); }
    if (want_spans) { res->span_start = reinterpret_cast<uint32_t*>(hs + o_ss); res->span_end = reinterpret_cast<uint32_t*>(hs + o_se); }
    res->freq_delta = reinterpret_cast<uint64_t*>(hs + o_delta);
    res->n_patterns = np;
    res->log_n_lines.resize(n_logs);
    for (uint32_t k = 0; k < n_logs; ++k) res->log_n_lines[k] = reinterpret_cast<uint32_t*>(hs + o_nl)[k];
    if (rgl && rgl->dev_range) {
        long long rw[3];
        memcpy(rw, hs + o_flag + 8, sizeof rw);
        if (rw[2] == 1) throw std::invalid_argument("halo_before exceeds the lines of the lower shards");
        if (rw[2] == 2) throw std::invalid_argument("shard summaries come from a different pattern set");
        res->log_n_lines[0] = (uint64_t)rw[1];
    } else if (rgl) res->log_n_lines[0] = (uint64_t)rgl->n_global;      // a shard reports the whole log's line count
    res->log_event_begin.assign(reinterpret_cast<uint64_t*>(hs + o_leb), reinterpret_cast<uint64_t*>(hs + o_leb) + n_logs + 1);
    if (options & LP_OPT_LINE_OFFSETS) {
        res->has_offsets = true;
        res->line_start.resize((size_t)G + 1);
        res->line_end.resize((size_t)G);
        CK(cudaMemcpyAsync(res->line_start.data(), c->line_start.p, (size_t)G * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(res->line_end.data(), c->line_end.p, (size_t)G * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        res->line_start[G] = (uint32_t)c->pend.len;
    }
    (void)line_cap;
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev[0], c->ev[4]); c->tm.total_ms = ms;
    cudaEventElapsedTime(&ms, c->ev[1], c->ev[2]); c->tm.scan_ms = ms;
    cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]); c->tm.verify_ms = ms;
    cudaEventElapsedTime(&ms, c->ev[3], c->ev[4]); c->tm.score_ms = ms;
    cudaEventElapsedTime(&ms, c->ev[6], c->ev[7]); c->tm.dfa_scan_ms = ms;
    cudaEventElapsedTime(&ms, c->ev[2], c->ev[9]); c->tm.lines_ms = ms;
    c->tm.kernel_launches = c->launches;
    c->tm.candidates = h.n_cand;
    c->tm.verify_tasks = (uint64_t)h.n_tasks + h.n_vm_tasks + (uint64_t)G * (c->plan.always_list.size() + c->plan.always_vm_list.size());
    // k_filter_count: the log read once, one count per tile and one position per candidate written
    // algorithmic bytes of the scan kernel timed as scan_ms. k_fscan: the log read once; written: one 2-byte list entry per
    // newline, 1 + 16 counts per tile, one position per candidate. Legacy k_filter_count: log, tile counts, candidates.
    if (c->scan_mode == 0) c->tm.scan_bytes = c->pend.len + 2ull * h.total_nl + 36ull * ((c->pend.len + fs::TILE) / fs::TILE) + 4ull * h.n_cand;
    else c->tm.scan_bytes = c->pend.len + 4ull * ((c->pend.len + TILE_BYTES) / TILE_BYTES) + 4ull * h.n_cand;
    c->tm.dfa_scan_bytes = c->d_group_blob.empty() ? 0 : c->pend.len + 8ull * ((uint64_t)G + 1) + (uint64_t)G;
    // The filter is tuned once, against the first log. It is tuned again (at most twice) only when logs arrive that
    // produce several times the candidates per line of the log it was tuned on.
    if (!c->plan.lit_off.empty() && c->calibrated && G > 0) {
        const double rate = (double)h.n_cand / (double)G;
        if (c->cal_rate < 0) c->cal_rate = rate;
        else if (rate > 4.0 * c->cal_rate + 0.5 && c->recalibrations < 2) { c->calibrated = false; ++c->recalibrations; }
    }
    c->pend.valid = false;
    res->hold = c->stage;          // ownership moves to the result; the next call draws a buffer from the pool
    res->pool = c->pool;
    c->stage = PinnedBuf();
    return res.release();
}

// scan + score with the re-run loop for under-estimated buffers
template <class Stage>
lp_result* run_all(lp_ctx* c, Stage&& stage_text, uint64_t len, const uint64_t* offsets, uint32_t n_logs, const uint64_t* prior, int options) {
    for (;;) {
        try {
            const uint8_t* d = stage_text();
            run_scan(c, d, len, offsets, n_logs, options);
            return run_score(c, prior);
        } catch (const Retry&) {
            continue;
        } catch (...) {
            c->host_src = nullptr;      // never leave a caller's host pointer behind
            throw;
        }
    }
}

template <class F>
int guarded(lp_ctx* c, F&& f) {
    try {
        if (c) CK(cudaSetDevice(c->device));
        f();
        return LP_OK;
    } catch (const lpx::RegexSyntaxError& e) {
        g_last_error = std::string("PatternSyntaxException: ") + e.what() + " near index " + std::to_string(e.index);
        return LP_ERR_SYNTAX;
    } catch (const lpx::RegexUnsupported& e) {
        g_last_error = std::string("unsupported regex construct: ") + e.what();
        return LP_ERR_UNSUPPORTED;
    } catch (const CudaFail& e) {
        g_last_error = e.what();
        return LP_ERR_CUDA;
    } catch (const LimitFail& e) {
        g_last_error = e.what();
        return LP_ERR_LIMIT;
    } catch (const std::overflow_error& e) {
        g_last_error = e.what();
        return LP_ERR_BACKTRACK;
    } catch (const std::invalid_argument& e) {
        g_last_error = e.what();
        return LP_ERR_INVALID;
    } catch (const std::exception& e) {
        g_last_error = e.what();
        return LP_ERR_INVALID;
    }
}

const uint8_t* stage_host_log(lp_ctx* c, const uint8_t* host, uint64_t len) {
    c->text.ensure(len + 64);
    c->host_src = nullptr;
    if (c->calibrated && !c->plan.lit_off.empty() && len >= (8u << 20)) {
        c->host_src = host;             // run_scan copies it piecewise, overlapped with the prefilter scan
        return c->text.as<uint8_t>();
    }
    CK(cudaEventRecord(c->ev[5], c->stream));
    if (len) CK(cudaMemcpyAsync(c->text.p, host, len, cudaMemcpyHostToDevice, c->stream));
    CK(cudaEventRecord(c->ev[8], c->stream));
    return c->text.as<uint8_t>();
}

// A free worker of the context (itself or one of its extras), locked; when all are busy, the next one in turn.
struct WorkerLock {
    lp_ctx* w;
    std::unique_lock<std::mutex> lk;
};
WorkerLock pick_worker(lp_ctx* c) {
    const size_t n = 1 + c->extra.size();
    for (size_t i = 0; i < n; ++i) {            // the context itself first: a single caller always lands on the same, warm scratch
        lp_ctx* w = i == 0 ? c : c->extra[i - 1];
        std::unique_lock<std::mutex> lk(w->mu, std::try_to_lock);
        if (lk.owns_lock()) return WorkerLock{w, std::move(lk)};
    }
    const uint32_t turn = c->next_worker.fetch_add(1, std::memory_order_relaxed) % n;
    lp_ctx* w = turn == 0 ? c : c->extra[turn - 1];
    return WorkerLock{w, std::unique_lock<std::mutex>(w->mu)};
}

}  // namespace

// ============================================================================================
//                                         C ABI
// ============================================================================================
extern "C" {

const char* lp_last_error(void) { return g_last_error.c_str(); }
const char* lp_version(void) { return "logparser_b200 0.1 (sm_100a)"; }

static lp_ctx* create_worker(const lp_scoring_cfg* cfg, int device) {
    std::unique_ptr<lp_ctx> ctx(new lp_ctx());
    ctx->owner = ctx.get();
    ctx->device = device;
    ctx->cfg = *cfg;
    ctx->cfg.decay_lut = nullptr;
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (auto& e : ctx->ev) CK(cudaEventCreate(&e));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    ctx->n_sm = prop.multiProcessorCount;
    if (const char* m = getenv("LP_SCAN_MODE")) ctx->scan_mode = strcmp(m, "legacy") == 0 ? 1 : 0;
    if (const char* v = getenv("LP_ANCHOR_BUDGET")) ctx->anchor_budget = (uint32_t)std::max(1, std::min(4096, atoi(v)));
    if (const char* v = getenv("LP_VM_BLOCKS")) ctx->vm_blocks_per_sm = (uint32_t)std::max(1, std::min(16, atoi(v)));
    if (const char* v = getenv("LP_FSCAN_VARIANT")) ctx->fscan_variant = std::max(0, std::min(3, atoi(v)));
    std::vector<double> lut((size_t)cfg->max_window + 1);
    for (int d = 0; d <= cfg->max_window; ++d)
        lut[d] = cfg->decay_lut ? cfg->decay_lut[d] : std::exp(-(double)d / cfg->decay_constant);   // SS:177
    ctx->d_lut.ensure(lut.size() * 8);
    CK(cudaMemcpy(ctx->d_lut.p, lut.data(), lut.size() * 8, cudaMemcpyHostToDevice));
    ScoreCfg& s = ctx->sc;
    s.early = cfg->early_bonus_threshold; s.max_early = cfg->max_early_bonus; s.penalty = cfg->penalty_threshold;
    s.max_ctx = cfg->max_context_factor; s.freq_thr = cfg->frequency_threshold; s.max_pen = cfg->max_penalty;
    s.max_window = cfg->max_window; s.hours = cfg->time_window_hours; s.decay_lut = ctx->d_lut.as<double>();
    return ctx.release();
}

lp_ctx* lp_create(const lp_scoring_cfg* cfg, int device) {
    if (device == -1) {     // every visible device
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) { g_last_error = "no CUDA device visible"; return nullptr; }
        std::vector<int> all(n);
        for (int i = 0; i < n; ++i) all[i] = i;
        return lp_create_multi(cfg, all.data(), all.size());
    }
    lp_ctx* c = nullptr;
    int rc = guarded(nullptr, [&] {
        if (!cfg) throw std::invalid_argument("cfg is NULL");
        if (cfg->max_window < 0) throw std::invalid_argument("scoring.proximity.max-window < 0");
        if (cfg->time_window_hours <= 0) throw std::invalid_argument("scoring.frequency.time-window-hours <= 0");
        int n = 0;
        CK(cudaGetDeviceCount(&n));
        if (n <= 0) throw CudaFail("no CUDA device visible");
        if (device < 0 || device >= n) throw std::invalid_argument("device index out of range");
        CK(cudaSetDevice(device));
        CK(cudaFuncSetAttribute(k_filter_count, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(GRAM_WORDS * 4)));
        CK(cudaFuncSetAttribute(k_fscan<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fs::SMEM_BYTES));
        CK(cudaFuncSetAttribute(k_fscan<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fs::SMEM_BYTES));
        CK(cudaFuncSetAttribute(k_fscan<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fs::SMEM_BYTES));
        CK(cudaFuncSetAttribute(k_fscan<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fs::SMEM_BYTES));
        CK(cudaFuncSetAttribute(k_scan_tiles, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 4));
        CK(cudaFuncSetAttribute(k_tile_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SsSorter::TempStorage)));
        CK(cudaFuncSetAttribute(k_dfa_scan<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
        CK(cudaFuncSetAttribute(k_dfa_scan<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
        CK(cudaFuncSetAttribute(k_dfa_scan<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
        CK(cudaFuncSetAttribute(k_dfa_scan<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
        std::unique_ptr<lp_ctx, void (*)(lp_ctx*)> ctx(create_worker(cfg, device), lp_destroy);
        int workers = 2;        // LP_WORKERS: requests that may be in flight at once on this context
        if (const char* w = getenv("LP_WORKERS")) workers = std::max(1, std::min(8, atoi(w)));
        for (int k = 1; k < workers; ++k) { ctx->extra.push_back(create_worker(cfg, device)); ctx->extra.back()->owner = ctx.get(); }
        c = ctx.release();
    });
    return rc == LP_OK ? c : nullptr;
}

lp_ctx* lp_create_multi(const lp_scoring_cfg* cfg, const int* devices, size_t n_devices) {
    if (!devices || n_devices == 0) { g_last_error = "no devices given"; return nullptr; }
    lp_ctx* c = lp_create(cfg, devices[0]);
    if (!c) return nullptr;
    for (size_t i = 1; i < n_devices; ++i) {
        lp_ctx* d = devices[i] >= 0 ? lp_create(cfg, devices[i]) : nullptr;
        if (!d) { if (devices[i] < 0) g_last_error = "device index out of range"; lp_destroy(c); return nullptr; }
        c->shards.push_back(d);
    }
    return c;
}

static void nccl_comm_destroy(void* comm);
void lp_destroy(lp_ctx* c) {
    if (!c) return;
    for (lp_ctx* d : c->shards) lp_destroy(d);
    c->shards.clear();
    for (lp_ctx* e : c->extra) lp_destroy(e);
    c->extra.clear();
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    for (auto& b : c->plan_bufs) b.release();
    DevBuf* all[] = {&c->text, &c->log_off, &c->tile_counts, &c->counters, &c->line_start, &c->line_end, &c->first_line, &c->n_lines,
                     &c->cand, &c->tasks, &c->group_hist, &c->vm_tasks, &c->vm_tasks_alt, &c->vm_stack, &c->recs, &c->recs_alt, &c->cub_tmp, &c->rx_begin, &c->ctx_flags, &c->ev_key, &c->ev_key_alt,
                     &c->ev_log, &c->slot_key, &c->slot_key_alt, &c->ev_idx, &c->ev_idx_alt, &c->rank_in_slot, &c->freq_delta, &c->prior,
                     &c->out_arena, &c->sort_key, &c->sort_key_alt, &c->order, &c->order_alt,
                     &c->log_sorted, &c->log_sorted_alt, &c->log_event_begin, &c->n_unique, &c->d_lut,
                     &c->f_bitmap, &c->f_key, &c->f_lits, &c->f_bytes, &c->f_bits, &c->f_gkeys, &c->fs_lists, &c->scan_list, &c->line_skip, &c->items, &c->ss_keys, &c->ss_idx, &c->seq_tab_off, &c->seq_tab, &c->seq_T, &c->prior_eff, &c->any_flag, &c->dist_mine, &c->dist_all, &c->range_dev};
    for (DevBuf* b : all) b->release();
    if (c->nccl_comm) nccl_comm_destroy(c->nccl_comm);
    c->stage.release();
    c->pool->close();
    for (auto& e : c->ev) if (e) cudaEventDestroy(e);
    for (auto& e : c->chunk_ev) cudaEventDestroy(e);
    for (auto& e : c->dist_ev) if (e) cudaEventDestroy(e);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int lp_load_patterns(lp_ctx* c, const lp_pattern_desc* pats, size_t n, int64_t* bad_pattern) {
    if (bad_pattern) *bad_pattern = -1;
    if (!c) { g_last_error = "ctx is NULL"; return LP_ERR_INVALID; }
    for (lp_ctx* d : c->shards) {       // a multi-device context: the other devices first (a failure leaves the message of the first one)
        const int rc = lp_load_patterns(d, pats, n, bad_pattern);
        if (rc != LP_OK) return rc;
    }
    return guarded(c, [&] {
        if (n && !pats) throw std::invalid_argument("patterns is NULL");
        std::lock_guard<std::mutex> lk(c->mu);
        std::vector<lpx::PlanPatternIn> in = lpx::plan_input_from_desc(pats, n, bad_pattern);
        c->loaded = false;
        c->calibrated = false;
        c->recalibrations = 0;
        c->f_mul = fs::MUL1;
        c->cal_seen = 0; c->cal_rate = -1.0;
        { std::lock_guard<std::mutex> lc(c->shared_cal.mu); c->shared_cal.version = 0; }
        lpx::build_plan(in, c->plan, bad_pattern);
        upload_plan(c);
        // line-range shards: where each sequence's chain table sits in a shard summary
        c->n_seq = (uint32_t)c->plan.seq_bonus.size();
        c->h_seq_tab_off.assign(c->n_seq + 1, 0);
        for (uint32_t q = 0; q < c->n_seq; ++q) {
            const uint32_t k = c->plan.seq_ev_begin[q + 1] - c->plan.seq_ev_begin[q];
            c->h_seq_tab_off[q + 1] = c->h_seq_tab_off[q] + (k >= 2 ? k - 1 : 0);
        }
        upload_into(c->seq_tab_off, c->h_seq_tab_off);
        c->loaded = true;
        for (lp_ctx* e : c->extra) {       // the other workers get their own copy of the (small) tables
            std::lock_guard<std::mutex> le(e->mu);
            e->loaded = false; e->calibrated = false; e->recalibrations = 0; e->f_mul = fs::MUL1; e->cal_seen = 0; e->cal_rate = -1.0;
            e->plan = c->plan;
            upload_plan(e);
            e->n_seq = c->n_seq; e->h_seq_tab_off = c->h_seq_tab_off;
            upload_into(e->seq_tab_off, e->h_seq_tab_off);
            e->est_lines = e->est_cand = e->est_tasks = e->est_vm_tasks = e->est_recs = e->est_items = e->est_events = 0;
            e->loaded = true;
        }
    });
}

int lp_batch_scan_device(lp_ctx* c, const uint8_t* d_logs, const uint64_t* offsets, size_t n_logs, int options, uint64_t* counts_out) {
    if (!c) { g_last_error = "ctx is NULL"; return LP_ERR_INVALID; }
    return guarded(c, [&] {
        if (!offsets) throw std::invalid_argument("offsets is NULL");
        std::lock_guard<std::mutex> lk(c->mu);
        std::vector<uint64_t> tmp(std::max<uint32_t>(c->dp.n_pat, 1));
        for (;;) {
            try {
                run_scan(c, d_logs, offsets[n_logs], offsets, (uint32_t)n_logs, options);
                per_pattern_counts(c, tmp.data());
                break;
            } catch (const Retry&) { continue; }
        }
        if (counts_out) memcpy(counts_out, tmp.data(), (size_t)c->dp.n_pat * 8);
    });
}
int lp_batch_score(lp_ctx* c, const uint64_t* prior_count, lp_result** out) {
    if (!c || !out) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    return guarded(c, [&] {
        std::lock_guard<std::mutex> lk(c->mu);
        try {
            *out = run_score(c, prior_count);
        } catch (const Retry&) {
            throw LimitFail("internal: estimates validated by lp_batch_scan_device were exceeded in lp_batch_score");
        }
    });
}

// ---- NCCL inside the library: scan -> ncclAllGather -> score in one call, all on the context's stream.
// libnccl.so.2 is resolved at run time (dlopen), so the library has no link-time dependency on it and a
// single-GPU host needs no NCCL at all; in a process that already loaded NCCL (torch) the same copy is used.
namespace {
struct NcclApi {
    typedef struct { char internal[128]; } UniqueId;
    int (*GetUniqueId)(UniqueId*) = nullptr;
    int (*CommInitRank)(void**, int, UniqueId, int) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
    std::string why;
    NcclApi() {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { why = std::string("libnccl.so.2 could not be loaded: ") + dlerror(); return; }
        GetUniqueId = reinterpret_cast<decltype(GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
        CommInitRank = reinterpret_cast<decltype(CommInitRank)>(dlsym(h, "ncclCommInitRank"));
        AllGather = reinterpret_cast<decltype(AllGather)>(dlsym(h, "ncclAllGather"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(dlsym(h, "ncclCommDestroy"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(dlsym(h, "ncclGetErrorString"));
        ok = GetUniqueId && CommInitRank && AllGather && CommDestroy && GetErrorString;
        if (!ok) why = "libnccl.so.2 lacks an expected symbol";
    }
};
NcclApi& nccl() { static NcclApi api; return api; }
void nccl_check(int rc, const char* what) {
    if (rc != 0) throw CudaFail(std::string(what) + ": " + (nccl().GetErrorString ? nccl().GetErrorString(rc) : "NCCL error"));
}
constexpr int NCCL_UINT64 = 5;      // ncclUint64 (nccl.h)
}  // namespace
static void nccl_comm_destroy(void* comm) { if (nccl().ok) nccl().CommDestroy(comm); }

int lp_comm_unique_id(uint8_t* id128) {
    if (!id128) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    return guarded(nullptr, [&] {
        if (!nccl().ok) throw CudaFail(nccl().why);
        NcclApi::UniqueId id;
        nccl_check(nccl().GetUniqueId(&id), "ncclGetUniqueId");
        memcpy(id128, id.internal, 128);
    });
}
int lp_comm_init(lp_ctx* c, const uint8_t* id128, uint32_t rank, uint32_t world) {
    if (!c || !id128) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    return guarded(c, [&] {
        std::lock_guard<std::mutex> lk(c->mu);
        if (rank >= world) throw std::invalid_argument("rank >= world");
        if (!nccl().ok) throw CudaFail(nccl().why);
        if (c->nccl_comm) { nccl().CommDestroy(c->nccl_comm); c->nccl_comm = nullptr; }
        NcclApi::UniqueId id;
        memcpy(id.internal, id128, 128);
        nccl_check(nccl().CommInitRank(&c->nccl_comm, (int)world, id, (int)rank), "ncclCommInitRank");
        c->comm_rank = rank; c->comm_world = world;
    });
}
// This rank's block of a batch that is sharded by log over the communicator of lp_comm_init: scan,
// all-gather of [per-pattern counts, status] over NVLink, exclusive prefix, score. Collective: every
// rank calls it once per step (with n_logs >= 1). A rank whose scratch estimates were too small makes
// all ranks repeat the step, which they agree on through the gathered status words.
static int analyze_batch_dist(lp_ctx* c, const uint8_t* logs, bool on_host, const uint64_t* offsets, size_t n_logs, const uint64_t* prior_count,
                              int options, lp_result** out) {
    if (!c || !out) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    *out = nullptr;
    return guarded(c, [&] {
        if (!offsets) throw std::invalid_argument("offsets is NULL");
        std::lock_guard<std::mutex> lk(c->mu);
        if (on_host && offsets[n_logs] && !logs) throw std::invalid_argument("logs is NULL");
        // host input: the copy to the device is cut in pieces and overlapped with the scan of the pieces already there
        const uint8_t* d_logs = on_host ? stage_host_log(c, logs, offsets[n_logs]) : logs;
        struct DropHostSrc { lp_ctx* c; ~DropHostSrc() { c->host_src = nullptr; } } drop{c};
        if (!c->nccl_comm) throw std::invalid_argument("lp_comm_init has not been called on this context");
        const uint32_t np = c->dp.n_pat, world = c->comm_world, rank = c->comm_rank;
        const size_t words = (size_t)np + 1;
        c->dist_mine.ensure(words * 8);
        c->dist_all.ensure(words * 8 * world);
        c->prior_eff.ensure((size_t)std::max<uint32_t>(np, 1) * 8);
        c->any_flag.ensure(8);
        for (int k = 0; k < 2; ++k) if (!c->dist_ev[k]) CK(cudaEventCreateWithFlags(&c->dist_ev[k], cudaEventDisableTiming));
        const unsigned long long* prior_in = nullptr;
        if (prior_count && np) {
            c->prior.ensure((size_t)np * 8);
            CK(cudaMemcpyAsync(c->prior.p, prior_count, (size_t)np * 8, cudaMemcpyHostToDevice, c->stream));
            prior_in = c->prior.as<unsigned long long>();
        }
        // The exchange starts as soon as the unique records exist and runs on the second stream, next to
        // the kernels that do not need the other ranks' counts (context flags, event list, its sort);
        // k_score waits for it. The status word covers every queue, the events queue through the sum of
        // the counts, so all ranks always agree on whether the step is repeated.
        struct Unhook { lp_ctx* c; ~Unhook() { c->after_unique = nullptr; } } unhook{c};
        c->after_unique = [&, c] {
            unsigned long long* mine = c->dist_mine.as<unsigned long long>();
            CK(cudaMemsetAsync(mine, 0, words * 8, c->stream));
            k_count_per_pattern<<<c->n_sm * 4, 256, 0, c->stream>>>(c->recs.as<unsigned long long>(), c->counters.as<Counters>(), c->dp, mine, 0u, 0u);
            k_publish_overflow<<<1, 128, 0, c->stream>>>(c->counters.as<Counters>(), est_caps(c), mine, np, mine + np);
            CK(cudaEventRecord(c->dist_ev[0], c->stream));
            CK(cudaStreamWaitEvent(c->copy_stream, c->dist_ev[0], 0));
            nccl_check(nccl().AllGather(mine, c->dist_all.p, words, NCCL_UINT64, c->nccl_comm, c->copy_stream), "ncclAllGather");
            k_prior_prefix<<<blocks_for(std::max<uint32_t>(np, 1), 128), 128, 0, c->copy_stream>>>(
                c->dist_all.as<unsigned long long>(), np, rank, world, prior_in, c->prior_eff.as<unsigned long long>(),
                c->any_flag.as<unsigned long long>(), c->dp);
            CK(cudaEventRecord(c->dist_ev[1], c->copy_stream));
            c->launches += 4;
        };
        for (int round = 0;; ++round) {
            if (round > 16) throw LimitFail("scratch estimates did not settle");
            run_scan(c, d_logs, offsets[n_logs], offsets, (uint32_t)n_logs, options);
            std::unique_ptr<lp_result> res;
            bool again = false;
            unsigned long long flag = 0;
            try {
                res.reset(run_score(c, nullptr, nullptr, c->prior_eff.as<unsigned long long>(), c->any_flag.as<unsigned long long>(), &flag,
                                    c->dist_ev[1]));
            } catch (const Retry&) { again = true; }
            if (again) {
                // this rank's own status word was non-zero, so the others repeat as well; read it to be sure
                CK(cudaMemcpy(&flag, c->any_flag.p, 8, cudaMemcpyDeviceToHost));
                if (!flag) throw LimitFail("internal: a repeat was requested that the other ranks cannot see");
                continue;
            }
            if ((flag & 16u) && (options & LP_OPT_STRICT_BUDGET)) throw std::overflow_error("backtracking verifier exceeded its step/stack budget");
            if (flag & ~16ull) continue;                 // every rank sees the same words: all of them repeat
            *out = res.release();
            if (on_host) {
                float ms = 0;
                cudaEventElapsedTime(&ms, c->ev[5], c->ev[8]);
                c->tm.h2d_ms = ms;
            }
            return;
        }
    });
}
int lp_analyze_batch_dist_device(lp_ctx* c, const uint8_t* d_logs, const uint64_t* offsets, size_t n_logs, const uint64_t* prior_count,
                                 int options, lp_result** out) {
    return analyze_batch_dist(c, d_logs, false, offsets, n_logs, prior_count, options, out);
}
int lp_analyze_batch_dist(lp_ctx* c, const uint8_t* logs, const uint64_t* offsets, size_t n_logs, const uint64_t* prior_count,
                          int options, lp_result** out) {
    return analyze_batch_dist(c, logs, true, offsets, n_logs, prior_count, options, out);
}

// Two-phase form with the exchange left on the device: nothing is read back between the phases.
int lp_batch_scan_device_async(lp_ctx* c, const uint8_t* d_logs, const uint64_t* offsets, size_t n_logs, int options, uint64_t* d_counts_out) {
    if (!c || !d_counts_out) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    return guarded(c, [&] {
        if (!offsets) throw std::invalid_argument("offsets is NULL");
        std::lock_guard<std::mutex> lk(c->mu);
        try {
            run_scan(c, d_logs, offsets[n_logs], offsets, (uint32_t)n_logs, options);
        } catch (const Retry&) { throw LimitFail("internal: run_scan does not synchronise"); }
        const uint32_t np = c->dp.n_pat;
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(d_counts_out);
        CK(cudaMemsetAsync(dst, 0, ((size_t)np + 1) * 8, c->stream));
        k_count_per_pattern<<<c->n_sm * 4, 256, 0, c->stream>>>(c->recs.as<unsigned long long>(), c->counters.as<Counters>(), c->dp, dst, 0u, 0u);
        k_publish_overflow<<<1, 128, 0, c->stream>>>(c->counters.as<Counters>(), est_caps(c), dst, np, dst + np);
        c->launches += 2;
        CK(cudaGetLastError());
    });
}
int lp_batch_score_gathered(lp_ctx* c, const uint64_t* d_all_counts, uint32_t rank, uint32_t world, const uint64_t* prior_count, lp_result** out) {
    if (!c || !out || !d_all_counts) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    *out = nullptr;
    bool again = false;
    const int rc = guarded(c, [&] {
        std::lock_guard<std::mutex> lk(c->mu);
        if (!c->pend.valid) throw std::invalid_argument("no scan is pending on this context");
        if (rank >= world) throw std::invalid_argument("rank >= world");
        const uint32_t np = c->dp.n_pat;
        c->prior_eff.ensure((size_t)std::max<uint32_t>(np, 1) * 8);
        c->any_flag.ensure(8);
        const unsigned long long* prior_in = nullptr;
        if (prior_count && np) {
            c->prior.ensure((size_t)np * 8);
            CK(cudaMemcpyAsync(c->prior.p, prior_count, (size_t)np * 8, cudaMemcpyHostToDevice, c->stream));
            prior_in = c->prior.as<unsigned long long>();
        }
        k_prior_prefix<<<blocks_for(std::max<uint32_t>(np, 1), 128), 128, 0, c->stream>>>(
            reinterpret_cast<const unsigned long long*>(d_all_counts), np, rank, world, prior_in, c->prior_eff.as<unsigned long long>(),
            c->any_flag.as<unsigned long long>(), c->dp);
        c->launches += 1;
        std::unique_ptr<lp_result> res;
        unsigned long long flag = 0;
        try {
            res.reset(run_score(c, nullptr, nullptr, c->prior_eff.as<unsigned long long>(), c->any_flag.as<unsigned long long>(), &flag));
        } catch (const Retry&) { again = true; return; }      // this rank's buffers were too small (now regrown)
        if ((flag & 16u) && (c->pend.options & LP_OPT_STRICT_BUDGET)) throw std::overflow_error("backtracking verifier exceeded its step/stack budget on another rank");
        if (flag & ~16ull) { again = true; return; }                   // another rank's counts were incomplete
        *out = res.release();
    });
    return rc != LP_OK ? rc : (again ? LP_RETRY : LP_OK);
}

// Host-side helper for callers that run the exchange themselves: frequency counts are kept per pattern id
// (FTS:46-50), so the count a pattern sees is the sum over every pattern carrying its id.
int lp_counts_by_id(const lp_ctx* c, const uint64_t* in, uint64_t* out) {
    if (!c || !in || !out) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    if (!c->loaded) { g_last_error = "lp_load_patterns has not been called"; return LP_ERR_INVALID; }
    for (uint32_t p = 0; p < c->plan.n_pat; ++p) {
        const int slot = c->plan.pat_slot[p];
        uint64_t v = 0;
        if (slot >= 0) for (uint32_t j = c->plan.slot_begin[slot]; j < c->plan.slot_begin[slot + 1]; ++j) v += in[c->plan.slot_pats[j]];
        out[p] = v;
    }
    return LP_OK;
}

// ---- one log sharded by line range (second mode)
uint32_t lp_range_halo_lines(const lp_ctx* c) {
    if (!c || !c->loaded) return 0;
    long long h = 5;                                                   // SS:275 looks 5 lines either side
    for (int32_t w : c->plan.sec_window) h = std::max<long long>(h, std::min<long long>(c->sc.max_window, w));   // SS:319
    for (uint32_t p = 0; p < c->plan.n_pat; ++p)
        if (c->plan.pat_has_ctx[p]) h = std::max<long long>(h, std::max(c->plan.pat_before[p], c->plan.pat_after[p]));   // AS:142-150
    return (uint32_t)std::min<long long>(h, 0x7FFFFFFF);
}
size_t lp_range_summary_words(const lp_ctx* c) {
    if (!c || !c->loaded) return 0;
    return 2 + (size_t)c->dp.n_pat + c->h_seq_tab_off[c->n_seq];
}
int lp_range_scan_device(lp_ctx* c, const uint8_t* d_text, size_t len, uint32_t halo_before_lines, uint32_t halo_after_lines, int options,
                         uint64_t* summary_out) {
    if (!c || !summary_out) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    return guarded(c, [&] {
        std::lock_guard<std::mutex> lk(c->mu);
        if (!c->loaded) throw std::invalid_argument("lp_load_patterns has not been called");
        const uint64_t off[2] = {0, (uint64_t)len};
        const RangeShard rg{halo_before_lines, halo_after_lines};
        const uint32_t np = c->dp.n_pat, n_tab = c->h_seq_tab_off[c->n_seq];
        std::vector<uint64_t> counts(std::max<uint32_t>(np, 1));
        std::vector<uint64_t> tab(std::max<uint32_t>(n_tab, 1));
        for (;;) {
            try {
                run_scan(c, d_text, len, off, 1, options, &rg);
                if (n_tab) {
                    c->seq_tab.ensure((size_t)n_tab * 8);
                    CK(cudaMemsetAsync(c->seq_tab.p, 0, (size_t)n_tab * 8, c->stream));
                    k_range_tables<<<blocks_for(c->n_seq, 128), 128, 0, c->stream>>>(
                        c->recs.as<unsigned long long>(), c->rx_begin.as<uint32_t>(), c->dp, c->counters.as<Counters>(), c->n_seq,
                        c->seq_tab_off.as<uint32_t>(), rg.halo_before, rg.halo_after, c->seq_tab.as<unsigned long long>());
                    c->launches += 1;
                    CK(cudaMemcpyAsync(tab.data(), c->seq_tab.p, (size_t)n_tab * 8, cudaMemcpyDeviceToHost, c->stream));
                }
                per_pattern_counts(c, counts.data());     // synchronises, validates the estimates
                break;
            } catch (const Retry&) { continue; }
        }
        Counters h{};
        CK(cudaMemcpy(&h, c->counters.p, sizeof h, cudaMemcpyDeviceToHost));
        const uint64_t lines = h.total_nl + 1;
        if ((uint64_t)halo_before_lines + halo_after_lines > lines) throw std::invalid_argument("halo line counts exceed the lines of the text");
        summary_out[0] = lines - halo_before_lines - halo_after_lines;
        summary_out[1] = np;
        for (uint32_t p = 0; p < np; ++p) summary_out[2 + p] = counts[p];
        for (uint32_t t = 0; t < n_tab; ++t) summary_out[2 + np + t] = tab[t];
    });
}
int lp_range_score(lp_ctx* c, const uint64_t* summaries, uint32_t rank, uint32_t world, const uint64_t* prior_count, lp_result** out) {
    if (!c || !out || !summaries) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    return guarded(c, [&] {
        std::lock_guard<std::mutex> lk(c->mu);
        if (!c->pend.valid || !c->pend.range) throw std::invalid_argument("no line-range scan is pending on this context");
        if (rank >= world) throw std::invalid_argument("rank >= world");
        const uint32_t np = c->dp.n_pat;
        const size_t words = 2 + (size_t)np + c->h_seq_tab_off[c->n_seq];
        long long base = 0, n_global = 0;
        std::vector<long long> core_base(world);
        for (uint32_t r = 0; r < world; ++r) {
            const uint64_t* s = summaries + (size_t)r * words;
            if (s[1] != np) throw std::invalid_argument("shard summaries come from a different pattern set");
            core_base[r] = n_global;
            if (r < rank) base += (long long)s[0];
            n_global += (long long)s[0];
        }
        if (base < (long long)c->pend.halo_before) throw std::invalid_argument("halo_before exceeds the lines of the lower shards");
        // counts seen before this shard's lines: the window's prior plus every lower shard (FTS read-before-increment order)
        std::vector<uint64_t> prior(std::max<uint32_t>(np, 1), 0);
        for (uint32_t p = 0; p < np; ++p) {
            prior[p] = prior_count ? prior_count[p] : 0;
            const int slot = c->plan.pat_slot[p];          // counts are kept per pattern id (FTS:46-50): add the whole id's events
            if (slot < 0) continue;
            for (uint32_t j = c->plan.slot_begin[slot]; j < c->plan.slot_begin[slot + 1]; ++j)
                for (uint32_t r = 0; r < rank; ++r) prior[p] += summaries[(size_t)r * words + 2 + c->plan.slot_pats[j]];
        }
        // per sequence: the line where its last-but-one event is first reachable, walking the shards in order
        std::vector<long long> seq_T(std::max<uint32_t>(c->n_seq, 1), -1);
        for (uint32_t q = 0; q < c->n_seq; ++q) {
            const uint32_t k = c->plan.seq_ev_begin[q + 1] - c->plan.seq_ev_begin[q];
            if (k < 2) { seq_T[q] = -1; continue; }             // no earlier events: condition holds for every line
            uint32_t state = 0;
            long long T = std::numeric_limits<long long>::max();  // never reached
            for (uint32_t r = 0; r < world && state + 1 < k; ++r) {
                const uint64_t e = summaries[(size_t)r * words + 2 + np + c->h_seq_tab_off[q] + state];
                const uint32_t exit_state = (uint32_t)(e >> 32), at = (uint32_t)(e & 0xFFFFFFFFull);
                if (exit_state + 1 >= k && at) { T = core_base[r] + (long long)at - 1; state = exit_state; break; }
                state = exit_state;
            }
            seq_T[q] = T;
        }
        const RangeGlobal rgl{base - (long long)c->pend.halo_before, n_global, seq_T.data()};
        try {
            *out = run_score(c, prior.data(), &rgl);
        } catch (const Retry&) {
            throw LimitFail("internal: estimates validated by lp_range_scan_device were exceeded in lp_range_score");
        }
    });
}

// One log cut by line range over the communicator of lp_comm_init, exchange on the device: this rank scans its piece
// (core lines + halos), its summary [per-pattern counts of the core lines, status, core lines, chain tables] is
// all-gathered over NVLink on the second stream, k_range_compose turns the gathered summaries into this shard's prior
// counts, place in the log and sequence thresholds, and k_score waits for that. Nothing comes back to the host before
// the events do. Collective: every rank calls it once per step; a rank whose scratch estimates were too small makes
// all ranks repeat the step (they see it in the gathered status words). Same results as lp_range_scan_device +
// lp_range_score, which remain for callers that run the exchange themselves.
static int analyze_range_dist(lp_ctx* c, const uint8_t* text, bool on_host, size_t len, uint32_t halo_before, uint32_t halo_after,
                              const uint64_t* prior_count, int options, lp_result** out) {
    if (!c || !out) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    *out = nullptr;
    return guarded(c, [&] {
        std::lock_guard<std::mutex> lk(c->mu);
        if (!c->loaded) throw std::invalid_argument("lp_load_patterns has not been called");
        if (!c->nccl_comm) throw std::invalid_argument("lp_comm_init has not been called on this context");
        if (on_host && len && !text) throw std::invalid_argument("text is NULL");
        const uint8_t* d_text = on_host ? stage_host_log(c, text, len) : text;
        struct DropHostSrc { lp_ctx* c; ~DropHostSrc() { c->host_src = nullptr; } } drop{c};
        const uint32_t np = c->dp.n_pat, world = c->comm_world, rank = c->comm_rank, n_tab = c->h_seq_tab_off[c->n_seq];
        const size_t words = (size_t)np + 3 + n_tab;
        c->dist_mine.ensure(words * 8);
        c->dist_all.ensure(words * 8 * world);
        c->prior_eff.ensure((size_t)std::max<uint32_t>(np, 1) * 8);
        c->any_flag.ensure(8);
        c->seq_T.ensure((size_t)std::max<uint32_t>(c->n_seq, 1) * 8);
        c->range_dev.ensure(32);
        for (int k = 0; k < 2; ++k) if (!c->dist_ev[k]) CK(cudaEventCreateWithFlags(&c->dist_ev[k], cudaEventDisableTiming));
        const unsigned long long* prior_in = nullptr;
        if (prior_count && np) {
            c->prior.ensure((size_t)np * 8);
            CK(cudaMemcpyAsync(c->prior.p, prior_count, (size_t)np * 8, cudaMemcpyHostToDevice, c->stream));
            prior_in = c->prior.as<unsigned long long>();
        }
        const uint64_t off[2] = {0, (uint64_t)len};
        const RangeShard rg{halo_before, halo_after};
        for (int round = 0;; ++round) {
            if (round > 16) throw LimitFail("scratch estimates did not settle");
            run_scan(c, d_text, len, off, 1, options, &rg);
            unsigned long long* mine = c->dist_mine.as<unsigned long long>();
            Counters* ctr = c->counters.as<Counters>();
            CK(cudaMemsetAsync(mine, 0, words * 8, c->stream));
            k_count_per_pattern<<<c->n_sm * 4, 256, 0, c->stream>>>(c->recs.as<unsigned long long>(), ctr, c->dp, mine, rg.halo_before, rg.halo_after);
            k_publish_overflow<<<1, 128, 0, c->stream>>>(ctr, est_caps(c), mine, np, mine + np);
            k_range_header<<<1, 32, 0, c->stream>>>(ctr, np, rg.halo_before, rg.halo_after, mine + np + 1);
            if (n_tab)
                k_range_tables<<<blocks_for(c->n_seq, 128), 128, 0, c->stream>>>(c->recs.as<unsigned long long>(), c->rx_begin.as<uint32_t>(), c->dp, ctr, c->n_seq,
                                                                                  c->seq_tab_off.as<uint32_t>(), rg.halo_before, rg.halo_after, mine + np + 3);
            CK(cudaEventRecord(c->dist_ev[0], c->stream));
            CK(cudaStreamWaitEvent(c->copy_stream, c->dist_ev[0], 0));
            nccl_check(nccl().AllGather(mine, c->dist_all.p, words, NCCL_UINT64, c->nccl_comm, c->copy_stream), "ncclAllGather");
            k_range_compose<<<blocks_for(std::max<uint32_t>(std::max(np, c->n_seq), 1), 128), 128, 0, c->copy_stream>>>(
                c->dist_all.as<unsigned long long>(), np, n_tab, rank, world, prior_in, rg.halo_before, c->n_seq, c->seq_tab_off.as<uint32_t>(), c->dp,
                c->prior_eff.as<unsigned long long>(), c->any_flag.as<unsigned long long>(), c->range_dev.as<long long>(), c->seq_T.as<long long>());
            CK(cudaEventRecord(c->dist_ev[1], c->copy_stream));
            c->launches += 5;
            std::unique_ptr<lp_result> res;
            bool again = false;
            unsigned long long flag = 0;
            RangeGlobal rgl{0, 0, nullptr};
            rgl.dev_range = c->range_dev.as<long long>(); rgl.dev_seq_T = c->seq_T.as<long long>();
            try {
                res.reset(run_score(c, nullptr, &rgl, c->prior_eff.as<unsigned long long>(), c->any_flag.as<unsigned long long>(), &flag, c->dist_ev[1]));
            } catch (const Retry&) { again = true; }
            if (again) {
                CK(cudaMemcpy(&flag, c->any_flag.p, 8, cudaMemcpyDeviceToHost));
                if (!flag) throw LimitFail("internal: a repeat was requested that the other ranks cannot see");
                continue;
            }
            if ((flag & 16u) && (options & LP_OPT_STRICT_BUDGET)) throw std::overflow_error("backtracking verifier exceeded its step/stack budget");
            if (flag & ~16ull) continue;                 // every rank sees the same words: all of them repeat
            *out = res.release();
            if (on_host) {
                float ms = 0;
                cudaEventElapsedTime(&ms, c->ev[5], c->ev[8]);
                c->tm.h2d_ms = ms;
            }
            return;
        }
    });
}
int lp_analyze_range_dist_device(lp_ctx* c, const uint8_t* d_text, size_t len, uint32_t halo_before_lines, uint32_t halo_after_lines,
                                 const uint64_t* prior_count, int options, lp_result** out) {
    return analyze_range_dist(c, d_text, false, len, halo_before_lines, halo_after_lines, prior_count, options, out);
}
int lp_analyze_range_dist(lp_ctx* c, const uint8_t* text, size_t len, uint32_t halo_before_lines, uint32_t halo_after_lines,
                          const uint64_t* prior_count, int options, lp_result** out) {
    return analyze_range_dist(c, text, true, len, halo_before_lines, halo_after_lines, prior_count, options, out);
}

int lp_analyze_batch_device(lp_ctx* c, const uint8_t* d_logs, const uint64_t* offsets, size_t n_logs, const uint64_t* prior_count,
                            int options, lp_result** out) {
    if (!c || !out) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    return guarded(c, [&] {
        if (!offsets) throw std::invalid_argument("offsets is NULL");
        WorkerLock wl = pick_worker(c);
        *out = run_all(wl.w, [&] { return d_logs; }, offsets[n_logs], offsets, (uint32_t)n_logs, prior_count, options);
    });
}
// ---- one process, several GPUs (the reference is ONE JVM; lp_create(cfg, -1) / lp_create_multi)
// A batch is cut into contiguous blocks of logs, one per device, balanced by bytes. Every block is copied to its device
// and scanned on a host thread of its own; the per-pattern event counts of the blocks meet on the host (P x 8 bytes per
// device: the only exchange, FTS:41-93 is the only cross-request state), every block is scored with the counts of the
// blocks in front of it as its prior, and the results are put together in batch order. Same events, scores, per-log
// ranks and counts as lp_analyze_batch on one device.
namespace {
lp_result* merge_results(std::vector<std::unique_ptr<lp_result>>& parts, uint32_t np) {
    std::unique_ptr<lp_result> R(new lp_result());
    uint64_t M = 0;
    bool factors = true, spans = true;
    for (auto& p : parts) { M += p->n_events; factors = factors && p->has_factors; spans = spans && p->span_start && p->span_end; }
    const size_t w_u32 = (size_t)(M + 1) / 2 + 1;       // 8-byte words holding M uint32
    size_t words = 3 * w_u32 + M + (factors ? 7 * M : 0) + (spans ? 2 * w_u32 : 0) + np + 1;
    R->own.assign(words, 0ull);
    unsigned long long* at = R->own.data();
    uint32_t* line = reinterpret_cast<uint32_t*>(at); at += w_u32;
    uint32_t* pat = reinterpret_cast<uint32_t*>(at); at += w_u32;
    uint32_t* rank = reinterpret_cast<uint32_t*>(at); at += w_u32;
    double* score = reinterpret_cast<double*>(at); at += M;
    double* fac = nullptr;
    if (factors) { fac = reinterpret_cast<double*>(at); at += 7 * M; }
    uint32_t *ss = nullptr, *se = nullptr;
    if (spans) { ss = reinterpret_cast<uint32_t*>(at); at += w_u32; se = reinterpret_cast<uint32_t*>(at); at += w_u32; }
    uint64_t* delta = reinterpret_cast<uint64_t*>(at);
    uint64_t e0 = 0;
    R->log_event_begin.push_back(0);
    for (auto& p : parts) {
        const uint64_t m = p->n_events;
        if (m) {
            memcpy(line + e0, p->line, m * 4); memcpy(pat + e0, p->pattern, m * 4); memcpy(rank + e0, p->rank, m * 4);
            memcpy(score + e0, p->score, m * 8);
            if (fac) memcpy(fac + 7 * e0, p->factors, m * 56);
            if (ss) { memcpy(ss + e0, p->span_start, m * 4); memcpy(se + e0, p->span_end, m * 4); }
        }
        for (uint32_t q = 0; q < np; ++q) delta[q] += p->freq_delta[q];
        for (uint64_t k = 0; k < p->n_logs; ++k) {
            R->log_n_lines.push_back(p->log_n_lines[k]);
            R->log_event_begin.push_back(e0 + p->log_event_begin[k + 1]);
        }
        e0 += m;
        R->n_logs += p->n_logs; R->total_lines += p->total_lines; R->budget_exceeded += p->budget_exceeded;
    }
    R->n_events = M; R->n_patterns = np;
    R->line = line; R->pattern = pat; R->rank = rank; R->score = score; R->freq_delta = delta;
    R->has_factors = factors; R->factors = fac; R->span_start = ss; R->span_end = se;
    return R.release();
}

void analyze_batch_multi(lp_ctx* c, const uint8_t* logs, const uint64_t* offsets, size_t n_logs, const uint64_t* prior_count, int options,
                         lp_result** out) {
    if (options & LP_OPT_LINE_OFFSETS) throw std::invalid_argument("LP_OPT_LINE_OFFSETS is not available for a batch cut over several devices");
    if (!c->loaded) throw std::invalid_argument("lp_load_patterns has not been called");
    std::vector<lp_ctx*> devs{c};
    devs.insert(devs.end(), c->shards.begin(), c->shards.end());
    const size_t S = std::min(devs.size(), n_logs);
    const uint64_t total = offsets[n_logs];
    if (total && !logs) throw std::invalid_argument("logs is NULL");
    std::vector<size_t> cut(S + 1, 0);        // block s = logs [cut[s], cut[s+1]): contiguous, at least one log each, balanced by bytes
    cut[S] = n_logs;
    for (size_t s = 1; s < S; ++s) {
        size_t k = std::max(cut[s - 1] + 1, (size_t)(std::lower_bound(offsets, offsets + n_logs + 1, total / S * s) - offsets));
        cut[s] = std::min(k, n_logs - (S - s));
    }
    const uint32_t np = c->dp.n_pat;
    std::vector<std::unique_lock<std::mutex>> locks;      // held across both phases: a scan is pending on every device in between
    for (size_t s = 0; s < S; ++s) locks.emplace_back(devs[s]->mu);
    std::vector<std::vector<uint64_t>> counts(S, std::vector<uint64_t>(std::max<uint32_t>(np, 1), 0)), off(S);
    std::vector<std::exception_ptr> err(S);
    auto run_phase = [&](const std::function<void(size_t)>& body) {
        std::vector<std::thread> th;
        for (size_t s = 0; s < S; ++s)
            th.emplace_back([&, s] {
                try { CK(cudaSetDevice(devs[s]->device)); body(s); } catch (...) { err[s] = std::current_exception(); }
            });
        for (auto& t : th) t.join();
        for (size_t s = 0; s < S; ++s) if (err[s]) std::rethrow_exception(err[s]);
    };
    run_phase([&](size_t s) {
        lp_ctx* d = devs[s];
        const uint64_t b0 = offsets[cut[s]], len = offsets[cut[s + 1]] - b0;
        const uint32_t n = (uint32_t)(cut[s + 1] - cut[s]);
        off[s].resize(n + 1);
        for (uint32_t k = 0; k <= n; ++k) off[s][k] = offsets[cut[s] + k] - b0;
        const uint8_t* dptr = stage_host_log(d, logs + b0, len);
        struct DropHostSrc { lp_ctx* c; ~DropHostSrc() { c->host_src = nullptr; } } drop{d};
        for (;;) {
            try {
                run_scan(d, dptr, len, off[s].data(), n, options);
                per_pattern_counts(d, counts[s].data());
                break;
            } catch (const Retry&) { continue; }
        }
    });
    std::vector<std::vector<uint64_t>> prior(S, std::vector<uint64_t>(std::max<uint32_t>(np, 1), 0));
    std::vector<uint64_t> run(std::max<uint32_t>(np, 1), 0), by_id(std::max<uint32_t>(np, 1), 0);
    for (uint32_t p = 0; p < np; ++p) run[p] = prior_count ? prior_count[p] : 0;
    for (size_t s = 0; s < S; ++s) {
        prior[s] = run;
        lp_counts_by_id(c, counts[s].data(), by_id.data());       // counts are kept per pattern id (FTS:46-50)
        for (uint32_t p = 0; p < np; ++p) run[p] += by_id[p];
    }
    std::vector<std::unique_ptr<lp_result>> parts(S);
    run_phase([&](size_t s) {
        try {
            parts[s].reset(run_score(devs[s], prior[s].data()));
        } catch (const Retry&) {
            throw LimitFail("internal: estimates validated by the scan phase were exceeded in the score phase");
        }
    });
    *out = merge_results(parts, np);
    CK(cudaSetDevice(c->device));
}
}  // namespace

int lp_analyze_batch(lp_ctx* c, const uint8_t* logs, const uint64_t* offsets, size_t n_logs, const uint64_t* prior_count, int options,
                     lp_result** out) {
    if (!c || !out) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    if (!c->shards.empty() && offsets && n_logs >= 2)
        return guarded(c, [&] { analyze_batch_multi(c, logs, offsets, n_logs, prior_count, options, out); });
    return guarded(c, [&] {
        if (!offsets) throw std::invalid_argument("offsets is NULL");
        WorkerLock wl = pick_worker(c);
        lp_ctx* const w = wl.w;
        const uint64_t len = offsets[n_logs];
        if (len && !logs) throw std::invalid_argument("logs is NULL");
        bool staged = false;
        const uint8_t* d = nullptr;
        *out = run_all(w, [&] { if (!staged) { d = stage_host_log(w, logs, len); staged = true; } return d; }, len, offsets, (uint32_t)n_logs,
                       prior_count, options);
        float ms = 0;
        cudaEventElapsedTime(&ms, w->ev[5], w->ev[8]);
        w->tm.h2d_ms = ms;
    });
}
int lp_analyze_device(lp_ctx* c, const uint8_t* d_log, size_t len, const uint64_t* prior_count, int options, lp_result** out) {
    const uint64_t off[2] = {0, (uint64_t)len};
    return lp_analyze_batch_device(c, d_log, off, 1, prior_count, options, out);
}
int lp_analyze(lp_ctx* c, const uint8_t* log, size_t len, const uint64_t* prior_count, int options, lp_result** out) {
    const uint64_t off[2] = {0, (uint64_t)len};
    return lp_analyze_batch(c, log, off, 1, prior_count, options, out);
}

// Page-locked host memory for the caller's log buffers. cudaMemcpyAsync from pageable memory is staged through a
// driver buffer and blocks the calling thread; from page-locked memory the piecewise copy of lp_analyze* really
// overlaps the scan. The JVM either lets the library allocate (and wraps the pointer in a MemorySegment) or
// registers a segment it already owns.
void* lp_host_alloc(size_t bytes) {
    void* p = nullptr;
    const int rc = guarded(nullptr, [&] { CK(cudaHostAlloc(&p, std::max<size_t>(bytes, 1), cudaHostAllocPortable)); });
    return rc == LP_OK ? p : nullptr;
}
void lp_host_free(void* p) { if (p) cudaFreeHost(p); }
int lp_host_register(void* p, size_t bytes) {
    if (!p || !bytes) { g_last_error = "NULL or empty range"; return LP_ERR_INVALID; }
    return guarded(nullptr, [&] { CK(cudaHostRegister(p, bytes, cudaHostRegisterPortable)); });
}
int lp_host_unregister(void* p) {
    if (!p) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    return guarded(nullptr, [&] { CK(cudaHostUnregister(p)); });
}

int lp_result_get(const lp_result* r, lp_result_view* v) {
    if (!r || !v) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    memset(v, 0, sizeof *v);
    v->n_logs = r->n_logs;
    v->n_events = r->n_events;
    v->log_n_lines = r->log_n_lines.data();
    v->log_event_begin = r->log_event_begin.data();
    v->line = r->line;
    v->pattern = r->pattern;
    v->score = r->score;
    v->rank = r->rank;
    v->factors = r->has_factors ? r->factors : nullptr;
    v->freq_delta = r->freq_delta;
    v->n_patterns = r->n_patterns;
    v->line_start = r->has_offsets ? r->line_start.data() : nullptr;
    v->line_end = r->has_offsets ? r->line_end.data() : nullptr;
    v->total_lines = r->total_lines;
    v->span_start = r->span_start;
    v->span_end = r->span_end;
    v->budget_exceeded = r->budget_exceeded;
    return LP_OK;
}
void lp_result_free(lp_result* r) { delete r; }

int lp_get_timings(const lp_ctx* c, lp_timings* out) {
    if (!c || !out) { g_last_error = "NULL argument"; return LP_ERR_INVALID; }
    const lp_ctx* w = c->last_used.load();       // the worker that ran the most recent stateless request, if any
    *out = w ? w->tm : c->tm;
    return LP_OK;
}

void* lp_stream(const lp_ctx* c) { return c ? (void*)c->stream : nullptr; }

}  // extern "C"
