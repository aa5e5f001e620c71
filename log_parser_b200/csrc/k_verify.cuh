// Matcher.find() kernels: DFA walkers, k_confirm, k_resolve, k_verify*, k_spans, k_dfa_scan.
// Part of engine.cu's translation unit (included inside its anonymous namespace); not a stand-alone header.
// ------------------------------------------------------------------ DFA walker (same loop as lpx::Dfa::match)
__device__ __forceinline__ long tail_terminator(const uint8_t* __restrict__ s, uint32_t ls, uint32_t le) {
    const uint32_t n = le - ls;
    if (n >= 1 && s[le - 1] == '\r') return (long)le - 1;
    if (n >= 2 && s[le - 2] == 0xC2 && s[le - 1] == 0x85) return (long)le - 2;
    if (n >= 3 && s[le - 3] == 0xE2 && s[le - 2] == 0x80 && (s[le - 1] == 0xA8 || s[le - 1] == 0xA9)) return (long)le - 3;
    return -1;
}
// NG automata over one line. The line is consumed 8 bytes per (aligned) load with the eight
// transitions unrolled; unaligned head/tail bytes are stepped one by one. `LOWFINAL` selects the
// single-DFA convention (states 0/1 = DEAD/MATCH are final) or the product-DFA one (states >= ff).
template <int NG, bool LOWFINAL>
struct LineWalker {
    const uint8_t* __restrict__ cm[NG];
    const uint16_t* __restrict__ T[NG];
    uint32_t nc[NG], ff[NG], st[NG];

    __device__ __forceinline__ bool live(int q) const { return LOWFINAL ? st[q] > 1u : st[q] < ff[q]; }
    __device__ __forceinline__ bool step(uint32_t b) {
        bool any = false;
#pragma unroll
        for (int q = 0; q < NG; ++q)
            if (live(q)) { st[q] = T[q][st[q] * nc[q] + cm[q][b]]; any = true; }
        return any;
    }
    __device__ __forceinline__ void symbol(uint32_t col_from_end) {     // 2 = TPOS, 1 = EOT
#pragma unroll
        for (int q = 0; q < NG; ++q) if (live(q)) st[q] = T[q][st[q] * nc[q] + nc[q] - col_from_end];
    }
    // bytes [a, b) of text; returns false once every automaton is final. Aligned 8-byte words,
    // one word prefetched ahead; partial first/last words are stepped byte by byte from registers.
    __device__ __forceinline__ bool segment(const uint8_t* __restrict__ text, uint64_t len, uint32_t a, uint32_t b) {
        if (a >= b) return true;
        uint32_t base = a & ~7u;
        unsigned long long cur = load8_any(text, base, len);
        unsigned long long nxt = (base + 8 < b) ? load8_any(text, (uint64_t)base + 8, len) : 0ull;
        uint32_t lo = a - base;
        for (;;) {
            const uint32_t hi = min(8u, b - base);
            if constexpr (NG == 1) {
                // one automaton: the eight class lookups do not depend on the state, so they are issued
                // together and only the transition loads form a dependent chain
                uint32_t cls[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) cls[k] = cm[0][(uint32_t)(cur >> (8 * k)) & 0xFFu];
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if ((uint32_t)k >= lo && (uint32_t)k < hi && live(0)) st[0] = T[0][st[0] * nc[0] + cls[k]];
                if (!live(0)) return false;
            } else if (lo == 0 && hi == 8) {
                const uint32_t x = (uint32_t)cur, y = (uint32_t)(cur >> 32);
                bool any = true;
#pragma unroll
                for (int k = 0; k < 4; ++k) any = step((x >> (8 * k)) & 0xFFu);
#pragma unroll
                for (int k = 0; k < 4; ++k) any = step((y >> (8 * k)) & 0xFFu);
                if (!any) return false;
            } else {
                for (uint32_t k = lo; k < hi; ++k)
                    if (!step((uint32_t)(cur >> (8 * k)) & 0xFFu)) return false;
            }
            base += 8;
            if (base >= b) return true;
            lo = 0;
            cur = nxt;
            nxt = (base + 8 < b) ? load8_any(text, (uint64_t)base + 8, len) : 0ull;
        }
    }
    // bytes [a, b) of text from b-1 down to a (automaton of the reversed language); NG == 1 only
    __device__ __forceinline__ bool segment_rev(const uint8_t* __restrict__ text, uint64_t len, uint32_t a, uint32_t b) {
        if (a >= b) return true;
        uint32_t base = (b - 1) & ~7u;
        unsigned long long cur = load8_any(text, base, len);
        unsigned long long prv = (base >= 8 && base > a) ? load8_any(text, (uint64_t)base - 8, len) : 0ull;
        for (;;) {
            const uint32_t hi = min(8u, b - base), lo = a > base ? a - base : 0u;
            uint32_t cls[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) cls[k] = cm[0][(uint32_t)(cur >> (8 * k)) & 0xFFu];
#pragma unroll
            for (int k = 7; k >= 0; --k)
                if ((uint32_t)k >= lo && (uint32_t)k < hi && live(0)) st[0] = T[0][st[0] * nc[0] + cls[k]];
            if (!live(0)) return false;
            if (base <= a) return true;
            base -= 8;
            b = base + 8;
            cur = prv;
            prv = (base >= 8 && base > a) ? load8_any(text, (uint64_t)base - 8, len) : 0ull;
        }
    }
    __device__ __forceinline__ void run(const uint8_t* __restrict__ text, uint64_t len, uint32_t ls, uint32_t le) {
        if (le > ls) {
            const long tpos = tail_terminator(text, ls, le);
            if (tpos < 0) segment(text, len, ls, le);
            else if (segment(text, len, ls, (uint32_t)tpos)) { symbol(2); segment(text, len, (uint32_t)tpos, le); }
        }
        symbol(1);
    }
};

// One regex over one line (tables in global memory, L1/L2-cached).
__device__ bool dfa_walk(const DevPlan& P, uint32_t r, const uint8_t* __restrict__ text, uint64_t len, uint32_t ls, uint32_t le) {
    LineWalker<1, true> w;
    w.nc[0] = P.rx_ncls[r];
    w.T[0] = P.dfa_trans + P.rx_dfa_off[r];
    w.cm[0] = P.rx_cmap + (size_t)r * 256;
    w.ff[0] = 0;
    w.st[0] = P.rx_start[r];
    w.run(text, len, ls, le);
    return w.st[0] == 1;
}

__device__ __forceinline__ void emit_record(Counters* ctr, unsigned long long* recs, uint32_t cap, uint32_t gb, bool hit, uint32_t r, uint32_t g) {
    const unsigned mask = 0xffffffffu;   // callers keep all 32 lanes in the loop
    const unsigned votes = __ballot_sync(mask, hit);
    if (!votes) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(votes) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(&ctr->n_recs, (unsigned)__popc(votes));
    base = __shfl_sync(mask, base, leader);
    if (hit) {
        const uint32_t idx = base + __popc(votes & ((1u << lane) - 1u));
        if (idx < cap) recs[idx] = ((unsigned long long)r << gb) | g;
        else atomicOr(&ctr->overflow, 4u);
    }
}

// Block-level staging of queue output: entries are collected in shared memory (shared-memory
// atomics only) and published with one global atomic per flush, so no warp waits on a global atomic.
constexpr uint32_t Q_STAGE = 2048;
struct QStage {
    unsigned long long buf[Q_STAGE];
    uint32_t n, base;
};
__device__ __forceinline__ void stage_put(QStage& sg, Counters* ctr, unsigned int* gcount, unsigned long long* gbuf, uint32_t cap,
                                          uint32_t overflow_bit, unsigned long long key) {
    const uint32_t idx = atomicAdd(&sg.n, 1u);
    if (idx < Q_STAGE) sg.buf[idx] = key;
    else {      // staging full: straight to the global queue
        const uint32_t gi = atomicAdd(gcount, 1u);
        if (gi < cap) gbuf[gi] = key; else atomicOr(&ctr->overflow, overflow_bit);
    }
}
// all threads of the block; `force` flushes whatever is staged, otherwise only a nearly full buffer
__device__ __forceinline__ void stage_flush(QStage& sg, Counters* ctr, unsigned int* gcount, unsigned long long* gbuf, uint32_t cap,
                                            uint32_t overflow_bit, bool force) {
    __syncthreads();
    const uint32_t staged = min(sg.n, Q_STAGE);
    if (staged == 0 || (!force && staged < Q_STAGE - 2 * blockDim.x)) return;     // block-uniform
    if (threadIdx.x == 0) sg.base = atomicAdd(gcount, staged);
    __syncthreads();
    const uint32_t base = sg.base;
    for (uint32_t q = threadIdx.x; q < staged; q += blockDim.x) {
        if (base + q < cap) gbuf[base + q] = sg.buf[q]; else atomicOr(&ctr->overflow, overflow_bit);
    }
    __syncthreads();
    if (threadIdx.x == 0) sg.n = 0;
    __syncthreads();
}

// ------------------------------------------------------------------ K3: confirm gram hits, resolve literal occurrences
// Two kernels with a queue between them, so that the lanes of a warp do the same kind of work
// (one thread per candidate doing everything serially ran with ~6 of 32 lanes active):
//   k_confirm  thread per candidate: gram table probe, literal comparison in 4-byte words; every
//              (occurrence, regex) pair becomes an item in the queue of its verification class
//   k_resolve  thread per item, a block works on one class at a time: line lookup, then
//              class 0 boundary test | 1 anchored automaton forwards | 2 backwards | 3 hand over to
//              k_verify / k_verify_vm (walk of the whole line)
constexpr int N_CLASSES = 4;
constexpr uint32_t ITEM_STAGE = 1024;      // items staged per class and block
struct ItemStage { unsigned long long buf[N_CLASSES][ITEM_STAGE]; uint32_t n[N_CLASSES], base[N_CLASSES]; };

struct ConfirmArgs {
    const uint8_t* text;
    uint64_t len;
    const uint32_t* cand;
    uint32_t cand_cap;
    unsigned long long* items;       // [N_CLASSES][item_cap]: s << 32 | literal length << 24 | entry of lit_rx_packed
    uint32_t item_cap;
    Counters* ctr;
};
__device__ __forceinline__ bool dev_word(uint32_t b) { return (b >= '0' && b <= '9') || (b >= 'A' && b <= 'Z') || b == '_' || (b >= 'a' && b <= 'z'); }
// 4 bytes at any position p; bytes at or beyond len read as zero
__device__ __forceinline__ uint32_t load4_unaligned(const uint8_t* __restrict__ text, uint32_t p, uint64_t len) {
    const uint32_t b = p & ~3u;
    if ((uint64_t)b + 8 <= len) {
        const uint32_t lo = __ldg(reinterpret_cast<const uint32_t*>(text + b)), hi = __ldg(reinterpret_cast<const uint32_t*>(text + b + 4));
        return __funnelshift_r(lo, hi, (p & 3u) * 8);
    }
    uint32_t w = 0;
    for (int i = 0; i < 4; ++i) if ((uint64_t)p + i < len) w |= (uint32_t)text[p + i] << (8 * i);
    return w;
}
// ASCII 'A'..'Z' -> lower case in all four bytes, everything else unchanged
__device__ __forceinline__ uint32_t lower4(uint32_t t) {
    const uint32_t y = t & 0x7F7F7F7Fu;
    const uint32_t up = (y + 0x3F3F3F3Fu) & ~(y + 0x25252525u) & ~t & 0x80808080u;
    return t | (up >> 2);
}
__global__ void __launch_bounds__(256) k_confirm(ConfirmArgs a, DevPlan P) {
    __shared__ ItemStage sg;
    if (threadIdx.x < N_CLASSES) sg.n[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t n = min(a.ctr->n_cand, a.cand_cap);
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t tmask = (1u << P.table_bits) - 1u;
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < n; c += stride) {
        const uint32_t pos = a.cand[c];
        const uint32_t w = load4_unaligned(a.text, pos, a.len);
        const uint32_t f = w | 0x20202020u;
        uint32_t slot = (f * GRAM_MUL2) >> (32 - P.table_bits);
        uint4 gs;
        while ((gs = __ldg(&P.gram_slots[slot])).x != 0 && gs.x != f) slot = (slot + 1) & tmask;
        for (uint32_t q = 0; q < gs.z; ++q) {       // gs.z == 0 on an empty slot
            const uint4 ld = __ldg(&P.lit_desc[gs.y + q]);
            const uint32_t ll = ld.y & 0xFFu, go = (ld.y >> 8) & 0xFFu;
            if (pos < go) continue;
            const uint32_t s = pos - go;
            if ((uint64_t)s + ll > a.len) continue;
            const uint32_t* __restrict__ lit = reinterpret_cast<const uint32_t*>(P.lit_bytes8 + ld.x);
            const bool icase = (ld.y >> 16) & 1u;
            bool ok = true;
            for (uint32_t j = 0; j < ll && ok; j += 4) {
                uint32_t t = (j == go) ? w : load4_unaligned(a.text, s + j, a.len);
                if (icase) t = lower4(t);
                const uint32_t rem = ll - j;
                const uint32_t mask = rem >= 4 ? 0xFFFFFFFFu : (1u << (8 * rem)) - 1u;
                ok = ((t ^ __ldg(lit + (j >> 2))) & mask) == 0;
            }
            if (!ok) continue;
            for (uint32_t j = ld.z; j < ld.z + ld.w; ++j) {
                const uint32_t mode = __ldg(&P.lit_rx_packed[j]) >> 24;
                const uint32_t cls = (mode & 1u) ? 0u : (mode & 8u) ? 1u : (mode & 16u) ? 2u : 3u;
                const unsigned long long item = ((unsigned long long)s << 32) | ((unsigned long long)ll << 24) | j;
                const uint32_t idx = atomicAdd(&sg.n[cls], 1u);
                if (idx < ITEM_STAGE) sg.buf[cls][idx] = item;
                else {      // staging full: straight to the global queue
                    const uint32_t gi = atomicAdd(&a.ctr->n_items[cls], 1u);
                    if (gi < a.item_cap) a.items[(size_t)cls * a.item_cap + gi] = item; else atomicOr(&a.ctr->overflow, 128u);
                }
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < N_CLASSES) {
        const uint32_t staged = min(sg.n[threadIdx.x], ITEM_STAGE);
        sg.base[threadIdx.x] = staged ? atomicAdd(&a.ctr->n_items[threadIdx.x], staged) : 0u;
    }
    __syncthreads();
    for (int cls = 0; cls < N_CLASSES; ++cls) {
        const uint32_t staged = min(sg.n[cls], ITEM_STAGE), base = sg.base[cls];
        for (uint32_t q = threadIdx.x; q < staged; q += blockDim.x) {
            if (base + q < a.item_cap) a.items[(size_t)cls * a.item_cap + base + q] = sg.buf[cls][q];
            else atomicOr(&a.ctr->overflow, 128u);
        }
    }
}

struct ResolveArgs {
    const uint8_t* text;
    uint64_t len;
    const unsigned long long* items;
    uint32_t item_cap;
    const uint32_t* tile_prefix;     // exclusive newline count before each tile of the scan
    uint32_t n_tiles;
    uint32_t tile_bytes;             // tile size of the scan that wrote tile_prefix
    const uint16_t* span_counts;     // [n_tiles][16] newlines per 1 KiB span of a tile (k_fscan), or NULL
    const uint64_t* log_off;
    uint32_t n_logs;
    const uint32_t* line_start;
    const uint32_t* line_end;
    uint32_t n_lines_total;
    uint4* tasks;                    // {regex | flags, line, position, state}, see k_verify
    uint32_t task_cap;
    unsigned long long* vm_tasks;
    uint32_t vm_task_cap;
    unsigned long long* recs;
    uint32_t rec_cap;
    Counters* ctr;
    uint32_t budget;                 // anchored-walk steps per item (ANCHOR_BUDGET)
};
constexpr uint32_t ANCHOR_BUDGET = 12;     // default of ResolveArgs::budget: automaton steps an item may spend in k_resolve before it is handed to k_verify
constexpr uint32_t T_STAGE = 1024;         // tasks staged per block of k_resolve
constexpr uint32_t VT_FWD = 1u << 30, VT_REV = 1u << 31;      // task.x flags: continue an anchored walk forwards / backwards (else: whole-line walk)
__global__ void __launch_bounds__(256) k_resolve(ResolveArgs a, DevPlan P) {
    __shared__ QStage s_rec;
    __shared__ uint4 s_tbuf[T_STAGE];        // tasks for k_verify, staged like the records (16-byte entries)
    __shared__ uint32_t s_tn, s_tbase;
    if (threadIdx.x == 0) { s_rec.n = 0; s_tn = 0; }
    __syncthreads();
    auto put_task = [&](const uint4& t) {
        const uint32_t idx = atomicAdd(&s_tn, 1u);
        if (idx < T_STAGE) s_tbuf[idx] = t;
        else {      // staging full: straight to the global queue
            const uint32_t gi = atomicAdd(&a.ctr->n_tasks, 1u);
            if (gi < a.task_cap) a.tasks[gi] = t; else atomicOr(&a.ctr->overflow, 2u);
        }
    };
    const uint32_t n0 = min(a.ctr->n_items[0], a.item_cap), n1 = min(a.ctr->n_items[1], a.item_cap),
                   n2 = min(a.ctr->n_items[2], a.item_cap), n3 = min(a.ctr->n_items[3], a.item_cap);
    const uint32_t c1 = (n0 + blockDim.x - 1) / blockDim.x, c2 = c1 + (n1 + blockDim.x - 1) / blockDim.x,
                   c3 = c2 + (n2 + blockDim.x - 1) / blockDim.x, c4 = c3 + (n3 + blockDim.x - 1) / blockDim.x;
    const uint32_t gmax = min(a.ctr->n_lines_total, a.n_lines_total);
    for (uint32_t ch = blockIdx.x; ch < c4; ch += gridDim.x) {
        // a chunk of 256 items never straddles two classes: the block does one kind of work at a time
        const uint32_t cls = ch >= c3 ? 3u : ch >= c2 ? 2u : ch >= c1 ? 1u : 0u;
        const uint32_t first = cls == 3 ? c3 : cls == 2 ? c2 : cls == 1 ? c1 : 0u;
        const uint32_t count = cls == 3 ? n3 : cls == 2 ? n2 : cls == 1 ? n1 : n0;
        const uint32_t idx = (ch - first) * blockDim.x + threadIdx.x;
        if (idx < count) do {
            const unsigned long long item = a.items[(size_t)cls * a.item_cap + idx];
            const uint32_t s = (uint32_t)(item >> 32), ll = (uint32_t)(item >> 24) & 0xFFu, e2 = s + ll;
            const uint32_t packed = __ldg(&P.lit_rx_packed[(uint32_t)item & 0xFFFFFFu]);
            const uint32_t r = packed & 0xFFFFFFu, mode = packed >> 24;
            // line of the occurrence: the newline prefix of its tile (and its neighbour) brackets the index
            const uint32_t t = s / a.tile_bytes;
            uint32_t glo = a.tile_prefix[t], ghi = (t + 1 < a.n_tiles ? a.tile_prefix[t + 1] : a.ctr->n_lines_total) + 1;
            if (a.span_counts) {        // the scan also counted the newlines of every 1 KiB span: bracket the line within its span
                const uint4* row = reinterpret_cast<const uint4*>(a.span_counts + (size_t)t * 16);
                const uint4 c0 = __ldg(row), c1 = __ldg(row + 1);
                const uint32_t cw[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
                const uint32_t w = (s - t * a.tile_bytes) >> 10;
                uint32_t before = 0, mine = 0;
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const uint32_t v = (cw[q >> 1] >> ((q & 1) * 16)) & 0xFFFFu;
                    before += (uint32_t)q < w ? v : 0u;
                    mine = (uint32_t)q == w ? v : mine;
                }
                glo += before;
                ghi = glo + mine + 1;
            }
            if (a.n_logs > 1) {
                glo += upper_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, (uint64_t)t * a.tile_bytes) - 1;
                ghi += upper_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, (uint64_t)(t + 1) * a.tile_bytes);
            }
            ghi = min(ghi, gmax);
            if (glo >= ghi) break;                    // beyond the (under-estimated) line arrays: the call is re-run
            const uint32_t g = glo + upper_bound_dev<uint32_t>(a.line_start + glo, ghi - glo, s) - 1;
            if (g >= gmax || g + 1 == glo) break;
            const uint32_t ls = a.line_start[g], le = a.line_end[g];
            if (e2 > le) break;                       // the occurrence runs into the line separator
            const unsigned long long rec = ((unsigned long long)r << P.gb) | g;
            if (cls == 0) {
                // literal-with-boundaries pattern (mode bits 2/4: \b required left/right; 32/64: the literal's
                // first/last byte is a word byte)
                bool hit = true;
                if (mode & 2u) hit = (s > ls && dev_word(a.text[s - 1])) != ((mode & 32u) != 0);
                if (hit && (mode & 4u)) hit = ((mode & 64u) != 0) != (e2 < le && dev_word(a.text[e2]));
                if (hit) stage_put(s_rec, a.ctr, &a.ctr->n_recs, a.recs, a.rec_cap, 4u, rec);
            } else if (cls == 3) {
                if (P.rx_kind[r] == 0) put_task(make_uint4(r, g, 0u, 0u));
                else {
                    const uint32_t vi = atomicAdd(&a.ctr->n_vm_tasks, 1u);
                    if (vi < a.vm_task_cap) a.vm_tasks[vi] = ((unsigned long long)r << 32) | g;
                    else atomicOr(&a.ctr->overflow, 32u);
                }
            } else {
                // every match starts (class 1) / ends (class 2) with this literal: anchored automaton from the
                // occurrence. Walks die or match within a few bytes; one that is still undecided after
                // `budget` steps (".*" tails) is handed to k_verify, which carries the same walk on.
                const uint4 ad = __ldg(&P.anchor_desc[r]);
                const uint32_t nc = ad.y;
                const uint16_t* __restrict__ T = P.dfa_trans + ad.x;
                const uint8_t* __restrict__ cm = P.rx_acmap + (size_t)r * 256;
                LineWalker<1, true> w;
                w.nc[0] = nc; w.T[0] = T; w.cm[0] = cm; w.ff[0] = 0;
                if (cls == 1) {
                    w.st[0] = s == ls ? (ad.z & 0xFFFFu) : dev_word(a.text[s - 1]) ? ad.w : (ad.z >> 16);
                    const uint32_t stop = min(le, s + a.budget);
                    const long tpos = tail_terminator(a.text, ls, le);
                    if (tpos < 0 || (long)s > tpos || (long)stop <= tpos) w.segment(a.text, a.len, s, stop);
                    else if (w.segment(a.text, a.len, s, (uint32_t)tpos)) { w.symbol(2); w.segment(a.text, a.len, (uint32_t)tpos, stop); }
                    if (stop == le) w.symbol(1);
                } else {
                    w.st[0] = e2 == le ? (ad.z & 0xFFFFu) : dev_word(a.text[e2]) ? ad.w : (ad.z >> 16);
                    const uint32_t stop = e2 - ls > a.budget ? e2 - a.budget : ls;
                    w.segment_rev(a.text, a.len, stop, e2);
                    if (stop == ls) w.symbol(1);
                }
                const uint32_t st = w.st[0];
                if (st == 1u) stage_put(s_rec, a.ctr, &a.ctr->n_recs, a.recs, a.rec_cap, 4u, rec);
                else if (st > 1u)       // undecided: k_verify carries this walk on from where it stopped, in the state it stopped in
                    put_task(make_uint4(r | (cls == 1 ? VT_FWD : VT_REV), g, cls == 1 ? min(le, s + a.budget) : (e2 - ls > a.budget ? e2 - a.budget : ls), st));
            }
        } while (false);
        // no barrier per chunk: a block handles a few chunks and its output fits the staging buffers
        // (entries beyond them go straight to the global queues)
    }
    stage_flush(s_rec, a.ctr, &a.ctr->n_recs, a.recs, a.rec_cap, 4u, true);
    const uint32_t staged = min(s_tn, T_STAGE);
    if (staged) {
        if (threadIdx.x == 0) s_tbase = atomicAdd(&a.ctr->n_tasks, staged);
        __syncthreads();
        for (uint32_t q = threadIdx.x; q < staged; q += blockDim.x) {
            if (s_tbase + q < a.task_cap) a.tasks[s_tbase + q] = s_tbuf[q]; else atomicOr(&a.ctr->overflow, 2u);
        }
    }
}

// ------------------------------------------------------------------ K4: verify tasks
// ------------------------------------------------------------------ tasks grouped by regex
// k_verify_vm interprets per-regex bytecode: with tasks in arrival order the 32 lanes of a warp hold 32 different
// programs and diverge at once. A counting sort on the regex id (histogram, scan, scatter: three tiny kernels; the
// order inside a regex does not matter, the records are sorted later) puts neighbours on the same program; it
// replaced a two-pass device radix sort. (For k_verify's table-driven walks the same grouping, even with the
// automata of a block copied to shared memory, did not pay: 15 + 4 + 15 us of grouping for 72 -> 50 us of walking;
// profiles/r02_scan_variants.jsonl.)
__global__ void k_group_count(const unsigned long long* __restrict__ tasks, const unsigned int* n_ptr, uint32_t cap, uint32_t n_rx, uint32_t* __restrict__ hist) {
    const uint32_t n = min(*n_ptr, cap);
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x)
        atomicAdd(&hist[min((uint32_t)(tasks[t] >> 32), n_rx)], 1u);
}
// one block: hist[0..n_rx] -> exclusive prefix in place, a second copy as the scatter cursors
__global__ void __launch_bounds__(1024) k_group_scan(uint32_t* __restrict__ hist, uint32_t n_bins, uint32_t* __restrict__ cursor) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_bins; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < n_bins ? hist[i] : 0u;
        uint32_t inc = v;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o); if ((threadIdx.x & 31) >= o) inc += y; }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
        __syncthreads();
        if (threadIdx.x < 32) {
            const uint32_t w = warp_tot[threadIdx.x];
            uint32_t wi = w;
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, wi, o); if (threadIdx.x >= o) wi += y; }
            warp_tot[threadIdx.x] = wi - w;
        }
        __syncthreads();
        const uint32_t excl = carry + warp_tot[threadIdx.x >> 5] + inc - v;
        if (i < n_bins) { hist[i] = excl; cursor[i] = excl; }
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
}
__global__ void k_group_scatter(const unsigned long long* __restrict__ tasks, const unsigned int* n_ptr, uint32_t cap, uint32_t n_rx,
                                uint32_t* __restrict__ cursor, unsigned long long* __restrict__ out) {
    const uint32_t n = min(*n_ptr, cap);
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        const unsigned long long task = tasks[t];
        out[atomicAdd(&cursor[min((uint32_t)(task >> 32), n_rx)], 1u)] = task;
    }
}

struct VerifyArgs {
    const uint8_t* text;
    uint64_t len;
    const uint32_t* line_start;
    const uint32_t* line_end;
    const uint4* tasks;
    uint32_t task_cap;
    unsigned long long* recs;
    uint32_t rec_cap;
    Counters* ctr;
    // always-candidates pass
    uint32_t n_lines_total;
    const uint32_t* first_line;
    const uint32_t* n_lines;
    uint32_t n_logs;
    // backtracking verifier
    const unsigned long long* vm_tasks;
    uint32_t vm_task_cap;
    uint32_t* vm_stack;               // pool of deep backtracking stacks, VM_STACK_WORDS words each
    uint32_t vm_pool_n;               // stacks in the pool
};
__global__ void __launch_bounds__(256) k_verify(VerifyArgs a, DevPlan P) {
    // records leave through a warp-wide vote and one atomic per warp: a block-wide staging buffer made every warp wait at
    // its barrier for the block's longest walk (28 % of the warp samples)
    const uint32_t n = min(a.ctr->n_tasks, a.task_cap);
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t rounds = (n + stride - 1) / stride;
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    for (uint32_t it = 0; it < rounds; ++it, t += stride) {
        bool hit_ = false;
        uint32_t r_ = 0, g_ = 0;
        if (t < n) {
            const uint4 task = a.tasks[t];
            const uint32_t r = task.x & ~(VT_FWD | VT_REV), g = task.y;
            const uint32_t ls = a.line_start[g], le = a.line_end[g];
            bool hit = false;
            if (task.x & (VT_FWD | VT_REV)) {
                // the anchored automaton of k_resolve, from the position and state its budget ran out in: the rest of the
                // line (forwards) or the part before the occurrence (backwards), not the whole line from its start
                const uint4 ad = __ldg(&P.anchor_desc[r]);
                LineWalker<1, true> w;
                w.nc[0] = ad.y; w.T[0] = P.dfa_trans + ad.x; w.cm[0] = P.rx_acmap + (size_t)r * 256; w.ff[0] = 0;
                w.st[0] = task.w;
                const uint32_t at = task.z;
                if (task.x & VT_FWD) {
                    const long tpos = tail_terminator(a.text, ls, le);
                    if (tpos < 0 || (long)at > tpos) w.segment(a.text, a.len, at, le);
                    else if (w.segment(a.text, a.len, at, (uint32_t)tpos)) { w.symbol(2); w.segment(a.text, a.len, (uint32_t)tpos, le); }
                } else {
                    w.segment_rev(a.text, a.len, ls, at);
                }
                w.symbol(1);
                hit = w.st[0] == 1u;
            } else if (P.rx_kind[r] == 0) {
                hit = dfa_walk(P, r, a.text, a.len, ls, le);
            }
            hit_ = hit; r_ = r; g_ = g;
        }
        emit_record(a.ctr, a.recs, a.rec_cap, P.gb, hit_, r_, g_);
    }
}
// regexes without a usable literal: every (untrimmed) line is a candidate
__global__ void __launch_bounds__(256) k_verify_always(VerifyArgs a, DevPlan P) {
    a.n_lines_total = min(a.ctr->n_lines_total, a.n_lines_total);     // host passes the capacity
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t rounds = (a.n_lines_total + stride - 1) / stride;
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    for (uint32_t it = 0; it < rounds; ++it, g += stride) {
        bool live = g < a.n_lines_total;
        uint32_t ls = 0, le = 0;
        if (live) {
            uint32_t k = a.n_logs == 1 ? 0 : upper_bound_dev<uint32_t>(a.first_line, a.n_logs + 1, g) - 1;
            live = k < a.n_logs && g - a.first_line[k] < a.n_lines[k];
            if (live) { ls = a.line_start[g]; le = a.line_end[g]; }
        }
        for (uint32_t q = 0; q < P.n_always; ++q) {
            const uint32_t r = P.always_list[q];
            bool hit = false;
            if (live && P.rx_kind[r] == 0) hit = dfa_walk(P, r, a.text, a.len, ls, le);
            emit_record(a.ctr, a.recs, a.rec_cap, P.gb, hit, r, g);
        }
    }
}

// backtracking verifier: look-around, back-references, possessive/atomic, MULTILINE anchors, DFA blow-ups
// Backtracking stack of one evaluation on the device: the first VM_TOP_WORDS words live in the thread's local memory
// (cached write-back in L1); an evaluation that needs more takes one deep stack (lpx::VM_STACK_WORDS words of global
// memory) from a pool and gives up, counted in n_budget like an exhausted step budget, when the pool is empty.
// Greedy loops push one range entry, not one entry per position, so almost every evaluation stays inside the
// top part and the threads of the VM kernels cost no global memory.
constexpr uint32_t VM_TOP_WORDS = 256;
constexpr uint32_t VM_POOL_STACKS = 4096;
struct VmStack {
    uint32_t* top;          // VM_TOP_WORDS words, local memory
    uint32_t** deep;        // the thread's deep stack once it has one (index i >= VM_TOP_WORDS lives there)
    uint32_t* pool;
    uint32_t pool_n;
    Counters* ctr;
    __device__ __forceinline__ uint32_t& operator[](uint32_t i) const { return i < VM_TOP_WORDS ? top[i] : (*deep)[i]; }
};
__device__ __noinline__ bool vm_stack_grow(VmStack& s, uint32_t& cap) {
    if (cap >= lpx::VM_STACK_WORDS) return false;
    if (!*s.deep) {
        const uint32_t k = atomicAdd(&s.ctr->vm_pool_next, 1u);
        if (k >= s.pool_n) return false;
        *s.deep = s.pool + (size_t)k * lpx::VM_STACK_WORDS;
    }
    cap = lpx::VM_STACK_WORDS;
    return true;
}
__device__ __forceinline__ int vm_eval_span(const DevPlan& P, uint32_t r, const uint8_t* __restrict__ text, uint32_t ls, uint32_t le,
                                            uint32_t* pool, uint32_t pool_n, Counters* ctr, uint32_t* span) {
    lpx::VmView v;
    v.code = reinterpret_cast<const lpx::VmCode*>(P.vm_code) + P.rx_vm_off[r];
    v.set_off = P.vm_set_off + P.rx_vm_set_base[r];
    v.set_data = P.vm_set_ranges;
    v.n_slots = P.rx_vm_slots[r];
    v.first = P.rx_vm_first + (size_t)r * 8;
    uint32_t top[VM_TOP_WORDS];
    uint32_t* deep = nullptr;
    const int res = lpx::vm_find(v, text, ls, le, VmStack{top, &deep, pool, pool_n, ctr}, VM_TOP_WORDS, lpx::VM_MAX_STEPS, span);
    if (res < 0) { atomicOr(&ctr->overflow, 16u); atomicAdd(&ctr->n_budget, 1u); }
    return res;
}
__device__ __forceinline__ bool vm_eval(const DevPlan& P, uint32_t r, const uint8_t* __restrict__ text, uint32_t ls, uint32_t le,
                                        uint32_t* pool, uint32_t pool_n, Counters* ctr) {
    return vm_eval_span(P, r, text, ls, le, pool, pool_n, ctr, nullptr) > 0;
}
// LP_OPT_SPANS: Matcher.start()/end() of the first find() of each event's primary regex on its line.
__global__ void __launch_bounds__(128) k_spans(const unsigned long long* __restrict__ ev_key, const Counters* ctr, uint32_t m_cap, DevPlan P,
                                               const uint8_t* __restrict__ text, const uint32_t* __restrict__ line_start,
                                               const uint32_t* __restrict__ line_end, uint32_t* __restrict__ vm_pool, uint32_t vm_pool_n, Counters* ctr_w,
                                               uint32_t* __restrict__ span_start, uint32_t* __restrict__ span_end) {
    const uint32_t m = min(ctr->n_events, m_cap);
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < m; e += stride) {
        const unsigned long long k = ev_key[e];
        const uint32_t g = (uint32_t)(k >> P.pb), p = (uint32_t)(k & ((1ull << P.pb) - 1ull));
        const uint32_t r = P.pat_primary[p], ls = line_start[g], le = line_end[g];
        uint32_t span[2] = {ls, ls};
        const int res = vm_eval_span(P, r, text, ls, le, vm_pool, vm_pool_n, ctr_w, span);
        span_start[e] = res > 0 ? span[0] - ls : 0xFFFFFFFFu;     // the automaton said the line matches; a miss here would be a bug
        span_end[e] = res > 0 ? span[1] - ls : 0xFFFFFFFFu;
    }
}
// One evaluation per lane, but only L lanes of a warp take one: the lanes of a warp interpret in lock step, so a warp
// costs the SUM of its lanes' divergent paths, and with a few ten thousand evaluations on an otherwise idle GPU the
// longest warp is the kernel's duration (424 us with 32 evaluations per warp on C3, whose longest single one takes
// 215 interpreter steps). L = evaluations / resident warps, rounded up.
__global__ void __launch_bounds__(128) k_verify_vm(VerifyArgs a, DevPlan P) {
    const uint32_t n = min(a.ctr->n_vm_tasks, a.vm_task_cap);
    const uint32_t warps = gridDim.x * (blockDim.x >> 5);
    const uint32_t L = min(32u, max(1u, (n + warps - 1) / warps));
    const uint32_t lane = threadIdx.x & 31, w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    for (uint32_t base = 0; base < n; base += warps * L) {       // uniform trip count: emit_record is a warp collective
        const uint32_t t = base + w * L + lane;
        bool hit = false;
        uint32_t r = 0, g = 0;
        if (lane < L && t < n) {
            const unsigned long long task = a.vm_tasks[t];
            r = (uint32_t)(task >> 32);
            g = (uint32_t)task;
            hit = vm_eval(P, r, a.text, a.line_start[g], a.line_end[g], a.vm_stack, a.vm_pool_n, a.ctr);
        }
        emit_record(a.ctr, a.recs, a.rec_cap, P.gb, hit, r, g);
    }
}
__global__ void __launch_bounds__(128) k_verify_vm_always(VerifyArgs a, DevPlan P) {
    a.n_lines_total = min(a.ctr->n_lines_total, a.n_lines_total);
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t rounds = (a.n_lines_total + stride - 1) / stride;
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    for (uint32_t it = 0; it < rounds; ++it, g += stride) {
        bool live = g < a.n_lines_total;
        uint32_t ls = 0, le = 0;
        if (live) {
            uint32_t k = a.n_logs == 1 ? 0 : upper_bound_dev<uint32_t>(a.first_line, a.n_logs + 1, g) - 1;
            live = k < a.n_logs && g - a.first_line[k] < a.n_lines[k];
            if (live) { ls = a.line_start[g]; le = a.line_end[g]; }
        }
        for (uint32_t q = 0; q < P.n_always_vm; ++q) {
            const uint32_t r = P.always_vm_list[q];
            const bool hit = live && vm_eval(P, r, a.text, ls, le, a.vm_stack, a.vm_pool_n, a.ctr);
            emit_record(a.ctr, a.recs, a.rec_cap, P.gb, hit, r, g);
        }
    }
}


// ------------------------------------------------------------------ always-on DFA scan
// The multi-pattern DFA scan proper: every line is walked once through up to DS_MAXG product
// automata (context regexes of CAS:27-34 + every DFA pattern without a usable literal, AS:93-95).
// Transition tables, class maps and result masks are resident in shared memory for the life of the
// CTA. One thread walks one line; it streams the line through registers with 8-byte loads, one
// word prefetched ahead. (Lines are ~135 bytes and a walk is one dependent shared-memory lookup per
// byte, so the kernel is bound by that chain, not by how the text arrives; the TMA-staged tiles of
// the scan are used by k_fscan, where every byte of a tile is needed.)
constexpr uint32_t DS_THREADS = 256, DS_MAXG = 4;

struct DfaScanArgs {
    const uint8_t* text;
    uint64_t len;
    const uint32_t* line_start;
    const uint32_t* line_end;
    uint32_t n_lines_total;
    const uint32_t* first_line;
    const uint32_t* n_lines;
    uint32_t n_logs;
    const uint32_t* blob[DS_MAXG];
    uint32_t blob_words[DS_MAXG];
    uint32_t n_groups;
    const uint8_t* rx_record;
    uint8_t* ctx_flags;
    int write_flags;             // 0 leave alone, 1 store every line's flag byte, 2 OR this launch's bits in
    uint32_t skip1[8];           // first bytes on which every group of this launch is decided without a hit: such lines are not walked
    const uint8_t* line_skip;    // [lines] 1 = k_fscan saw such a first byte while the text was in shared memory (NULL: look at the text here)
    const uint32_t* scan_list;   // the lines to walk, compacted by k_expand (ctr->n_scan of them), or NULL: every line is visited;
                                 // with the list the flag bytes of the other lines must have been cleared beforehand
    unsigned long long* recs;
    uint32_t rec_cap;
    uint32_t gb;
    Counters* ctr;
};

template <int NG>
__global__ void __launch_bounds__(DS_THREADS) k_dfa_scan(DfaScanArgs a) {
    extern __shared__ __align__(16) uint32_t ds_tabs[];
    __shared__ uint32_t s_goff[DS_MAXG], s_recmask[DS_MAXG];
    __shared__ int s_ctxbit[DS_MAXG][4];
    {
        uint32_t off = 0;
        for (int g = 0; g < NG; ++g) {
            for (uint32_t i = threadIdx.x; i < a.blob_words[g]; i += DS_THREADS) ds_tabs[off + i] = __ldg(a.blob[g] + i);
            if (threadIdx.x == 0) s_goff[g] = off;
            off += (a.blob_words[g] + 3u) & ~3u;
        }
    }
    __syncthreads();
    if (threadIdx.x < NG) {
        const uint32_t* h = ds_tabs + s_goff[threadIdx.x];
        uint32_t rm = 0;
        int cb[4] = {-1, -1, -1, -1};
        for (uint32_t k = 0; k < h[4]; ++k) {
            const uint32_t r = h[5 + k];
            if (a.rx_record[r]) rm |= 1u << k;
            if (r < 4) cb[r] = (int)k;
        }
        s_recmask[threadIdx.x] = rm;
        for (int c = 0; c < 4; ++c) s_ctxbit[threadIdx.x][c] = cb[c];
    }
    __syncthreads();

    const uint8_t* cm[NG];
    const uint16_t* T[NG];
    const uint32_t* smask[NG];
    uint32_t nc[NG], ff[NG], st0[NG];
#pragma unroll
    for (int g = 0; g < NG; ++g) {
        const uint32_t* h = ds_tabs + s_goff[g];
        nc[g] = h[1]; st0[g] = h[2]; ff[g] = h[3];
        cm[g] = reinterpret_cast<const uint8_t*>(h + 37);
        smask[g] = h + 101;
        T[g] = reinterpret_cast<const uint16_t*>(h + 101 + h[0]);
    }

    const uint32_t stride = gridDim.x * DS_THREADS;
    const uint32_t n_total = min(a.ctr->n_lines_total, a.n_lines_total);     // host passes the capacity
    const uint32_t n_iter = a.scan_list ? min(a.ctr->n_scan, a.n_lines_total) : n_total;
    for (uint32_t it = blockIdx.x * DS_THREADS + threadIdx.x; it < n_iter; it += stride) {
        const uint32_t g = a.scan_list ? a.scan_list[it] : it;
        if (g >= n_total) continue;
        const uint32_t k = a.n_logs == 1 ? 0 : upper_bound_dev<uint32_t>(a.first_line, a.n_logs + 1, g) - 1;
        const bool live = k < a.n_logs && g - a.first_line[k] < a.n_lines[k];
        uint32_t flags = 0;
        bool walk = live;
        uint32_t ls = 0, le = 0;
        if (live) {
            ls = a.line_start[g]; le = a.line_end[g];
            if (a.scan_list) walk = true;
            else if (a.line_skip) walk = a.line_skip[g] == 0;
            else if (le > ls) { const uint32_t b = a.text[ls]; walk = ((a.skip1[b >> 5] >> (b & 31u)) & 1u) == 0; }
        }
        if (walk) {
            LineWalker<NG, false> w;
#pragma unroll
            for (int q = 0; q < NG; ++q) { w.cm[q] = cm[q]; w.T[q] = T[q]; w.nc[q] = nc[q]; w.ff[q] = ff[q]; w.st[q] = st0[q]; }
            w.run(a.text, a.len, ls, le);
#pragma unroll
            for (int q = 0; q < NG; ++q) {
                const uint32_t mask = smask[q][w.st[q]];
#pragma unroll
                for (int c = 0; c < 4; ++c) { const int bit = s_ctxbit[q][c]; if (bit >= 0) flags |= ((mask >> bit) & 1u) << c; }
                uint32_t pend = mask & s_recmask[q];
                const uint32_t* rxs = ds_tabs + s_goff[q] + 5;
                while (pend) {        // records of literal-free user patterns are sparse: one atomic each
                    const uint32_t r = rxs[__ffs(pend) - 1];
                    pend &= pend - 1;
                    const uint32_t idx = atomicAdd(&a.ctr->n_recs, 1u);
                    if (idx < a.rec_cap) a.recs[idx] = ((unsigned long long)r << a.gb) | g;
                    else atomicOr(&a.ctr->overflow, 4u);
                }
            }
        }
        // every line's flag byte is stored by the first launch (it doubles as the clearing pass); further
        // launches, when the groups do not fit one, add their bits (a line belongs to one thread)
        if (a.write_flags == 1 && (!a.scan_list || flags)) a.ctx_flags[g] = (uint8_t)flags;
        else if (a.write_flags == 2 && flags) a.ctx_flags[g] = (uint8_t)(a.ctx_flags[g] | flags);
    }
}
