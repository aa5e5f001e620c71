// Record post-processing, event list, scoring (ScoringService / ContextAnalysisService / FrequencyTrackingService arithmetic), multi-GPU composition kernels.
// Part of engine.cu's translation unit (included inside its anonymous namespace); not a stand-alone header.
// ------------------------------------------------------------------ record post-processing
// rx_begin[r] = first record of regex r in the sorted unique records. One warp per regex: a 32-ary search, so the chain
// of dependent loads is 4 long for 10^6 records instead of 20 (the kernel is pure load latency).
__global__ void k_rx_bounds(const unsigned long long* __restrict__ keys, const Counters* ctr, uint32_t n_rx, uint32_t gb, uint32_t* __restrict__ rx_begin) {
    const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (r > n_rx) return;       // warp-uniform
    const unsigned long long want = (unsigned long long)r << gb;
    uint32_t lo = 0, hi = ctr->n_unique;          // answer in [lo, hi]
    while (hi - lo > 32) {
        const uint32_t step = (hi - lo + 32) / 33;                                  // probes at lo + step * (lane + 1) - 1
        const uint32_t at = min(lo + step * (lane + 1) - 1, hi - 1);
        const unsigned below = __ballot_sync(0xffffffffu, keys[at] < want);       // monotone: a prefix of the lanes
        const uint32_t k = __popc(below);
        const uint32_t nlo = k ? min(lo + step * k, hi) : lo;                     // keys[nlo - 1] < want
        hi = k < 32 ? min(lo + step * (k + 1) - 1, hi - 1) : hi;                  // keys[hi] >= want (or hi is the end)
        lo = nlo;
    }
    const unsigned below = __ballot_sync(0xffffffffu, lo + lane < hi && keys[lo + lane] < want);
    if (lane == 0) rx_begin[r] = lo + __popc(below);
}
__global__ void k_ctx_flags(const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ rx_begin, uint32_t gb, uint32_t* __restrict__ flags32) {
    const uint32_t n = rx_begin[4];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned long long k = keys[i];
        const uint32_t r = (uint32_t)(k >> gb), g = (uint32_t)(k & ((1ull << gb) - 1ull));
        atomicOr(&flags32[g >> 2], (1u << r) << ((g & 3) * 8));
    }
}
// Events are emitted for lines [core_lo, G - halo_after): the whole text unless it is a line-range
// shard, whose halo lines only feed the windows of its own lines.
__global__ void k_emit_events(const unsigned long long* __restrict__ keys, DevPlan P,
                              unsigned long long* __restrict__ ev_key, uint32_t ev_cap, Counters* ctr, uint32_t core_lo, uint32_t halo_after) {
    const uint32_t n = ctr->n_unique;
    const uint32_t core_hi = ctr->n_lines_total > halo_after ? ctr->n_lines_total - halo_after : 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned long long k = keys[i];
        const uint32_t r = (uint32_t)min(k >> P.gb, 0xFFFFFFFFull), g = (uint32_t)(k & ((1ull << P.gb) - 1ull));
        if (r >= P.n_rx) continue;     // padding sentinel
        if (g < core_lo || g >= core_hi) continue;
        const uint32_t pb = P.rxp_begin[r], pe = P.rxp_begin[r + 1];
        if (pb == pe) continue;
        const uint32_t base = atomicAdd(&ctr->n_events, pe - pb);
        for (uint32_t j = pb; j < pe; ++j) {
            const uint32_t idx = base + (j - pb);
            if (idx < ev_cap) ev_key[idx] = ((unsigned long long)g << P.pb) | P.rxp_pat[j];
            else atomicOr(&ctr->overflow, 8u);
        }
    }
}
__global__ void k_log_event_begin(const unsigned long long* __restrict__ ev_key, const Counters* ctr, const uint32_t* __restrict__ first_line,
                                  uint32_t n_logs, uint32_t pb, unsigned long long* __restrict__ begin) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > n_logs) return;
    const uint32_t m = ctr->n_events;
    begin[k] = lower_bound_dev<unsigned long long>(ev_key, m, (unsigned long long)first_line[k] << pb);
}

// ------------------------------------------------------------------ K5: scoring (fp64, no contraction)
struct ScoreArgs {
    const unsigned long long* ev_key;
    uint32_t m_cap;                        // threads launched; the live count comes from ctr->n_events
    const Counters* ctr;
    uint32_t* ev_log;                     // out: log of each event (padding: n_logs), for the per-log ranking
    uint32_t* ev_idx;                     // out: identity permutation, the value column of the ranking sort
    unsigned long long* freq_delta;       // out: events per pattern (zeroed before the launch)
    uint32_t n_logs;
    const uint32_t* rank_in_slot;
    const unsigned long long* prior;      // [n_pat] in-window counts at call start (+ lower ranks), may be null
    const uint32_t* first_line;
    const uint32_t* n_lines;
    const unsigned long long* keys;       // unique (regex, line) records
    const uint32_t* rx_begin;
    const uint8_t* ctx_flags;
    uint32_t* out_line;
    uint32_t* out_pat;
    double* out_score;
    double* out_factors;                  // may be null
    unsigned long long* sort_key;         // descending-score key for ranking
    // line-range shard (n_global != 0): position of text line 0 in the whole log, the whole log's line
    // count, the first own line, and the per-sequence chain thresholds composed over all shards
    long long line_base;
    long long n_global;
    const long long* range_dev;      // {line_base, n_global} composed on the device from the gathered shard summaries (overrides the two above)
    uint32_t core_lo;
    const long long* seq_T;
};

// lower_bound_dev by a group of SG consecutive lanes (all of them call it with the same arguments and get the same
// answer): SG probes per round split the range into SG + 1 parts, so a list of 5*10^4 records is searched in 5
// dependent loads instead of 16. k_score is such searches chained one after the other (70 % of its warp samples
// wait on them); a whole warp per event made the chains shorter still but the kernel slower (32 times the warps).
template <int SG, class T>
__device__ __forceinline__ uint32_t group_lower_bound(const T* __restrict__ a, uint32_t n, T v) {
    if constexpr (SG == 1) return lower_bound_dev<T>(a, n, v);
    const uint32_t lane = threadIdx.x & 31, sub = lane & (SG - 1), gbase = lane & ~(uint32_t)(SG - 1);
    const unsigned gmask = ((1u << SG) - 1u) << gbase;
    uint32_t lo = 0, hi = n;                    // the answer lies in [lo, hi]
    while (hi - lo > SG) {
        const uint32_t step = (hi - lo + SG) / (SG + 1);
        const uint32_t idx = lo + (sub + 1) * step - 1;
        const bool less = idx < hi && a[idx] < v;
        const uint32_t k = __popc(__ballot_sync(gmask, less));        // the list is sorted: the first k probes are < v
        const uint32_t nlo = lo + k * step, nhi = k < SG ? min(hi, lo + (k + 1) * step - 1) : hi;
        lo = nlo; hi = nhi;
    }
    const uint32_t idx = lo + sub;
    const bool less = idx < hi && a[idx] < v;
    return lo + __popc(__ballot_sync(gmask, less));
}
template <int SG>
__device__ __forceinline__ bool group_list_any_in(const unsigned long long* __restrict__ keys, uint32_t b, uint32_t e, uint32_t r, uint32_t gb,
                                                  uint32_t lo, uint32_t hi /* [lo,hi) global lines */) {
    if (lo >= hi) return false;
    const uint32_t i = b + group_lower_bound<SG, unsigned long long>(keys + b, e - b, ((unsigned long long)r << gb) | lo);
    return i < e && (uint32_t)(keys[i] & ((1ull << gb) - 1ull)) < hi;
}
// context factor, CAS:46-117, accumulated in the reference's per-line order. The walk over the context lines is a
// serial chain of fp64 additions that cannot be split between lanes without changing the rounding; it is done with
// one thread per event (a warp of the block takes the block's events) so that a wide window (C4: +-500 lines) costs
// its issue slots once per 32 events, not once per group.
__device__ __forceinline__ double ctx_factor(const ScoreArgs& a, const DevPlan& P, const ScoreCfg& C, uint32_t g, uint32_t p) {
    const uint32_t k = a.n_logs == 1 ? 0u : upper_bound_dev<uint32_t>(a.first_line, a.n_logs + 1, g) - 1;
    const uint32_t first = a.first_line[k];
    const long long N = a.n_lines[k];
    const long long i = (long long)g - first;
    long long b0 = i, a1 = i + 1;
    if (P.pat_has_ctx[p]) { b0 = max(0ll, i - (long long)P.pat_before[p]); a1 = min(N, i + 1 + (long long)P.pat_after[p]); }
    double score = 0.0;
    int err = 0, stack = 0;
    const uint8_t* __restrict__ fl = a.ctx_flags + first;
    for (long long l = b0; l < a1; ++l) {
        const uint32_t f = fl[l];
        if (f & 1u) { ++err; score = __dadd_rn(score, 0.4); }
        else if (f & 2u) { score = __dadd_rn(score, 0.2); }
        if (f & 4u) { ++stack; score = __dadd_rn(score, 0.1); }
        if (f & 8u) { score = __dadd_rn(score, 0.3); }
    }
    if (stack > 0) score = __dadd_rn(score, fmin(__dmul_rn((double)stack, 0.1), 0.5));
    const int total = (int)(a1 - b0);
    if (total > 10 && (double)(stack + err) > __dmul_rn((double)total, 0.7)) score = __dmul_rn(score, 0.8);
    double ctx = __dadd_rn(1.0, score);
    if (ctx > C.max_ctx) ctx = C.max_ctx;
    if (total <= 0) ctx = 1.0;
    return ctx;
}
// SG lanes per event: every lane of the group runs the same code on the same values (the searches are cooperative, see
// group_lower_bound), its first lane writes. The arithmetic is the reference's, in its order (fp64, no contraction).
template <int SG, int SC_THREADS>
__global__ void __launch_bounds__(SC_THREADS) k_score(ScoreArgs a, DevPlan P, ScoreCfg C) {
    constexpr int SC_EVENTS = SC_THREADS / SG;       // events of one block
    __shared__ double s_ctx[SG > 1 ? SC_EVENTS : 1];
    if (SG > 1 && threadIdx.x < SC_EVENTS) {      // the block's context factors, one thread per event
        const uint32_t e2 = blockIdx.x * SC_EVENTS + threadIdx.x;
        if (e2 < a.m_cap && e2 < a.ctr->n_events) {
            const unsigned long long ek2 = a.ev_key[e2];
            s_ctx[threadIdx.x] = ctx_factor(a, P, C, (uint32_t)(ek2 >> P.pb), (uint32_t)(ek2 & ((1ull << P.pb) - 1ull)));
        }
    }
    if (SG > 1) __syncthreads();
    const uint32_t e = blockIdx.x * SC_EVENTS + threadIdx.x / SG;
    const bool writer = (threadIdx.x & (SG - 1)) == 0;
    if (e >= a.m_cap) return;
    if (writer) a.ev_idx[e] = e;
    if (e >= a.ctr->n_events) { if (writer) { a.sort_key[e] = ~0ull; a.ev_log[e] = a.n_logs; } return; }      // padding sorts last
    if (a.range_dev) { a.line_base = a.range_dev[0]; a.n_global = a.range_dev[1]; }
    const unsigned long long ek = a.ev_key[e];
    const uint32_t g = (uint32_t)(ek >> P.pb), p = (uint32_t)(ek & ((1ull << P.pb) - 1ull));
    const unsigned long long gmask = (1ull << P.gb) - 1ull;
    const uint32_t k = a.n_logs == 1 ? 0u : upper_bound_dev<uint32_t>(a.first_line, a.n_logs + 1, g) - 1;
    if (writer) { a.ev_log[e] = k; atomicAdd(&a.freq_delta[p], 1ull); }        // FTS:41-56: one recordPatternMatch per event, applied by the host after the call
    const uint32_t first = a.first_line[k];
    const long long N = a.n_lines[k];
    const long long i = (long long)g - first;

    const double base = P.pat_conf[p];
    const double sev = P.pat_sev[p];

    // chronological factor, SS:123-151
    const double pos = a.n_global ? __ddiv_rn((double)(i + a.line_base), (double)a.n_global) : __ddiv_rn((double)i, (double)N);
    double chrono;
    if (pos <= C.early) {
        const double range = __dsub_rn(C.max_early, 1.5);
        chrono = __dadd_rn(1.5, __dmul_rn(__dsub_rn(C.early, pos), __ddiv_rn(range, C.early)));
    } else if (pos <= C.penalty) {
        const double mid = __dsub_rn(C.penalty, C.early);
        chrono = __dadd_rn(1.0, __dmul_rn(__dsub_rn(C.penalty, pos), __ddiv_rn(0.5, mid)));
    } else {
        chrono = __dadd_rn(0.5, __dsub_rn(1.0, pos));
    }

    // proximity factor, SS:161-190 + 315-347: closest hit of each secondary inside its window
    double prox = 1.0;
    {
        const uint32_t sb = P.sec_begin[p], se = P.sec_begin[p + 1];
        if (sb != se) {
            double total = 0.0;
            for (uint32_t s = sb; s < se; ++s) {
                const long long win = min((long long)C.max_window, (long long)P.sec_window[s]);
                const long long lo = max(0ll, i - win), hi = min(N, i + win + 1);   // local [lo,hi)
                if (lo >= hi) continue;
                const uint32_t r = P.sec_rx[s];
                const uint32_t b = a.rx_begin[r], en = a.rx_begin[r + 1];
                const unsigned long long* kk = a.keys + b;
                const uint32_t cnt = en - b;
                const unsigned long long self = ((unsigned long long)r << P.gb) | g;
                const uint32_t at = group_lower_bound<SG, unsigned long long>(kk, cnt, self);
                long long best = -1;
                if (at > 0) {
                    const long long l = (long long)(uint32_t)(kk[at - 1] & gmask) - first;
                    if (l >= lo) best = i - l;
                }
                uint32_t nx = at;
                if (nx < cnt && kk[nx] == self) ++nx;
                if (nx < cnt) {
                    const long long l = (long long)(uint32_t)(kk[nx] & gmask) - first;
                    if (l < hi) { const long long d = l - i; if (best < 0 || d < best) best = d; }
                }
                if (best >= 0) total = __dadd_rn(total, __dmul_rn(P.sec_weight[s], C.decay_lut[best]));
            }
            prox = __dadd_rn(1.0, total);
        }
    }

    // temporal factor, SS:199-305 (backward greedy walk == forward threshold)
    double temporal = 1.0;
    {
        const uint32_t qb = P.seq_begin[p], qe = P.seq_begin[p + 1];
        if (qb != qe) {
            double total = 0.0;
            for (uint32_t q = qb; q < qe; ++q) {
                const uint32_t eb = P.seq_ev_begin[q], ee = P.seq_ev_begin[q + 1];
                if (eb == ee) continue;                                  // SS:233-235
                const uint32_t rl = P.ev_rx[ee - 1];
                const long long nlo = max(0ll, i - 5), nhi = min(N, i + 6);  // SS:275-277
                if (!group_list_any_in<SG>(a.keys, a.rx_begin[rl], a.rx_begin[rl + 1], rl, P.gb, first + (uint32_t)nlo, first + (uint32_t)nhi)) continue;
                bool ok = true;
                long long T = -1;   // local line of the earliest feasible position of the previous event
                if (a.seq_T) ok = a.seq_T[q] < i + a.line_base;
                else for (uint32_t j = eb; j + 1 < ee && ok; ++j) {
                    const uint32_t r = P.ev_rx[j];
                    const uint32_t b = a.rx_begin[r], en = a.rx_begin[r + 1];
                    const unsigned long long want = ((unsigned long long)r << P.gb) | (uint32_t)(first + T + 1);
                    const uint32_t at = b + group_lower_bound<SG, unsigned long long>(a.keys + b, en - b, want);
                    if (at >= en) { ok = false; break; }
                    const long long l = (long long)(uint32_t)(a.keys[at] & gmask) - first;
                    if (l >= i) { ok = false; break; }     // must end strictly before the primary line
                    T = l;
                }
                if (ok) total = __dadd_rn(total, P.seq_bonus[q]);
            }
            temporal = __dadd_rn(1.0, total);
        }
    }

    const double ctx = SG > 1 ? s_ctx[threadIdx.x / SG] : ctx_factor(a, P, C, g, p);      // groups: computed above, one thread per event

    // frequency penalty, FTS:64-93 with the count read before this match is recorded (SS:84-88)
    double pen = 0.0;
    {
        const int slot = P.pat_slot[p];
        if (slot >= 0) {
            const unsigned long long pr = a.prior ? a.prior[P.slot_first_pat[slot]] : 0ull;
            // events of this frequency slot that come earlier in discovery order (whole batch): for
            // every pattern q sharing the id, the records of its primary on earlier lines, plus the
            // one on this line when q precedes p
            unsigned long long earlier = 0;
            for (uint32_t j = P.slot_begin[slot]; j < P.slot_begin[slot + 1]; ++j) {
                const uint32_t q = P.slot_pats[j], rq = P.pat_primary[q];
                const uint32_t b = a.rx_begin[rq], cnt2 = a.rx_begin[rq + 1] - b;
                const unsigned long long want = ((unsigned long long)rq << P.gb) | g;
                const uint32_t at = group_lower_bound<SG, unsigned long long>(a.keys + b, cnt2, want);
                earlier += at;
                if (a.n_global) earlier -= group_lower_bound<SG, unsigned long long>(a.keys + b, cnt2, ((unsigned long long)rq << P.gb) | a.core_lo);
                if (q < p && at < cnt2 && a.keys[b + at] == want) earlier += 1;
            }
            const unsigned long long cnt = pr + earlier;
            if (cnt > 0) {
                const double rate = __ddiv_rn((double)cnt, (double)C.hours);
                if (rate > C.freq_thr) pen = fmin(C.max_pen, __ddiv_rn(__dsub_rn(rate, C.freq_thr), C.freq_thr));
            }
        }
    }

    double s = __dmul_rn(base, sev);
    s = __dmul_rn(s, chrono);
    s = __dmul_rn(s, prox);
    s = __dmul_rn(s, temporal);
    s = __dmul_rn(s, ctx);
    s = __dmul_rn(s, __dsub_rn(1.0, pen));

    if (!writer) return;
    a.out_line[e] = (uint32_t)(i + a.line_base + 1);
    a.out_pat[e] = p;
    a.out_score[e] = s;
    if (a.out_factors) {
        double* f = a.out_factors + (size_t)e * 7;
        f[0] = base; f[1] = sev; f[2] = chrono; f[3] = prox; f[4] = temporal; f[5] = ctx; f[6] = pen;
    }
    // order-preserving key: larger score first; -0.0/+0.0 and NaN keep IEEE total order
    unsigned long long bits = (unsigned long long)__double_as_longlong(s);
    bits = (bits & 0x8000000000000000ull) ? ~bits : (bits | 0x8000000000000000ull);
    a.sort_key[e] = ~bits;
}
__global__ void k_write_rank(const uint32_t* __restrict__ order, const uint32_t* __restrict__ ev_log,
                             const unsigned long long* __restrict__ log_begin, const Counters* ctr, uint32_t* __restrict__ rank) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ctr->n_events) return;
    const uint32_t e = order[j];
    rank[e] = j - (uint32_t)log_begin[ev_log[e]];
}
// order[final position] = event, from the positions k_tile_rank gave the first *n_valid events; the padding behind them
// keeps its place (its log index is n_logs, so the stable sort by log leaves it last)
__global__ void k_invert_rank(const uint32_t* __restrict__ rank, const Counters* ctr, uint32_t n, uint32_t* __restrict__ order) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    if (j < ctr->n_events) order[rank[j]] = j; else order[j] = j;
}
__global__ void k_gather_u32(const uint32_t* __restrict__ src, const uint32_t* __restrict__ idx, uint32_t n, uint32_t* __restrict__ dst) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) dst[j] = src[idx[j]];
}
// Line-range shards: the sequence chain of SS:250-258 (forward form) restricted to this
// shard's own lines, for every state the chain can be in when it enters the shard. tab[off[q] + s] =
// (state on exit) << 32 | (1 + own-line index where the last-but-one event was reached, 0 if not reached).
__global__ void k_range_tables(const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ rx_begin, DevPlan P, const Counters* ctr,
                               uint32_t n_seq, const uint32_t* __restrict__ tab_off, uint32_t core_lo, uint32_t halo_after,
                               unsigned long long* __restrict__ tab) {
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_seq) return;
    const uint32_t core_hi = ctr->n_lines_total > halo_after ? ctr->n_lines_total - halo_after : 0;
    const uint32_t eb = P.seq_ev_begin[q], ee = P.seq_ev_begin[q + 1];
    const uint32_t k = ee - eb;
    if (k < 2) return;
    const unsigned long long gmask = (1ull << P.gb) - 1ull;
    for (uint32_t s = 0; s + 1 < k; ++s) {
        uint32_t j = s;
        long long T = (long long)core_lo - 1;
        while (j + 1 < k) {
            const uint32_t r = P.ev_rx[eb + j];
            const uint32_t b = rx_begin[r], en = rx_begin[r + 1];
            const unsigned long long want = ((unsigned long long)r << P.gb) | (uint32_t)(T + 1);
            const uint32_t at = b + lower_bound_dev<unsigned long long>(keys + b, en - b, want);
            if (at >= en) break;
            const uint32_t l = (uint32_t)(keys[at] & gmask);
            if (l >= core_hi) break;
            T = l;
            ++j;
        }
        tab[tab_off[q] + s] = ((unsigned long long)j << 32) | (j + 1 == k ? (unsigned long long)(T - core_lo + 1) : 0ull);
    }
}
// Multi-GPU batches with the exchange kept on the device: `all` holds every rank's [n_pat counts, overflow word]
// (all-gathered); prior_out[p] = window prior + the counts of the lower ranks; *any_flag = OR of the overflow words.
__global__ void k_prior_prefix(const unsigned long long* __restrict__ all, uint32_t n_pat, uint32_t rank, uint32_t world,
                               const unsigned long long* __restrict__ prior_in, unsigned long long* __restrict__ prior_out,
                               unsigned long long* __restrict__ any_flag, DevPlan P) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)n_pat + 1;
    if (p < n_pat) {
        // FrequencyTrackingService keys its counts by pattern id: what the lower ranks add to pattern p's count
        // is the sum over every pattern that carries p's id
        unsigned long long v = prior_in ? prior_in[p] : 0ull;
        const int slot = P.pat_slot[p];
        if (slot >= 0)
            for (uint32_t j = P.slot_begin[slot]; j < P.slot_begin[slot + 1]; ++j) {
                const uint32_t q = P.slot_pats[j];
                for (uint32_t r = 0; r < rank; ++r) v += all[r * stride + q];
            }
        prior_out[p] = v;
    }
    if (p == 0) {
        unsigned long long f = 0;
        for (uint32_t r = 0; r < world; ++r) f |= all[r * stride + n_pat];
        *any_flag = f;
    }
}
// Line-range shards with the exchange kept on the device. Every rank's summary, all-gathered into `all`, is
// [n_pat core-line event counts, status word, core lines, n_pat, chain tables (k_range_tables)]; this kernel does for
// one rank what lp_range_score does on the host: the counts seen before the shard's lines (window prior + lower
// shards, per pattern id), the shard's place in the whole log, and per sequence the line where its last-but-one
// event is first reachable when the shards are walked in order (ScoringService.java:296-305, 315-347).
// range_out = {line_base, n_global, error}: error 1 = halo_before exceeds the lines of the lower shards, 2 = the
// summaries come from different pattern sets.
__global__ void k_range_header(const Counters* ctr, uint32_t n_pat, uint32_t halo_before, uint32_t halo_after, unsigned long long* __restrict__ dst) {
    if (blockIdx.x || threadIdx.x) return;
    const unsigned long long lines = ctr->total_nl + 1ull, halos = (unsigned long long)halo_before + halo_after;
    dst[0] = lines > halos ? lines - halos : 0ull;
    dst[1] = n_pat;
}
__global__ void k_range_compose(const unsigned long long* __restrict__ all, uint32_t n_pat, uint32_t n_tab, uint32_t rank, uint32_t world,
                                const unsigned long long* __restrict__ prior_in, uint32_t halo_before, uint32_t n_seq,
                                const uint32_t* __restrict__ tab_off, DevPlan P, unsigned long long* __restrict__ prior_out,
                                unsigned long long* __restrict__ any_flag, long long* __restrict__ range_out, long long* __restrict__ seq_T) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)n_pat + 3 + n_tab;
    if (t < n_pat) {
        unsigned long long v = prior_in ? prior_in[t] : 0ull;
        const int slot = P.pat_slot[t];
        if (slot >= 0)
            for (uint32_t j = P.slot_begin[slot]; j < P.slot_begin[slot + 1]; ++j) {
                const uint32_t q = P.slot_pats[j];
                for (uint32_t r = 0; r < rank; ++r) v += all[r * stride + q];
            }
        prior_out[t] = v;
    }
    if (t < n_seq) {
        const uint32_t k = P.seq_ev_begin[t + 1] - P.seq_ev_begin[t];
        long long T = -1;                                   // fewer than two events: the condition holds for every line
        if (k >= 2) {
            T = 0x7FFFFFFFFFFFFFFFll;                       // never reached
            uint32_t state = 0;
            long long core_base = 0;
            for (uint32_t r = 0; r < world && state + 1 < k; ++r) {
                const unsigned long long* s = all + r * stride;
                const unsigned long long e = s[n_pat + 3 + tab_off[t] + state];
                const uint32_t exit_state = (uint32_t)(e >> 32), at = (uint32_t)(e & 0xFFFFFFFFull);
                if (exit_state + 1 >= k && at) { T = core_base + (long long)at - 1; break; }
                state = exit_state;
                core_base += (long long)s[n_pat + 1];
            }
        }
        seq_T[t] = T;
    }
    if (t == 0) {
        unsigned long long f = 0;
        long long base = 0, n_global = 0, err = 0;
        for (uint32_t r = 0; r < world; ++r) {
            const unsigned long long* s = all + r * stride;
            f |= s[n_pat];
            if (s[n_pat + 2] != n_pat) err = 2;
            if (r < rank) base += (long long)s[n_pat + 1];
            n_global += (long long)s[n_pat + 1];
        }
        if (!err && base < (long long)halo_before) err = 1;
        *any_flag = f;
        range_out[0] = base - (long long)halo_before; range_out[1] = n_global; range_out[2] = err;
    }
}
// Status word of a scan for the other ranks: non-zero exactly when check_estimates() on this rank will ask for a re-run.
struct EstCaps { unsigned long long lines, cand, tasks, vm_tasks, recs, items, events; };
// `counts` are this rank's per-pattern event counts: their sum is the number of events k_emit_events
// produces (same records, same line range), so the events queue is judged even when this runs before it.
__global__ void __launch_bounds__(128) k_publish_overflow(const Counters* ctr, EstCaps cap, const unsigned long long* __restrict__ counts,
                                                          uint32_t n_pat, unsigned long long* __restrict__ dst) {
    __shared__ unsigned long long part[128];
    unsigned long long sum = 0;
    for (uint32_t p = threadIdx.x; p < n_pat; p += blockDim.x) sum += counts[p];
    part[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 64; o; o >>= 1) { if ((int)threadIdx.x < o) part[threadIdx.x] += part[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x != 0) return;
    unsigned long long st = ctr->overflow;
    const unsigned int mi = max(max(ctr->n_items[0], ctr->n_items[1]), max(ctr->n_items[2], ctr->n_items[3]));
    if ((unsigned long long)ctr->n_lines_total + 1 > cap.lines || ctr->n_cand > cap.cand || ctr->n_tasks > cap.tasks ||
        ctr->n_vm_tasks > cap.vm_tasks || ctr->n_recs > cap.recs || mi > cap.items ||
        max((unsigned long long)ctr->n_events, part[0]) > cap.events)
        st |= 1ull << 32;
    *dst = st;
}

// Copies the small read-back items behind the event columns of the output arena (one block).
__global__ void k_pack_tail(const Counters* ctr, const unsigned long long* __restrict__ freq_delta, uint32_t n_pat,
                            const uint32_t* __restrict__ n_lines, uint32_t n_logs, const unsigned long long* __restrict__ log_event_begin,
                            const unsigned long long* flag, const long long* range_dev, Counters* o_ctr, unsigned long long* __restrict__ o_delta,
                            uint32_t* __restrict__ o_nl, unsigned long long* __restrict__ o_leb, unsigned long long* o_flag) {
    if (threadIdx.x == 0) {     // o_flag: status word, then {line_base, n_global, error} of a line-range shard
        *o_ctr = *ctr; o_flag[0] = flag ? *flag : 0ull;
        for (int i = 0; i < 3; ++i) o_flag[1 + i] = range_dev ? (unsigned long long)range_dev[i] : 0ull;
    }
    for (uint32_t p = threadIdx.x; p < n_pat; p += blockDim.x) o_delta[p] = freq_delta[p];
    for (uint32_t k = threadIdx.x; k < n_logs; k += blockDim.x) o_nl[k] = n_lines[k];
    // a single log: its events are all of them (k_log_event_begin is not launched)
    for (uint32_t k = threadIdx.x; k <= n_logs; k += blockDim.x) o_leb[k] = n_logs == 1 ? (k ? (unsigned long long)ctr->n_events : 0ull) : log_event_begin[k];
}
