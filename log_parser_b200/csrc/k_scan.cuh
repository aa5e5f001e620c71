// String.split / prefilter kernels: helpers, k_filter_count, k_scan_tiles, k_lines, k_log_meta.
// Part of engine.cu's translation unit (included inside its anonymous namespace); not a stand-alone header.

__device__ __forceinline__ uint32_t zero_byte_flags(uint32_t x) {
    // exact per-byte "== 0" detector: bit 7 of each byte set iff that byte of x is zero
    uint32_t t = ((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x;
    return ~(t | 0x7F7F7F7Fu);
}
__device__ __forceinline__ uint32_t nl_mask4(uint32_t w) {
    uint32_t m = (zero_byte_flags(w ^ 0x0A0A0A0Au) >> 7);
    return ((m * 0x00204081u) >> 21) & 0xFu;
}
__device__ __forceinline__ uint32_t nl_mask16(const uint4& v) {
    return nl_mask4(v.x) | (nl_mask4(v.y) << 4) | (nl_mask4(v.z) << 8) | (nl_mask4(v.w) << 12);
}

// 16 bytes at `pos`; bytes at or beyond `len` read as zero. `text` is 16-byte aligned. The
// byte-wise tail is kept out of line so that it is not if-converted into every caller's hot loop.
__device__ __noinline__ uint4 load16_tail(const uint8_t* __restrict__ text, uint64_t pos, uint64_t len) {
    uint32_t w[4] = {0, 0, 0, 0};
    for (int i = 0; i < 16; ++i)
        if (pos + i < len) w[i >> 2] |= (uint32_t)text[pos + i] << ((i & 3) * 8);
    return make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ uint4 load16(const uint8_t* __restrict__ text, uint64_t pos, uint64_t len) {
    if (pos + 16 <= len) return __ldg(reinterpret_cast<const uint4*>(text + pos));
    return load16_tail(text, pos, len);
}
__device__ __forceinline__ uint4 load16_any(const uint8_t* __restrict__ text, uint64_t pos, uint64_t len) {
    if (pos + 16 <= len) return __ldg(reinterpret_cast<const uint4*>(text + pos));
    if (pos >= len) return make_uint4(0, 0, 0, 0);
    return load16(text, pos, len);
}
__device__ __noinline__ unsigned long long load8_tail(const uint8_t* __restrict__ text, uint64_t pos, uint64_t len) {
    unsigned long long w = 0;
    for (int i = 0; i < 8; ++i)
        if (pos + i < len) w |= (unsigned long long)text[pos + i] << (i * 8);
    return w;
}
// 8 bytes at the 8-aligned `pos`; bytes at or beyond `len` read as zero
__device__ __forceinline__ unsigned long long load8_any(const uint8_t* __restrict__ text, uint64_t pos, uint64_t len) {
    if (pos + 8 <= len) { const uint2 v = __ldg(reinterpret_cast<const uint2*>(text + pos)); return ((unsigned long long)v.y << 32) | v.x; }
    return load8_tail(text, pos, len);
}
__device__ __noinline__ uint32_t load4_tail(const uint8_t* __restrict__ text, uint64_t pos, uint64_t len) {
    uint32_t w = 0;
    for (int i = 0; i < 4; ++i)
        if (pos + i < len) w |= (uint32_t)text[pos + i] << (i * 8);
    return w;
}
// 4 bytes at any position p (two aligned loads + funnel shift); bytes at or beyond len read as zero
__device__ __forceinline__ uint32_t load4_any(const uint8_t* __restrict__ text, uint32_t p, uint64_t len) {
    const uint32_t b = p & ~3u;
    if ((uint64_t)b + 8 <= len) {
        const uint32_t lo = __ldg(reinterpret_cast<const uint32_t*>(text + b)), hi = __ldg(reinterpret_cast<const uint32_t*>(text + b + 4));
        return __funnelshift_r(lo, hi, (p & 3u) * 8);
    }
    uint32_t w = 0;
    for (int i = 0; i < 4; ++i) if ((uint64_t)p + i < len) w |= (uint32_t)text[p + i] << (8 * i);
    return w;
}
__device__ __forceinline__ uint32_t load4_safe(const uint8_t* __restrict__ text, uint64_t pos, uint64_t len) {
    if (pos + 4 <= len) return __ldg(reinterpret_cast<const uint32_t*>(text + pos));  // pos is 4-aligned here
    return load4_tail(text, pos, len);
}

// number of elements of a[0..n) that are <= v   (upper bound)
template <class T>
__device__ __forceinline__ uint32_t upper_bound_dev(const T* __restrict__ a, uint32_t n, T v) {
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (a[mid] <= v) lo = mid + 1; else hi = mid;
    }
    return lo;
}
// number of elements of a[0..n) that are < v   (lower bound)
template <class T>
__device__ __forceinline__ uint32_t lower_bound_dev(const T* __restrict__ a, uint32_t n, T v) {
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (a[mid] < v) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// ------------------------------------------------------------------ K1: newline count per tile + gram prefilter
// The streaming kernel of the scan: one pass over the log, 16 bytes per thread and iteration.
// It needs no line structure: newlines are only counted per tile (for the offsets pass that
// follows), and every 4-byte window is looked up in a one-byte-per-slot table of literal-gram hashes
// held in shared memory (first stage). First-stage hits are staged per block; at the end of a tile all
// threads check them against the two-probe Bloom bitmap (second stage) and the survivors go to the
// candidate queue with one global atomic per tile.
constexpr uint32_t CAND_STAGE = 2048;      // candidates staged per block and tile in shared memory
struct FilterArgs {
    const uint8_t* text;
    uint64_t len;
    uint32_t n_tiles;
    uint32_t tile_begin, tile_end;    // this launch's tiles (a log arriving from the host is scanned chunk by chunk)
    uint32_t* tile_counts;
    const uint32_t* bitmap;
    const uint8_t* byte_table;
    uint32_t* cand;                   // candidate byte positions
    uint32_t cand_cap;
    Counters* ctr;
    int use_filter;
};
__global__ void __launch_bounds__(SCAN_THREADS) k_filter_count(FilterArgs a) {
    extern __shared__ uint8_t s_tab[];      // 2^16 one-byte slots: 1 if some literal gram hashes here
    __shared__ uint32_t warp_sum[SCAN_THREADS / 32];
    __shared__ uint32_t s_cand[CAND_STAGE];
    __shared__ uint32_t s_ncand[2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { s_ncand[0] = 0; s_ncand[1] = 0; }
    uint32_t par = 0;
    if (a.use_filter) {
        const uint4* src = reinterpret_cast<const uint4*>(a.byte_table);
        uint4* dst = reinterpret_cast<uint4*>(s_tab);
        for (uint32_t i = threadIdx.x; i < 65536 / 16; i += SCAN_THREADS) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    for (uint32_t tile = a.tile_begin + blockIdx.x; tile < a.tile_end; tile += gridDim.x) {
        const uint64_t tile_base = (uint64_t)tile * TILE_BYTES;
        uint32_t cnt = 0;
        const uint64_t pos0 = tile_base + (uint64_t)threadIdx.x * 16;
        uint4 vnext = pos0 < a.len ? load16(a.text, pos0, a.len) : make_uint4(0, 0, 0, 0);
        // the 4 bytes after a warp's 512-byte span (needed by lane 31's last windows) are prefetched the same way
        uint32_t tnext = (lane == 31 && a.use_filter && pos0 + 16 < a.len) ? load4_safe(a.text, pos0 + 16, a.len) : 0u;
#pragma unroll 1
        for (int it = 0; it < SCAN_ITERS; ++it) {
            const uint64_t pos = tile_base + ((uint64_t)it * SCAN_THREADS + threadIdx.x) * 16;
            const uint4 v = vnext;
            const uint32_t tail = tnext;
            {   // the next iteration's 16 bytes are requested before this iteration's work starts
                const uint64_t pn = pos + (uint64_t)SCAN_THREADS * 16;
                const bool more = it + 1 < SCAN_ITERS && pn < a.len;
                vnext = more ? load16(a.text, pn, a.len) : make_uint4(0, 0, 0, 0);
                tnext = (more && lane == 31 && a.use_filter && pn + 16 < a.len) ? load4_safe(a.text, pn + 16, a.len) : 0u;
            }
            cnt += __popc(zero_byte_flags(v.x ^ 0x0A0A0A0Au)) + __popc(zero_byte_flags(v.y ^ 0x0A0A0A0Au)) +
                   __popc(zero_byte_flags(v.z ^ 0x0A0A0A0Au)) + __popc(zero_byte_flags(v.w ^ 0x0A0A0A0Au));
            if (!a.use_filter) continue;
            uint32_t nxt = __shfl_down_sync(0xffffffffu, v.x, 1);
            if (lane == 31) nxt = tail;
            const uint32_t w[5] = {v.x | 0x20202020u, v.y | 0x20202020u, v.z | 0x20202020u, v.w | 0x20202020u, nxt | 0x20202020u};
            uint32_t hits = 0;
#pragma unroll
            for (int i = 15; i >= 0; --i) {
                const uint32_t win = (i & 3) ? __funnelshift_r(w[i >> 2], w[(i >> 2) + 1], (i & 3) * 8) : w[i >> 2];
                hits = hits * 2u + s_tab[(win * GRAM_MUL) >> 16];       // 5 instructions per window: SHF, IMAD, SHF, LDS.U8, LEA
            }
            if (pos >= a.len) hits = 0;
            // first-stage hits (about 1-2 % of the positions) go to a per-block staging buffer in shared
            // memory (one shared-memory atomic per warp and iteration); at the end of the tile the whole
            // block checks them against the two-probe Bloom bitmap with every lane busy, and publishes the
            // survivors with ONE global atomic
            const uint32_t n = __popc(hits);
            if (__ballot_sync(0xffffffffu, n != 0)) {
                uint32_t hinc = n;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, hinc, o); if (lane >= o) hinc += t; }
                const uint32_t total = __shfl_sync(0xffffffffu, hinc, 31);
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(&s_ncand[par], total);
                base = __shfl_sync(0xffffffffu, base, 0) + hinc - n;
                uint32_t hm = hits;
                while (hm) {
                    const int i = __ffs(hm) - 1;
                    hm &= hm - 1;
                    if (base < CAND_STAGE) s_cand[base] = (uint32_t)(pos + i);
                    else {      // staging full (a tile dense with hits): straight to the global queue
                        const uint32_t gi = atomicAdd(&a.ctr->n_cand, 1u);
                        if (gi < a.cand_cap) a.cand[gi] = (uint32_t)(pos + i);
                        else atomicOr(&a.ctr->overflow, 1u);
                    }
                    ++base;
                }
            }
        }
        for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0) warp_sum[warp] = cnt;
        __syncthreads();
        const uint32_t staged = min(s_ncand[par], CAND_STAGE);
        if (threadIdx.x == 0) {
            uint32_t sum = 0;
            for (int q = 0; q < SCAN_THREADS / 32; ++q) sum += warp_sum[q];
            a.tile_counts[tile] = sum;
            s_ncand[par ^ 1] = 0;       // the next tile stages behind the other counter: no reset barrier
        }
        // second stage: both Bloom probes for every staged position (CAND_STAGE / SCAN_THREADS entries per
        // thread); each warp appends its survivors to the global queue with one atomic
#pragma unroll
        for (int k = 0; k < (int)(CAND_STAGE / SCAN_THREADS); ++k) {
            const uint32_t q = threadIdx.x + k * SCAN_THREADS;
            bool keep = false;
            uint32_t p = 0;
            if (q < staged) {
                p = s_cand[q];
                const uint32_t win = load4_any(a.text, p, a.len) | 0x20202020u;
                const uint32_t h1 = (win * GRAM_MUL) >> (32 - GRAM_BITS), h2 = (win * GRAM_MUL2) >> (32 - GRAM_BITS);
                keep = ((__ldg(a.bitmap + (h1 >> 5)) >> (h1 & 31u)) & (__ldg(a.bitmap + (h2 >> 5)) >> (h2 & 31u)) & 1u) != 0;
            }
            const unsigned votes = __ballot_sync(0xffffffffu, keep);
            if (votes) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(&a.ctr->n_cand, (uint32_t)__popc(votes));
                base = __shfl_sync(0xffffffffu, base, 0) + __popc(votes & ((1u << lane) - 1u));
                if (keep) {
                    if (base < a.cand_cap) a.cand[base] = p;
                    else atomicOr(&a.ctr->overflow, 1u);
                }
            }
        }
        __syncthreads();        // every staged entry has been read before the next tile overwrites the buffer
        par ^= 1;
    }
}

// exclusive scan of tile counts (one block); also publishes the total
// `span_counts` (k_fscan: 16 newline counts per tile, one per 1 KiB span) replaces tile_counts as the input when given.
__device__ __forceinline__ uint32_t tile_count_in(const uint32_t* __restrict__ tile_counts, const uint16_t* __restrict__ span_counts, uint32_t i) {
    if (!span_counts) return tile_counts[i];
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(span_counts) + 2 * (size_t)i), b = __ldg(reinterpret_cast<const uint4*>(span_counts) + 2 * (size_t)i + 1);
    const uint32_t lo = (a.x & 0xFFFFu) + (a.y & 0xFFFFu) + (a.z & 0xFFFFu) + (a.w & 0xFFFFu) + (b.x & 0xFFFFu) + (b.y & 0xFFFFu) + (b.z & 0xFFFFu) + (b.w & 0xFFFFu);
    const uint32_t hi = (a.x >> 16) + (a.y >> 16) + (a.z >> 16) + (a.w >> 16) + (b.x >> 16) + (b.y >> 16) + (b.z >> 16) + (b.w >> 16);
    return lo + hi;
}
__global__ void __launch_bounds__(1024) k_scan_tiles(uint32_t* __restrict__ tile_counts, const uint16_t* __restrict__ span_counts, uint32_t n_tiles, uint32_t n_logs, Counters* ctr) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry;
    if (n_tiles <= 1024 * 16) {
        // the usual case (texts up to 256 MiB): every thread owns up to 16 consecutive tiles, one block-wide prefix in all.
        // The counts are fetched with coalesced loads first (a thread's own tiles are 32-byte rows 16 rows apart per lane).
        extern __shared__ uint32_t s_cnt[];
        for (uint32_t i = threadIdx.x; i < n_tiles; i += 1024) s_cnt[i] = tile_count_in(tile_counts, span_counts, i);
        __syncthreads();
        const uint32_t per = (n_tiles + 1023) / 1024, first = threadIdx.x * per;
        uint32_t v[16], sum = 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) { v[k] = ((uint32_t)k < per && first + k < n_tiles) ? s_cnt[first + k] : 0u; sum += v[k]; }
        uint32_t inc = sum;
        for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if ((threadIdx.x & 31) >= o) inc += t; }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_tot[threadIdx.x];
            uint32_t wi = w;
            for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, wi, o); if (threadIdx.x >= o) wi += t; }
            warp_tot[threadIdx.x] = wi - w;  // exclusive
            if (threadIdx.x == 31) { ctr->total_nl = wi; ctr->n_lines_total = wi + n_logs; }
        }
        __syncthreads();
        uint32_t run = warp_tot[threadIdx.x >> 5] + inc - sum;
#pragma unroll
        for (int k = 0; k < 16; ++k) if ((uint32_t)k < per && first + k < n_tiles) { tile_counts[first + k] = run; run += v[k]; }
        return;
    }
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_tiles; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint32_t v = i < n_tiles ? tile_count_in(tile_counts, span_counts, i) : 0;
        uint32_t inc = v;
        for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if ((threadIdx.x & 31) >= o) inc += t; }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_tot[threadIdx.x];
            uint32_t wi = w;
            for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, wi, o); if (threadIdx.x >= o) wi += t; }
            warp_tot[threadIdx.x] = wi - w;  // exclusive
        }
        __syncthreads();
        uint32_t excl = carry + warp_tot[threadIdx.x >> 5] + inc - v;
        if (i < n_tiles) tile_counts[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) { ctr->total_nl = carry; ctr->n_lines_total = carry + n_logs; }
}

// ------------------------------------------------------------------ K2: line offsets
// Second pass over the log: every '\n' yields line_end[g] / line_start[g+1] (with the \r?\n rule and
// the log boundaries of a batch), at the global line index given by the scanned tile counts. (The
// kernel can also run the gram filter itself, `use_filter`; the pipeline does that in k_filter_count.)
struct ScanArgs {
    const uint8_t* text;
    uint64_t len;
    uint32_t n_tiles;
    const uint32_t* tile_prefix;      // exclusive newline count before each tile
    const uint64_t* log_off;          // [n_logs+1]
    uint32_t n_logs;
    uint32_t* line_start;             // [G]
    uint32_t* line_end;               // [G]
    uint32_t* first_line;             // [n_logs+1]
    uint32_t line_cap;                // capacity of line_start / line_end (optimistic; overflow -> retry)
    const uint32_t* bitmap;           // GRAM_WORDS words
    unsigned long long* cand;         // candidates: (global line << 32) | byte position
    uint32_t cand_cap;
    Counters* ctr;
    int use_filter;
};

// Line offsets: every '\n' at p ends line g (at p, or p-1 after a '\r') and starts line g+1 at p+1;
// log boundaries of a batch act as line breaks. One thread owns 64 contiguous bytes, so a tile
// needs a single block-wide prefix of the per-thread newline counts.
constexpr uint32_t LN_SPAN = 64;                                   // bytes per thread
constexpr uint32_t LN_TILE_THREADS = TILE_BYTES / LN_SPAN;         // 512 threads cover one 32 KiB tile
__global__ void __launch_bounds__(LN_TILE_THREADS) k_lines(ScanArgs a) {
    __shared__ uint32_t s_warp[LN_TILE_THREADS / 32];
    __shared__ uint32_t s_k0, s_inner, s_events;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
        const uint64_t tile_base = (uint64_t)tile * TILE_BYTES;
        const uint64_t tile_end = tile_base + TILE_BYTES;
        if (threadIdx.x == 0) {
            uint32_t j0 = upper_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, tile_base);
            s_k0 = j0 - 1;
            s_inner = (j0 <= a.n_logs && a.log_off[j0] < tile_end) ? 1u : 0u;
            uint32_t j1 = lower_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, tile_base);
            s_events = (j1 <= a.n_logs && a.log_off[j1] < tile_end) ? 1u : 0u;
        }
        const uint64_t pos = tile_base + (uint64_t)threadIdx.x * LN_SPAN;
        unsigned long long nl = 0;        // bit i: byte pos+i is '\n'
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint64_t cp = pos + 16 * q;
            if (cp < a.len) {
                const uint4 v = load16(a.text, cp, a.len);
                const uint32_t f = zero_byte_flags(v.x ^ 0x0A0A0A0Au) | zero_byte_flags(v.y ^ 0x0A0A0A0Au) |
                                   zero_byte_flags(v.z ^ 0x0A0A0A0Au) | zero_byte_flags(v.w ^ 0x0A0A0A0Au);
                if (f) nl |= (unsigned long long)nl_mask16(v) << (16 * q);
            }
        }
        const uint32_t cnt = __popcll(nl);
        uint32_t inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        uint32_t wbase = 0;
#pragma unroll
        for (int w = 0; w < LN_TILE_THREADS / 32; ++w) { const uint32_t t = s_warp[w]; if (w < warp) wbase += t; }
        const uint32_t excl = a.tile_prefix[tile] + wbase + inc - cnt;     // newlines before this thread's bytes
        const uint32_t k0 = s_k0;
        const bool inner = s_inner != 0, events = s_events != 0;
        unsigned long long m = nl;
        while (m) {
            const int i = __ffsll((long long)m) - 1;
            m &= m - 1;
            const uint64_t p = pos + i;
            uint32_t k = k0;
            if (inner) k = upper_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, p) - 1;
            const uint32_t g = excl + __popcll(nl & ((1ull << i) - 1ull)) + k;
            const bool cr = p > a.log_off[k] && a.text[p - 1] == '\r';
            if (g + 1 < a.line_cap) {
                a.line_end[g] = (uint32_t)(p - (cr ? 1 : 0));
                a.line_start[g + 1] = (uint32_t)(p + 1);
            } else atomicOr(&a.ctr->overflow, 64u);
        }
        if (events && pos <= a.len) {     // start of log j / end of log j-1 inside this thread's bytes
            uint32_t j = lower_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, pos);
            while (j <= a.n_logs && a.log_off[j] < pos + LN_SPAN) {
                const uint64_t b = a.log_off[j];
                const uint32_t nlb = excl + __popcll(nl & ((1ull << (uint32_t)(b - pos)) - 1ull));
                a.first_line[j] = nlb + j;
                if (nlb + j < a.line_cap) {
                    if (j < a.n_logs) { a.line_start[nlb + j] = (uint32_t)b; }
                    if (j > 0) a.line_end[nlb + j - 1] = (uint32_t)b;
                } else atomicOr(&a.ctr->overflow, 64u);
                ++j;
            }
        }
        __syncthreads();
    }
}

// Java's Pattern.split(…, 0) drops trailing empty strings; no separator at all keeps [input]. One block per log: the
// trailing empty lines are looked for 256 at a time from the end (a log may end in a million blank lines; one thread
// stepping back two dependent loads at a time took 0.6 us per line).
constexpr int LM_THREADS = 256;
__global__ void __launch_bounds__(LM_THREADS) k_log_meta(const uint32_t* __restrict__ first_line, const uint32_t* __restrict__ line_start,
                                                         const uint32_t* __restrict__ line_end, uint32_t n_logs, uint32_t line_cap, uint32_t no_trim,
                                                         uint32_t* __restrict__ n_lines) {
    __shared__ uint32_t s_last;      // 1 + index (within the chunk) of the last non-empty line of the chunk, 0 = none
    for (uint32_t k = blockIdx.x; k < n_logs; k += gridDim.x) {
        const uint32_t f = first_line[k];
        uint32_t n = first_line[k + 1] - f;
        if (first_line[k + 1] > line_cap) { if (threadIdx.x == 0) n_lines[k] = 0; continue; }     // line arrays overflowed: the call is retried
        if (n > 1 && !no_trim) {      // a line-range shard is cut from a log whose trailing empty lines are already gone
            while (n > 0) {
                const uint32_t lo = n > LM_THREADS ? n - LM_THREADS : 0u;      // lines [lo, n) of the log
                __syncthreads();
                if (threadIdx.x == 0) s_last = 0;
                __syncthreads();
                const uint32_t g = lo + threadIdx.x;
                if (g < n && line_end[f + g] != line_start[f + g]) atomicMax(&s_last, threadIdx.x + 1);
                __syncthreads();
                if (s_last) { n = lo + s_last; break; }
                n = lo;
            }
        }
        if (threadIdx.x == 0) n_lines[k] = n;
        __syncthreads();
    }
}
