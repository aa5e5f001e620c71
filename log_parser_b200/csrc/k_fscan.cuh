// Fused scan: String.split("\\r?\\n") (AnalysisService.java:53) and the literal-gram prefilter of the
// N x P find() loop (AS:89-95) in ONE pass over the log. Replaces k_filter_count + k_scan_tiles + k_lines.
// Part of engine.cu's translation unit (included inside its anonymous namespace); not a stand-alone header.
//
// Shape: one CTA of 1024 threads per SM; every warp is a pipeline of its own. A warp owns two buffers of one
// 1 KiB span (+ 16 bytes of halo) of the log in shared memory, filled by 1-D TMA bulk copies
// (cp.async.bulk ... mbarrier::complete_tx; the source is always this GPU's own memory) while the warp works on
// the other one, so the bytes in flight do not depend on registers. Spans are dealt round-robin to the warps of
// the grid. After the first-stage table has been loaded there is no barrier between warps at all: the table
// lookups of one warp (integer-ALU bound) overlap the sparse, latency-bound loops of another. (Two earlier
// shapes, 16-KiB tiles shared by 16 warps behind a barrier or behind a last-arriver counter, lost 13-25 % of
// the issue slots to waiting: profiles/r02_fscan_spanlists_* and r02_fscan_lastarriver_*.)
//
// Per span, a lane owns 32 contiguous bytes:
//   phase A  newline mask by SWAR, and for each of its 32 positions one lookup of the 4-byte window in the
//            first-stage gram table. The table is a bit table replicated per lane: row = hash bits, the
//            word of lane L lives in bank L, so the 32 lookups of a warp never conflict (the byte table of
//            k_filter_count lost 55 % of its shared-memory wavefronts to bank conflicts). The wanted bit
//            is rotated to bit 31 and funnel-shifted into the hit mask: 6 instructions per position.
//   phase B  newlines: one warp scan; position | first-byte skip flag | '\r' flag of every newline goes, in
//            order, to the warp span's slot of the tile's newline lists (k_expand turns the lists into
//            line_start / line_end after k_scan_tiles has summed and scanned the spans' counts: the log is not read
//            again for the offsets). First-stage hits are checked against the exact gram keys (shared
//            memory when they fit; windows re-read from the staged span); survivors collect in the warp's
//            list and go to the candidate queue with one global atomic per ~32 of them.
namespace fs {
constexpr int THREADS = 1024;
constexpr uint32_t TILE = 16384, HALO = 16;                          // a tile is the unit of the newline prefix (k_scan_tiles, k_expand)
constexpr uint32_t SPAN = 1024, SPANS = TILE / SPAN;                 // a span is what one warp stages and scans at a time: 32 bytes per lane
constexpr uint32_t SPAN_STRIDE = 1152;                               // span + halo, rounded up to 128
constexpr uint32_t ROWS = 1024;                                      // first-stage table rows (one word per lane and row)
constexpr uint32_t MUL1 = 0x9E3779B1u;
constexpr size_t TABLE_BYTES = (size_t)ROWS * 128;
constexpr uint32_t WLIST = 64;                                       // newline list entries per span
constexpr uint32_t OUT_CAP = 32;                                     // survivors staged per warp (flushed when half full)
constexpr uint32_t HIT_CAP = 64;                                     // first-stage hits of a pair of spans compacted per warp
constexpr uint32_t KEY_BYTES = 8192;                                 // exact gram keys kept in shared memory when 4 << table_bits fits
// per warp: two span buffers, the pair's newline positions (2 x WLIST u16), first-stage hits (u16), survivors, state
constexpr size_t WARP_BYTES = 2 * SPAN_STRIDE + 2 * WLIST * 2 + HIT_CAP * 2 + OUT_CAP * 4 + 32;
constexpr size_t SMEM_BYTES = TABLE_BYTES + KEY_BYTES + (THREADS / 32) * WARP_BYTES;
static_assert(SMEM_BYTES <= 227 * 1024, "k_fscan shared memory");

// row = (H >> 7) & (ROWS - 1), rotation = H & 31; the table bit of a gram is bit (rotation - 1) & 31.
// `mul` is an odd multiplier picked per pattern set (and per first log, see pick_multiplier in engine.cu);
// MUL1 is the default.
template <int V>
__host__ __device__ __forceinline__ uint32_t hash(uint32_t folded, uint32_t mul) {
#ifdef __CUDA_ARCH__
    if (V == 0) return __umulhi(folded, mul);
#else
    if (V == 0) return (uint32_t)(((unsigned long long)folded * mul) >> 32);
#endif
    return (folded * mul) >> 15;
}
inline uint32_t host_hash(int variant, uint32_t folded, uint32_t mul) { return variant == 1 ? hash<1>(folded, mul) : hash<0>(folded, mul); }
// table row of a hash: bits 7..16 (the middle bits of the 64-bit product are the well-mixed ones; its top ten bits
// follow the window's last byte and gave four times the false hits). Variants 0/1 mask them in place (one LOP3
// builds the address); variant 3 isolates them with multiplies (H * 2^15, IMAD.HI by 1024, IMAD by 128 plus the
// lane's offset): the FMA pipe, which has slack, instead of the ALU pipe, which is the bound.
inline uint32_t host_row(uint32_t H) { return (H >> 7) & (ROWS - 1); }

struct WarpState {
    unsigned long long bar[2];      // mbarriers of the warp's two span buffers
    uint32_t ncand;                 // survivors waiting in the warp's list
    uint32_t nhit;                  // first-stage hits of the current pair of spans
};
static_assert(sizeof(WarpState) <= 32, "WarpState must fit its slot");
}  // namespace fs

struct FscanArgs {
    const uint8_t* text;
    uint64_t len;
    uint32_t n_tiles;                 // tiles of the whole text (the last one covers position `len`)
    uint32_t tile_begin, tile_end;    // this launch's tiles (a log arriving from the host is scanned piece by piece)
    uint32_t* tile_counts;            // [n_tiles] out, zeroed before the first launch: newlines per tile, summed from the spans by reduction atomics
    uint16_t* warp_counts;            // [n_tiles][16] out: newlines per 1 KiB warp span; above fs::WLIST the span's list was not written
    uint16_t* nl_lists;               // [n_tiles][16][fs::WLIST] out: the newlines of each warp span, see k_expand
    const uint32_t* bit_table;        // fs::ROWS words (variants 0/1) or the 64 KiB byte table (variant 2)
    const uint32_t* gram_keys;        // 2^table_bits folded grams, 0 = empty
    uint32_t table_bits;
    uint32_t mul;                     // first-stage hash multiplier (fs::hash)
    uint32_t k[8];                    // variant 3: {1, 1024, 128, 1<<25, 1<<26, 1<<27, 1<<28, 1<<15}, passed as arguments so that ptxas keeps the multiplies (FMA pipe) instead of turning them into shifts and adds (ALU pipe, the kernel's bound)
    uint32_t* cand;
    uint32_t cand_cap;
    Counters* ctr;
    int use_filter;
    uint32_t skip1[8];                // first bytes on which every always-on scan group is decided without a hit (k_dfa_scan skips such lines)
};

__device__ __forceinline__ uint32_t fs_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fs_mbar_init(void* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(fs_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fs_mbar_wait(void* bar, uint32_t parity) {
    const uint32_t addr = fs_smem_u32(bar);
    uint32_t ok;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    } while (!ok);
}
// 1-D bulk copy global -> shared of this CTA, completion counted in bytes on `bar`. No multicast; the source is this GPU's memory.
__device__ __forceinline__ void fs_tma_load(void* dst, const void* src, uint32_t bytes, void* bar) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fs_smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(fs_smem_u32(dst)), "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(fs_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fs_mbar_arrive(void* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(fs_smem_u32(bar)) : "memory");
}

// Bit 7 of each byte: set iff that byte of w is '\n' (exact). (y - 0x01..) & ~y flags zero bytes of y, and, through
// the borrow, a byte 0x01 that follows one; bit 0 of such a byte (y << 7) clears the false flag. Bits other
// than 7, 15, 23, 31 of the result are garbage: the caller masks. The subtraction and the left shift are
// written as multiply-adds so that they issue on the FMA pipe; the scan is bound by the integer ALU pipe.
__device__ __forceinline__ uint32_t fs_nl_flags_raw(uint32_t w) {
    const uint32_t y = w ^ 0x0A0A0A0Au;
    uint32_t t, u;
    asm("mad.lo.u32 %0, %1, 1, 0xFEFEFEFF;" : "=r"(t) : "r"(y));
    asm("mul.lo.u32 %0, %1, 128;" : "=r"(u) : "r"(y));
    return t & ~y & ~u;
}
// newline mask of 16 bytes in "spread" order: bit (q + 8 j) <-> byte j of word q
__device__ __forceinline__ uint32_t fs_nl_spread(const uint4& v) {
    return ((fs_nl_flags_raw(v.x) >> 7) & 0x01010101u) | ((fs_nl_flags_raw(v.y) >> 6) & 0x02020202u) |
           ((fs_nl_flags_raw(v.z) >> 5) & 0x04040404u) | ((fs_nl_flags_raw(v.w) >> 4) & 0x08080808u);
}
// the same with the subtraction and the right shifts as multiplies by kernel arguments (k[0] = 1, k[3..6] = 2^25..2^28)
__device__ __forceinline__ uint32_t fs_nl_flags_fma(uint32_t w, uint32_t one) {
    const uint32_t y = w ^ 0x0A0A0A0Au;
    uint32_t t, u;
    asm("mad.lo.u32 %0, %1, %2, 0xFEFEFEFF;" : "=r"(t) : "r"(y), "r"(one));
    asm("mul.lo.u32 %0, %1, 128;" : "=r"(u) : "r"(y));
    return t & ~y & ~u;
}
__device__ __forceinline__ uint32_t fs_nl_spread_fma(const uint4& v, const uint32_t (&k)[8]) {
    return (__umulhi(fs_nl_flags_fma(v.x, k[0]), k[3]) & 0x01010101u) | (__umulhi(fs_nl_flags_fma(v.y, k[0]), k[4]) & 0x02020202u) |
           (__umulhi(fs_nl_flags_fma(v.z, k[0]), k[5]) & 0x04040404u) | (__umulhi(fs_nl_flags_fma(v.w, k[0]), k[6]) & 0x08080808u);
}
// With two 16-byte chunks packed as lo | hi << 4, bit b stands for byte 4 (b & 7) + (b >> 3) of the 32 bytes.
// Mask of the packed bits whose byte position is below x (0..32).
__device__ __forceinline__ uint32_t fs_below(uint32_t x) {
    const uint32_t n = x >> 2;
    return (((1u << n) - 1u) * 0x01010101u) | ((0x01010101u << n) & ((1u << (8u * (x & 3u))) - 1u));
}

// first-stage lookups of the 16 windows starting in one 16-byte chunk; w4 = the word that follows the chunk
template <int V>
__device__ __forceinline__ uint32_t fs_probe16(const uint4& v, uint32_t w4, const uint8_t* __restrict__ tab, uint32_t lane4, uint32_t mul, const uint32_t (&k)[8]) {
    const uint32_t w[5] = {v.x | 0x20202020u, v.y | 0x20202020u, v.z | 0x20202020u, v.w | 0x20202020u, w4 | 0x20202020u};
    uint32_t acc = 0;
#pragma unroll
    for (int i = 15; i >= 0; --i) {
        const uint32_t win = (i & 3) ? __funnelshift_r(w[i >> 2], w[(i >> 2) + 1], (i & 3) * 8) : w[i >> 2];
        if (V == 2) {
            acc = acc * 2u + tab[(win * GRAM_MUL) >> 16];
        } else if (V == 3) {
            const uint32_t H = __umulhi(win, mul);
            const uint32_t word = *reinterpret_cast<const uint32_t*>(tab + (__umulhi(H * k[7], k[1]) * k[2] + lane4));
            acc = __funnelshift_l(__funnelshift_r(word, word, H), acc, 1);
        } else {
            const uint32_t H = fs::hash<V>(win, mul);
            const uint32_t word = *reinterpret_cast<const uint32_t*>(tab + ((H & ((fs::ROWS - 1u) << 7)) | lane4));
            acc = __funnelshift_l(__funnelshift_r(word, word, H), acc, 1);      // table bit -> bit 31 -> shifted into the mask
        }
    }
    return acc;
}

template <int V>
__global__ void __launch_bounds__(fs::THREADS, 1) k_fscan(FscanArgs a) {
    extern __shared__ __align__(128) uint8_t fs_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* const s_keys = reinterpret_cast<uint32_t*>(fs_smem + fs::TABLE_BYTES);          // exact gram keys (when they fit)
    uint8_t* const wbase = fs_smem + fs::TABLE_BYTES + fs::KEY_BYTES + (size_t)warp * fs::WARP_BYTES;
    uint16_t* const w_nl = reinterpret_cast<uint16_t*>(wbase + 2 * fs::SPAN_STRIDE);          // [2][WLIST] newline positions of the pair
    uint16_t* const w_hit = w_nl + 2 * fs::WLIST;                                             // [HIT_CAP] buffer << 10 | position
    uint32_t* const w_out = reinterpret_cast<uint32_t*>(w_hit + fs::HIT_CAP);                 // [OUT_CAP] this warp's survivors
    fs::WarpState* const ws = reinterpret_cast<fs::WarpState*>(w_out + fs::OUT_CAP);
    const uint32_t lane4 = (uint32_t)lane * 4u;
    const uint32_t full = 0xffffffffu;
    const bool keys_in_smem = (4u << a.table_bits) <= fs::KEY_BYTES;
    const uint32_t tmask = (1u << a.table_bits) - 1u, kshift = 32 - a.table_bits;

    if (lane == 0) {
        fs_mbar_init(&ws->bar[0], 1);
        fs_mbar_init(&ws->bar[1], 1);
        ws->ncand = 0;
        ws->nhit = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    // Spans of this launch, dealt round-robin to all warps of the grid: warp gw owns spans gw, gw + W, gw + 2W, ...
    // (131 552 spans over 4 736 warps on C2: 27 or 28 each, so a static deal is balanced to 1 %), and works on
    // two of them at a time, one per buffer.
    const uint32_t span_begin = a.tile_begin * fs::SPANS, span_end = a.tile_end * fs::SPANS;
    const uint32_t W = gridDim.x * (fs::THREADS / 32);
    uint32_t span = span_begin + blockIdx.x * (fs::THREADS / 32) + warp;
    // Fill buffer b with span sp: one bulk copy of the span and its halo when they lie inside the text, else (the
    // last spans of the text) a bounds-checked copy by the warp; bytes at or beyond `len` read as zero.
    auto fill = [&](uint32_t sp, int b) {
        if (sp >= span_end) return;
        uint8_t* const dst = wbase + b * fs::SPAN_STRIDE;
        const uint64_t pos = (uint64_t)sp * fs::SPAN;
        if (pos + fs::SPAN + fs::HALO <= a.len) {
            if (lane == 0) fs_tma_load(dst, a.text + pos, fs::SPAN + fs::HALO, &ws->bar[b]);
        } else {
            for (uint32_t i = lane; i < (fs::SPAN + fs::HALO) / 16; i += 32) reinterpret_cast<uint4*>(dst)[i] = load16_any(a.text, pos + 16ull * i, a.len);
            __syncwarp();
            if (lane == 0) fs_mbar_arrive(&ws->bar[b]);
        }
    };
    fill(span, 0);
    fill(span + W, 1);      // in flight while the table is loaded
    if (a.use_filter) {
        if (V == 2) {
            const uint4* src = reinterpret_cast<const uint4*>(a.bit_table);
            uint4* dst = reinterpret_cast<uint4*>(fs_smem);
            for (uint32_t i = threadIdx.x; i < 65536 / 16; i += fs::THREADS) dst[i] = __ldg(src + i);
        } else {
            uint32_t* dst = reinterpret_cast<uint32_t*>(fs_smem);
            for (uint32_t i = threadIdx.x; i < fs::ROWS * 32; i += fs::THREADS) dst[i] = __ldg(a.bit_table + (i >> 5));
        }
        if (keys_in_smem)
            for (uint32_t i = threadIdx.x; i <= tmask; i += fs::THREADS) s_keys[i] = __ldg(a.gram_keys + i);
    }
    __syncthreads();        // the only block-wide barrier: from here on every warp is a pipeline of its own

    // survivors of this warp -> candidate queue, one global atomic
    auto flush = [&]() {
        const uint32_t nc = min(ws->ncand, fs::OUT_CAP);
        if (nc) {
            uint32_t gb = 0;
            if (lane == 0) gb = atomicAdd(&a.ctr->n_cand, nc);
            gb = __shfl_sync(full, gb, 0);
            if ((uint32_t)lane < nc) {
                if (gb + lane < a.cand_cap) a.cand[gb + lane] = w_out[lane];
                else atomicOr(&a.ctr->overflow, 1u);
            }
        }
        __syncwarp();
        if (lane == 0) ws->ncand = 0;
        __syncwarp();
    };
    // one first-stage hit (buffer << 10 | position in the span) against the exact gram keys; the window is re-read
    // from the staged span
    auto check_hit = [&](uint32_t code, uint32_t span0) {
        const uint32_t bsel = code >> 10, pt = code & 1023u, al = pt & ~3u;
        const uint8_t* const st = wbase + bsel * fs::SPAN_STRIDE;
        const uint32_t wlo = *reinterpret_cast<const uint32_t*>(st + al), whi = *reinterpret_cast<const uint32_t*>(st + al + 4);
        const uint32_t f = __funnelshift_r(wlo, whi, (pt & 3u) * 8u) | 0x20202020u;
        uint32_t slot = (f * GRAM_MUL2) >> kshift;
        uint32_t key;
        if (keys_in_smem) { while ((key = s_keys[slot]) != 0u && key != f) slot = (slot + 1) & tmask; }
        else { while ((key = __ldg(a.gram_keys + slot)) != 0u && key != f) slot = (slot + 1) & tmask; }
        if (key == f) {
            const uint32_t pos = (span0 + bsel * W) * fs::SPAN + pt;          // texts are below 4 GiB
            const uint32_t ob = atomicAdd(&ws->ncand, 1u);
            if (ob < fs::OUT_CAP) w_out[ob] = pos;
            else {      // spans dense with gram occurrences: straight to the global queue
                const uint32_t gi = atomicAdd(&a.ctr->n_cand, 1u);
                if (gi < a.cand_cap) a.cand[gi] = pos;
                else atomicOr(&a.ctr->overflow, 1u);
            }
        }
    };

    for (uint32_t n = 0; span < span_end; ++n, span += 2 * W) {
        const bool two = span + W < span_end;      // the last round of a warp may hold one span only
        uint32_t P[2], hm[2];
        // ---- phase A on both buffers: 32 contiguous bytes per lane and span
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            P[b] = 0; hm[b] = 0;
            if (b == 1 && !two) break;
            const uint8_t* const stage = wbase + b * fs::SPAN_STRIDE;
            const uint64_t base = (uint64_t)(span + b * W) * fs::SPAN;
            fs_mbar_wait(&ws->bar[b], n & 1u);
            const uint4 v0 = *reinterpret_cast<const uint4*>(stage + 32 * lane);
            const uint4 v1 = *reinterpret_cast<const uint4*>(stage + 32 * lane + 16);
            P[b] = V == 3 ? fs_nl_spread_fma(v0, a.k) | (fs_nl_spread_fma(v1, a.k) << 4) : fs_nl_spread(v0) | (fs_nl_spread(v1) << 4);
            if (a.use_filter) {
                uint32_t nxt = __shfl_down_sync(full, v0.x, 1);
                if (lane == 31) nxt = *reinterpret_cast<const uint32_t*>(stage + fs::SPAN);
                hm[b] = fs_probe16<V>(v0, v1.x, fs_smem, lane4, a.mul, a.k) | (fs_probe16<V>(v1, nxt, fs_smem, lane4, a.mul, a.k) << 16);
                if (base + fs::SPAN > a.len) {      // the text ends in this span: no windows at or beyond `len`
                    const uint64_t p0 = base + 32ull * lane;
                    hm[b] = p0 >= a.len ? 0u : (a.len - p0 < 32 ? hm[b] & ((1u << (uint32_t)(a.len - p0)) - 1u) : hm[b]);
                }
            }
        }

        // ---- phase B, newlines. One warp scan ranks the newlines of both spans (counts packed 16 + 16 bits); every lane
        // drops the positions of its own into the warp's lists at their ranks; then the lists are finished by full
        // warps: position in the tile | skip flag << 14 | ('\r' before it) << 15, in position order, into the span's slot
        // of the global newline lists (coalesced). No block-wide rank is needed: k_expand adds the counts of the
        // spans in front. A span with more newlines than its slot holds writes only its count; k_expand then re-reads
        // that tile from the text.
        {
            const uint32_t cn = __popc(P[0]) | (__popc(P[1]) << 16);
            uint32_t inc = cn;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(full, inc, o); if (lane >= o) inc += y; }
            const uint32_t tot = __shfl_sync(full, inc, 31);
            const uint32_t t0 = tot & 0xFFFFu, t1 = tot >> 16;
            const uint32_t n0 = t0 <= fs::WLIST ? t0 : 0u, n1 = t1 <= fs::WLIST ? t1 : 0u;
            const uint32_t excl = inc - cn;
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                uint32_t m = (b ? n1 : n0) ? P[b] : 0u;
                const uint32_t ex = b ? excl >> 16 : excl & 0xFFFFu;
                while (m) {
                    const uint32_t bi = __ffs(m) - 1;
                    m &= m - 1;
                    const uint32_t nq = bi & 7u, j = bi >> 3;
                    const uint32_t lt = (((1u << nq) - 1u) * 0x01010101u) | ((0x01010101u << nq) & ((1u << (8u * j)) - 1u));
                    w_nl[b * fs::WLIST + ex + __popc(P[b] & lt)] = (uint16_t)(32 * lane + 4 * nq + j);
                }
            }
            __syncwarp();
            for (uint32_t e = lane; e < n0 + n1; e += 32) {
                const uint32_t bsel = e >= n0 ? 1u : 0u, idx = e - (bsel ? n0 : 0u);
                const uint32_t sp = span + bsel * W;
                const uint8_t* const st = wbase + bsel * fs::SPAN_STRIDE;
                const uint32_t pt = w_nl[bsel * fs::WLIST + idx];
                uint32_t prev;
                if (pt) prev = st[pt - 1];
                else prev = sp ? a.text[(uint64_t)sp * fs::SPAN - 1] : 0u;      // (two branches: a select would make both loads generic)
                const uint32_t nb = st[pt + 1];             // first byte of the line that starts here (the halo covers pt + 1 == SPAN)
                const uint32_t skip = (nb != '\n' && nb != 0u) ? ((a.skip1[nb >> 5] >> (nb & 31u)) & 1u) : 0u;
                a.nl_lists[(size_t)sp * fs::WLIST + idx] = (uint16_t)(((sp & (fs::SPANS - 1)) * fs::SPAN + pt) | (skip << 14) | (prev == '\r' ? 0x8000u : 0u));
            }
            if (lane == 0) {
                a.warp_counts[span] = (uint16_t)t0;
                if (t0) atomicAdd(&a.tile_counts[span / fs::SPANS], t0);
                if (two) {
                    a.warp_counts[span + W] = (uint16_t)t1;
                    if (t1) atomicAdd(&a.tile_counts[(span + W) / fs::SPANS], t1);
                }
            }
        }
        // ---- phase B, first-stage hits: compacted into the warp's list, then checked by full warps (the hits of a
        // lane are few and uneven: a loop per lane ran with 2-3 of 32 lanes for most of its trips)
        if (a.use_filter) {
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                uint32_t m = hm[b];
                while (m) {
                    const uint32_t i = __ffs(m) - 1;
                    m &= m - 1;
                    const uint32_t code = ((uint32_t)b << 10) | (32u * lane + i);
                    const uint32_t at = atomicAdd(&ws->nhit, 1u);
                    if (at < fs::HIT_CAP) w_hit[at] = (uint16_t)code;
                    else check_hit(code, span);
                }
            }
            __syncwarp();
            const uint32_t nh = min(ws->nhit, fs::HIT_CAP);
            for (uint32_t e = lane; e < nh; e += 32) check_hit(w_hit[e], span);
            __syncwarp();
            if (lane == 0) ws->nhit = 0;
        }
        __syncwarp();       // every lane is done with the buffers
        fill(span + 2 * W, 0);
        fill(span + 3 * W, 1);
        if (ws->ncand >= fs::OUT_CAP / 2) flush();
    }
    __syncwarp();
    flush();
}

// ------------------------------------------------------------------ line offsets from the newline lists
// After k_scan_tiles has turned the tiles' newline counts into exclusive prefixes: every '\n' at p ends line g
// (at p, or p-1 after a '\r') and starts line g+1 at p+1; the log boundaries of a batch act as line breaks.
// One block per tile. A tile whose 16 warp spans all fit their list slots is expanded from the lists (no access
// to the text); a denser tile is re-read from the text, 64 bytes per thread, with one block-wide prefix.
struct ExpandArgs {
    const uint8_t* text;
    uint64_t len;
    uint32_t n_tiles;
    const uint32_t* tile_prefix;      // exclusive newline count before each tile (tile_counts after k_scan_tiles)
    const uint16_t* warp_counts;      // [n_tiles][16]
    const uint16_t* nl_lists;         // [n_tiles][16][fs::WLIST]
    const uint64_t* log_off;          // [n_logs+1]
    uint32_t n_logs;
    uint32_t* line_start;
    uint32_t* line_end;
    uint8_t* line_skip;               // [line_cap] out (may be NULL): 1 = the always-on scan need not walk this line
    uint32_t* scan_list;              // [line_cap] out (with line_skip): the lines it has to walk, compacted (any order)
    uint32_t skip1[8];
    uint32_t* first_line;             // [n_logs+1]
    uint32_t line_cap;
    Counters* ctr;
};
constexpr int EX_THREADS = 256;
constexpr int EX_SCAN_STAGE = 1536;      // a tile holds at most 16 x 64 listed newlines; the rest is room for the first lines of logs
// Appends line g to the list of lines the always-on scan walks, one global atomic per warp. Called by every lane that is
// active at the call site (`want` false: nothing to add). A warp of k_dfa_scan that takes 32 consecutive lines nearly
// always holds a stack frame or two, so skipping inside it saved memory traffic but no time; with the list only the
// lines that need it are dealt out.
__device__ __forceinline__ void ex_append_scan(const ExpandArgs& a, uint32_t g, bool want) {
    const unsigned act = __activemask();
    const unsigned votes = __ballot_sync(act, want);
    if (!votes) return;
    const int lane = threadIdx.x & 31, leader = __ffs(votes) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(&a.ctr->n_scan, (unsigned)__popc(votes));
    base = __shfl_sync(act, base, leader);
    if (want) a.scan_list[base + __popc(votes & ((1u << lane) - 1u))] = g;
}
__global__ void __launch_bounds__(EX_THREADS) k_expand(ExpandArgs a) {
    __shared__ uint32_t s_warp[EX_THREADS / 32];
    __shared__ uint32_t s_pre[17];      // exclusive prefix of the 16 span counts; [16] = the tile's total
    __shared__ uint32_t s_dense;
    __shared__ uint32_t s_scan[EX_SCAN_STAGE], s_nscan, s_scan_base;      // the tile's lines for the always-on scan, published with one atomic
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_nscan = 0;
    auto stage_scan = [&](uint32_t g) {
        const uint32_t at = atomicAdd(&s_nscan, 1u);
        if (at < EX_SCAN_STAGE) s_scan[at] = g;
        else a.scan_list[atomicAdd(&a.ctr->n_scan, 1u)] = g;       // more than the stage holds (a tile of tiny logs): straight to the list
    };
    for (uint32_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
        const uint64_t base = (uint64_t)tile * fs::TILE, tile_end = base + fs::TILE;
        const uint32_t prefix = a.tile_prefix[tile];
        __syncthreads();        // the previous tile's shared values have been read
        if (threadIdx.x < 32) {
            const uint32_t c = lane < 16 ? a.warp_counts[(size_t)tile * 16 + lane] : 0u;
            uint32_t inc = c;
#pragma unroll
            for (int o = 1; o < 16; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
            if (lane < 16) s_pre[lane] = inc - c;
            if (lane == 15) s_pre[16] = inc;
            const unsigned over = __ballot_sync(0xffffffffu, c > fs::WLIST);
            if (lane == 0) s_dense = over != 0;
        }
        __syncthreads();
        const uint32_t count = s_pre[16];
        // where the logs begin relative to this tile (a single log: no loads)
        uint32_t k0 = 0;
        bool inner = false, events = false;
        if (a.n_logs == 1) {
            k0 = base == a.len ? 1u : 0u;
            events = base == 0 || a.len < tile_end;
        } else {
            const uint32_t j0 = upper_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, base);
            k0 = j0 - 1;
            inner = j0 <= a.n_logs && a.log_off[j0] < tile_end;
            const uint32_t j1 = lower_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, base);
            events = j1 <= a.n_logs && a.log_off[j1] < tile_end;
        }
        if (!s_dense) {
            const uint16_t* __restrict__ lists = a.nl_lists + (size_t)tile * 16 * fs::WLIST;
            for (uint32_t e = threadIdx.x; e < count; e += EX_THREADS) {
                uint32_t w = 0;         // the span that holds the tile's e-th newline
#pragma unroll
                for (int q = 8; q; q >>= 1) if (s_pre[w + q] <= e) w += q;
                const uint32_t ent = lists[w * fs::WLIST + (e - s_pre[w])];
                const uint64_t p = base + (ent & 0x3FFFu);
                uint32_t k = k0;
                if (inner) k = upper_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, p) - 1;
                const uint32_t g = prefix + e + k;
                const uint64_t log_begin = a.n_logs == 1 ? 0ull : a.log_off[k];
                const bool cr = (ent >> 15) != 0 && p > log_begin;
                if (g + 1 < a.line_cap) {
                    a.line_end[g] = (uint32_t)(p - (cr ? 1 : 0));
                    a.line_start[g + 1] = (uint32_t)(p + 1);
                    if (a.line_skip) a.line_skip[g + 1] = (uint8_t)((ent >> 14) & 1u);
                } else atomicOr(&a.ctr->overflow, 64u);
                if (a.scan_list && g + 1 < a.line_cap && !((ent >> 14) & 1u)) stage_scan(g + 1);
            }
            if (events) {       // start of log j / end of log j-1 inside this tile: newlines before it = spans in front + a binary search in its span
                uint32_t j = lower_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, base) + threadIdx.x;
                for (; j <= a.n_logs && a.log_off[j] < tile_end; j += EX_THREADS) {
                    const uint64_t b = a.log_off[j];
                    const uint32_t want = (uint32_t)(b - base), w = want >> 10;
                    const uint16_t* __restrict__ list = lists + w * fs::WLIST;
                    uint32_t lo = 0, hi = s_pre[w + 1] - s_pre[w];
                    while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if ((list[mid] & 0x3FFFu) < want) lo = mid + 1; else hi = mid; }
                    const uint32_t nlb = prefix + s_pre[w] + lo;
                    a.first_line[j] = nlb + j;
                    if (nlb + j < a.line_cap) {
                        if (j < a.n_logs) { a.line_start[nlb + j] = (uint32_t)b; if (a.line_skip) a.line_skip[nlb + j] = 0; }   // first line of a log: always walked
                        if (j > 0) a.line_end[nlb + j - 1] = (uint32_t)b;
                    } else atomicOr(&a.ctr->overflow, 64u);
                    if (a.scan_list && nlb + j < a.line_cap && j < a.n_logs) stage_scan(nlb + j);
                }
            }
            if (a.scan_list) {
                __syncthreads();
                const uint32_t staged = min(s_nscan, (uint32_t)EX_SCAN_STAGE);
                if (threadIdx.x == 0 && staged) s_scan_base = atomicAdd(&a.ctr->n_scan, staged);
                __syncthreads();
                for (uint32_t q = threadIdx.x; q < staged; q += EX_THREADS) a.scan_list[s_scan_base + q] = s_scan[q];
                __syncthreads();
                if (threadIdx.x == 0) s_nscan = 0;
            }
            continue;
        }
        // dense tile: from the text
        const uint64_t pos = base + (uint64_t)threadIdx.x * 64;
        unsigned long long nl = 0;        // bit i: byte pos+i is '\n'
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint64_t cp = pos + 16 * q;
            if (cp < a.len) nl |= (unsigned long long)nl_mask16(load16(a.text, cp, a.len)) << (16 * q);
        }
        const uint32_t cnt = __popcll(nl);
        uint32_t inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        uint32_t wbase = 0;
#pragma unroll
        for (int w = 0; w < EX_THREADS / 32; ++w) { const uint32_t y = s_warp[w]; if (w < warp) wbase += y; }
        const uint32_t excl = prefix + wbase + inc - cnt;     // newlines before this thread's bytes
        unsigned long long m = nl;
        while (m) {
            const int i = __ffsll((long long)m) - 1;
            m &= m - 1;
            const uint64_t p = pos + i;
            uint32_t k = k0;
            if (inner) k = upper_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, p) - 1;
            const uint32_t g = excl + __popcll(nl & ((1ull << i) - 1ull)) + k;
            const bool cr = p > a.log_off[k] && a.text[p - 1] == '\r';
            uint32_t sk = 0;
            if (g + 1 < a.line_cap) {
                a.line_end[g] = (uint32_t)(p - (cr ? 1 : 0));
                a.line_start[g + 1] = (uint32_t)(p + 1);
                if (a.line_skip) {
                    const uint32_t nb = p + 1 < a.len ? a.text[p + 1] : 0u;
                    sk = (nb != '\n' && nb != 0u) ? ((a.skip1[nb >> 5] >> (nb & 31u)) & 1u) : 0u;
                    a.line_skip[g + 1] = (uint8_t)sk;
                }
            } else atomicOr(&a.ctr->overflow, 64u);
            if (a.scan_list) ex_append_scan(a, g + 1, g + 1 < a.line_cap && !sk);
        }
        if (events && pos <= a.len) {
            uint32_t j = lower_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, pos);
            while (j <= a.n_logs && a.log_off[j] < pos + 64) {
                const uint64_t b = a.log_off[j];
                const uint32_t nlb = excl + __popcll(nl & ((1ull << (uint32_t)(b - pos)) - 1ull));
                a.first_line[j] = nlb + j;
                if (nlb + j < a.line_cap) {
                    if (j < a.n_logs) { a.line_start[nlb + j] = (uint32_t)b; if (a.line_skip) a.line_skip[nlb + j] = 0; }
                    if (j > 0) a.line_end[nlb + j - 1] = (uint32_t)b;
                } else atomicOr(&a.ctr->overflow, 64u);
                if (a.scan_list) ex_append_scan(a, nlb + j, nlb + j < a.line_cap && j < a.n_logs);
                ++j;
            }
        }
    }
}
