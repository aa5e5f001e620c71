// Fused scan: String.split("\\r?\\n") (AnalysisService.java:53) and the literal-gram prefilter of the
// N x P find() loop (AS:89-95) in ONE pass over the log. Replaces k_filter_count + k_scan_tiles + k_lines.
// Part of engine.cu's translation unit (included inside its anonymous namespace); not a stand-alone header.
//
// Shape: one CTA of 1024 threads per SM, split into two independent halves of 512 threads. A half owns a
// two-stage ring of 16 KiB log tiles in shared memory, filled by 1-D TMA bulk copies
// (cp.async.bulk ... mbarrier::complete_tx; the source is always this GPU's own memory) while the
// other tile is being worked on, so the bytes in flight do not depend on registers. Tiles are handed
// out by a ticket counter, which also makes the look-back below deadlock-free.
//
// Per tile, a thread owns 32 contiguous bytes:
//   phase A  newline mask by SWAR, and for each of its 32 positions one lookup of the 4-byte window in the
//            first-stage gram table. The table is a bit table replicated per lane: row = hash bits, the
//            word of lane L lives in bank L, so the 32 lookups of a warp never conflict (the byte table of
//            k_filter_count lost 55 % of its shared-memory wavefronts to bank conflicts). The wanted bit
//            is rotated to bit 31 and funnel-shifted into the hit mask: 6 instructions per position.
//   phase B  one block scan of (newlines, hits) per tile; the set bits are compacted into two lists in
//            shared memory so that the sparse work runs with full warps: first-stage hits are checked
//            against the exact gram key table (windows re-read from the staged tile) and survivors go to
//            the candidate queue with one global atomic per tile; newlines become line_end[g] /
//            line_start[g+1] at the global line index given by a decoupled look-back over the tiles'
//            newline counts (single pass: no second read of the log for the offsets).
namespace fs {
constexpr int HALF = 512, THREADS = 1024;
constexpr uint32_t TILE = 16384, HALO = 16, STAGE_BYTES = 16512;      // tile + halo, rounded up to 128
constexpr uint32_t ROWS = 1024;                                      // first-stage table rows (one word per lane and row)
constexpr uint32_t LIST = 1024;                                      // entries per compaction list and pass
constexpr uint32_t MUL1 = 0x9E3779B1u;
constexpr size_t TABLE_BYTES = (size_t)ROWS * 128;
constexpr size_t HALF_BYTES = 2 * STAGE_BYTES + LIST * 2 + LIST * 2 + LIST * 4 + 256;
constexpr size_t SMEM_BYTES = TABLE_BYTES + 2 * HALF_BYTES;

// row = (H >> 7) & (ROWS - 1), rotation = H & 31; the table bit of a gram is bit (rotation - 1) & 31
template <int V>
__host__ __device__ __forceinline__ uint32_t hash(uint32_t folded) {
#ifdef __CUDA_ARCH__
    if (V == 0) return __umulhi(folded, MUL1);
#else
    if (V == 0) return (uint32_t)(((unsigned long long)folded * MUL1) >> 32);
#endif
    return (folded * MUL1) >> 15;
}
inline uint32_t host_hash(int variant, uint32_t folded) { return variant == 0 ? hash<0>(folded) : hash<1>(folded); }

struct HalfState {
    unsigned long long bar[2];      // mbarriers of the two stages
    uint32_t tile[2];               // tile index loaded (or to be filled) into each stage
    uint32_t wtot[16], wex[16];
    uint32_t total, prefix, ncand;
    uint32_t k0, inner, events;
};
static_assert(sizeof(HalfState) <= 256, "HalfState must fit its slot");
}  // namespace fs

struct FscanArgs {
    const uint8_t* text;
    uint64_t len;
    uint32_t n_tiles;                 // tiles of the whole text (the last one covers position `len`)
    uint32_t tile_begin, tile_end;    // this launch's tiles (a log arriving from the host is scanned piece by piece)
    uint32_t* ticket;                 // zeroed counter of this launch
    unsigned long long* status;       // [n_tiles] look-back words, zeroed per call: flag << 32 | newline count
    uint32_t* tile_prefix;            // [n_tiles] out: newlines before each tile (k_resolve brackets its line search with it)
    const uint64_t* log_off;          // [n_logs+1]
    uint32_t n_logs;
    uint32_t* line_start;             // [line_cap]
    uint32_t* line_end;
    uint32_t* first_line;             // [n_logs+1]
    uint32_t line_cap;
    const uint32_t* bit_table;        // fs::ROWS words (variants 0/1) or the 64 KiB byte table (variant 2)
    const uint32_t* gram_keys;        // 2^table_bits folded grams, 0 = empty
    uint32_t table_bits;
    uint32_t* cand;
    uint32_t cand_cap;
    Counters* ctr;
    int use_filter;
};

__device__ __forceinline__ uint32_t fs_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fs_half_sync(int half) { asm volatile("bar.sync %0, %1;" ::"r"(1 + half), "n"(fs::HALF) : "memory"); }
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
__device__ __forceinline__ bool fs_tma_ok(uint32_t tile, uint64_t len) { return (uint64_t)tile * fs::TILE + fs::TILE + fs::HALO <= len; }

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
// With two 16-byte chunks packed as lo | hi << 4, bit b stands for byte 4 (b & 7) + (b >> 3) of the 32 bytes.
// Mask of the packed bits whose byte position is below x (0..32).
__device__ __forceinline__ uint32_t fs_below(uint32_t x) {
    const uint32_t n = x >> 2;
    return (((1u << n) - 1u) * 0x01010101u) | ((0x01010101u << n) & ((1u << (8u * (x & 3u))) - 1u));
}

// first-stage lookups of the 16 windows starting in one 16-byte chunk; w4 = the word that follows the chunk
template <int V>
__device__ __forceinline__ uint32_t fs_probe16(const uint4& v, uint32_t w4, const uint8_t* __restrict__ tab, uint32_t lane4) {
    const uint32_t w[5] = {v.x | 0x20202020u, v.y | 0x20202020u, v.z | 0x20202020u, v.w | 0x20202020u, w4 | 0x20202020u};
    uint32_t acc = 0;
#pragma unroll
    for (int i = 15; i >= 0; --i) {
        const uint32_t win = (i & 3) ? __funnelshift_r(w[i >> 2], w[(i >> 2) + 1], (i & 3) * 8) : w[i >> 2];
        if (V == 2) {
            acc = acc * 2u + tab[(win * GRAM_MUL) >> 16];
        } else {
            const uint32_t H = fs::hash<V>(win);
            const uint32_t word = *reinterpret_cast<const uint32_t*>(tab + ((H & ((fs::ROWS - 1u) << 7)) | lane4));
            acc = __funnelshift_l(__funnelshift_r(word, word, H), acc, 1);      // table bit -> bit 31 -> shifted into the mask
        }
    }
    return acc;
}

template <int V>
__global__ void __launch_bounds__(fs::THREADS, 1) k_fscan(FscanArgs a) {
    extern __shared__ __align__(128) uint8_t fs_smem[];
    const int half = threadIdx.x >> 9, t = threadIdx.x & (fs::HALF - 1), lane = t & 31, warp = t >> 5;
    uint8_t* const hbase = fs_smem + fs::TABLE_BYTES + (size_t)half * fs::HALF_BYTES;
    uint8_t* const stages = hbase;
    uint16_t* const s_nll = reinterpret_cast<uint16_t*>(hbase + 2 * fs::STAGE_BYTES);
    uint16_t* const s_hl = s_nll + fs::LIST;
    uint32_t* const s_out = reinterpret_cast<uint32_t*>(s_hl + fs::LIST);
    fs::HalfState* const hs = reinterpret_cast<fs::HalfState*>(s_out + fs::LIST);
    const uint32_t lane4 = (uint32_t)lane * 4u;
    const uint32_t full = 0xffffffffu;

    if (t == 0) {
        fs_mbar_init(&hs->bar[0], 1);
        fs_mbar_init(&hs->bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (a.use_filter) {
        if (V == 2) {
            const uint4* src = reinterpret_cast<const uint4*>(a.bit_table);
            uint4* dst = reinterpret_cast<uint4*>(fs_smem);
            for (uint32_t i = threadIdx.x; i < 65536 / 16; i += fs::THREADS) dst[i] = __ldg(src + i);
        } else {
            uint32_t* dst = reinterpret_cast<uint32_t*>(fs_smem);
            for (uint32_t i = threadIdx.x; i < fs::ROWS * 32; i += fs::THREADS) dst[i] = __ldg(a.bit_table + (i >> 5));
        }
    }
    __syncthreads();
    if (t == 0) {
        for (int s = 0; s < 2; ++s) {
            const uint32_t tile = a.tile_begin + atomicAdd(a.ticket, 1u);
            hs->tile[s] = tile;
            if (tile < a.tile_end && fs_tma_ok(tile, a.len))
                fs_tma_load(stages + s * fs::STAGE_BYTES, a.text + (uint64_t)tile * fs::TILE, fs::TILE + fs::HALO, &hs->bar[s]);
        }
    }
    fs_half_sync(half);

    uint32_t parity = 0;
    for (uint32_t n = 0;; ++n) {
        const int s = n & 1;
        const uint32_t tile = hs->tile[s];
        if (tile >= a.tile_end) break;
        const uint64_t base = (uint64_t)tile * fs::TILE;
        uint8_t* const stage = stages + s * fs::STAGE_BYTES;
        const bool tma = fs_tma_ok(tile, a.len);
        if (tma) {
            fs_mbar_wait(&hs->bar[s], (parity >> s) & 1u);
            parity ^= 1u << s;
        } else {        // the last tiles of the text: bounds-checked fill, bytes at or beyond `len` read as zero
            for (uint32_t i = t; i < fs::STAGE_BYTES / 16; i += fs::HALF)
                reinterpret_cast<uint4*>(stage)[i] = load16_any(a.text, base + 16ull * i, a.len);
            fs_half_sync(half);
        }
        if (t == 32) {      // where the logs of a batch begin, relative to this tile
            const uint64_t tile_end = base + fs::TILE;
            const uint32_t j0 = upper_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, base);
            hs->k0 = j0 - 1;
            hs->inner = (j0 <= a.n_logs && a.log_off[j0] < tile_end) ? 1u : 0u;
            const uint32_t j1 = lower_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, base);
            hs->events = (j1 <= a.n_logs && a.log_off[j1] < tile_end) ? 1u : 0u;
        }

        // ---- phase A: 32 contiguous bytes per thread
        const uint4 v0 = *reinterpret_cast<const uint4*>(stage + 32 * t);
        const uint4 v1 = *reinterpret_cast<const uint4*>(stage + 32 * t + 16);
        const uint32_t P = fs_nl_spread(v0) | (fs_nl_spread(v1) << 4);
        uint32_t hm = 0;
        if (a.use_filter) {
            uint32_t nxt = __shfl_down_sync(full, v0.x, 1);
            if (lane == 31) nxt = *reinterpret_cast<const uint32_t*>(stage + 32 * (t + 1));
            hm = fs_probe16<V>(v0, v1.x, fs_smem, lane4) | (fs_probe16<V>(v1, nxt, fs_smem, lane4) << 16);
            if (!tma) {     // windows that start at or beyond `len`
                const uint64_t p0 = base + 32ull * t;
                if (p0 >= a.len) hm = 0;
                else if (a.len - p0 < 32) hm &= (1u << (uint32_t)(a.len - p0)) - 1u;
            }
        }

        // ---- phase B: block scan of (newlines, hits)
        const uint32_t cn = __popc(P), ch = __popc(hm);
        const uint32_t packed = cn | (ch << 16);
        uint32_t inc = packed;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(full, inc, o); if (lane >= o) inc += y; }
        if (lane == 31) hs->wtot[warp] = inc;
        fs_half_sync(half);
        if (warp == 0) {
            const uint32_t w = lane < 16 ? hs->wtot[lane] : 0u;
            uint32_t wi = w;
#pragma unroll
            for (int o = 1; o < 16; o <<= 1) { const uint32_t y = __shfl_up_sync(full, wi, o); if (lane >= o) wi += y; }
            if (lane < 16) hs->wex[lane] = wi - w;
            if (lane == 15) {
                hs->total = wi;
                hs->ncand = 0;
                // the tile's newline count is published as early as possible: later tiles look back at it
                volatile unsigned long long* stw = a.status + tile;
                if (tile != 0) *stw = (1ull << 32) | (wi & 0xFFFFu);
            }
        }
        fs_half_sync(half);
        const uint32_t excl = hs->wex[warp] + inc - packed;
        const uint32_t excl_nl = excl & 0xFFFFu, excl_hit = excl >> 16;
        const uint32_t n_nl = hs->total & 0xFFFFu, n_hit = hs->total >> 16;
        const uint32_t passes = max(1u, max((n_nl + fs::LIST - 1) / fs::LIST, (n_hit + fs::LIST - 1) / fs::LIST));
        uint32_t prefix = 0;
        for (uint32_t pass = 0; pass < passes; ++pass) {
            const uint32_t lo = pass * fs::LIST;
            // compaction: position-in-tile of every newline / first-stage hit, in position order
            {
                uint32_t m = P;
                while (m) {
                    const uint32_t b = __ffs(m) - 1;
                    m &= m - 1;
                    const uint32_t nq = b & 7u, j = b >> 3;
                    const uint32_t lt = (((1u << nq) - 1u) * 0x01010101u) | ((0x01010101u << nq) & ((1u << (8u * j)) - 1u));
                    const uint32_t r = excl_nl + __popc(P & lt) - lo;
                    if (r < fs::LIST) s_nll[r] = (uint16_t)(32 * t + 4 * nq + j);
                }
                m = hm;
                uint32_t r = excl_hit - lo;
                while (m) {
                    const uint32_t i = __ffs(m) - 1;
                    m &= m - 1;
                    if (r < fs::LIST) s_hl[r] = (uint16_t)(32 * t + i);
                    ++r;
                }
            }
            fs_half_sync(half);
            if (warp == 0) {
                if (pass == 0) {
                    // decoupled look-back over the tiles' newline counts. The halves run in lock step, so the
                    // nearest published prefix is usually a whole wave (~2 x SM count tiles) back: each round
                    // loads LB_U x 32 status words at once (independent loads, one L2 round trip) instead of
                    // walking back 32 tiles per round trip. Lane i of group u inspects tile look - 32 u - i.
                    const uint32_t aggregate = n_nl;
                    uint32_t run = 0;
                    if (tile != 0) {
                        constexpr int LB_U = 10;
                        int look = (int)tile - 1;
                        bool done = false;
                        while (!done) {
                            unsigned long long sw[LB_U];
#pragma unroll
                            for (int u = 0; u < LB_U; ++u) {
                                const int tt = look - 32 * u - lane;
                                sw[u] = tt >= 0 ? *(volatile unsigned long long*)(a.status + tt) : (2ull << 32);
                            }
                            uint32_t add = 0;       // per-lane partial sum of the groups accepted in this round
                            int accepted = 0;
                            bool stall = false;
#pragma unroll
                            for (int u = 0; u < LB_U; ++u) {
                                if (done || stall) continue;
                                const uint32_t flag = (uint32_t)(sw[u] >> 32), val = (uint32_t)sw[u];
                                const unsigned pm = __ballot_sync(full, flag == 2u), zm = __ballot_sync(full, flag == 0u);
                                if (pm) {
                                    const int fp = __ffs(pm) - 1;
                                    const unsigned upto = fp == 31 ? 0xFFFFFFFFu : ((2u << fp) - 1u);
                                    if ((zm & upto) == 0) { if (lane <= fp) add += val; done = true; }
                                    else stall = true;
                                } else if (zm == 0) { add += val; ++accepted; }
                                else stall = true;
                            }
#pragma unroll
                            for (int o = 16; o; o >>= 1) add += __shfl_xor_sync(full, add, o);
                            run += add;
                            look -= 32 * accepted;
                            if (stall) __nanosleep(100);
                        }
                    }
                    if (lane == 0) {
                        *(volatile unsigned long long*)(a.status + tile) = (2ull << 32) | (unsigned long long)(run + aggregate);
                        hs->prefix = run;
                        a.tile_prefix[tile] = run;
                        if (tile + 1 == a.n_tiles) { a.ctr->total_nl = run + aggregate; a.ctr->n_lines_total = run + aggregate + a.n_logs; }
                    }
                }
            } else {
                // first-stage hits against the exact gram keys; the window is re-read from the staged tile
                const uint32_t cnt = min(n_hit - min(n_hit, lo), fs::LIST);
                const uint32_t tmask = (1u << a.table_bits) - 1u;
                for (uint32_t e0 = (uint32_t)(warp - 1) * 32u; e0 < cnt; e0 += 15u * 32u) {
                    const uint32_t e = e0 + lane;
                    bool keep = false;
                    uint32_t pt = 0;
                    if (e < cnt) {
                        pt = s_hl[e];
                        const uint32_t al = pt & ~3u;
                        const uint32_t wlo = *reinterpret_cast<const uint32_t*>(stage + al), whi = *reinterpret_cast<const uint32_t*>(stage + al + 4);
                        const uint32_t f = __funnelshift_r(wlo, whi, (pt & 3u) * 8u) | 0x20202020u;
                        uint32_t slot = (f * GRAM_MUL2) >> (32 - a.table_bits);
                        uint32_t key;
                        while ((key = __ldg(a.gram_keys + slot)) != 0u && key != f) slot = (slot + 1) & tmask;
                        keep = key == f;
                    }
                    const unsigned votes = __ballot_sync(full, keep);
                    if (votes) {
                        uint32_t ob = 0;
                        if (lane == 0) ob = atomicAdd(&hs->ncand, (uint32_t)__popc(votes));
                        ob = __shfl_sync(full, ob, 0) + __popc(votes & ((1u << lane) - 1u));
                        if (keep) s_out[ob] = (uint32_t)(base + pt);
                    }
                }
            }
            fs_half_sync(half);
            if (pass == 0) prefix = hs->prefix;
            {   // newlines -> line_end[g] / line_start[g+1] (the \r?\n rule; log boundaries of a batch shift the index)
                const uint32_t cnt = min(n_nl - min(n_nl, lo), fs::LIST);
                const uint32_t k0 = hs->k0;
                const bool inner = hs->inner != 0;
                for (uint32_t e = t; e < cnt; e += fs::HALF) {
                    const uint32_t pt = s_nll[e];
                    const uint64_t p = base + pt;
                    uint32_t k = k0;
                    if (inner) k = upper_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, p) - 1;
                    const uint32_t g = prefix + lo + e + k;
                    bool cr = false;
                    if (p > a.log_off[k]) cr = (pt ? stage[pt - 1] : a.text[p - 1]) == '\r';
                    if (g + 1 < a.line_cap) {
                        a.line_end[g] = (uint32_t)(p - (cr ? 1 : 0));
                        a.line_start[g + 1] = (uint32_t)(p + 1);
                    } else atomicOr(&a.ctr->overflow, 64u);
                }
            }
            if (warp == 15) {   // survivors of this pass -> candidate queue, one global atomic
                const uint32_t nc = hs->ncand;
                if (nc) {
                    uint32_t gb = 0;
                    if (lane == 0) gb = atomicAdd(&a.ctr->n_cand, nc);
                    gb = __shfl_sync(full, gb, 0);
                    for (uint32_t q = lane; q < nc; q += 32) {
                        if (gb + q < a.cand_cap) a.cand[gb + q] = s_out[q];
                        else atomicOr(&a.ctr->overflow, 1u);
                    }
                }
            }
            if (pass + 1 < passes) {
                fs_half_sync(half);
                if (t == 0) hs->ncand = 0;
            }
        }
        if (hs->events && base + 32ull * t <= a.len) {      // start of log j / end of log j-1 inside this thread's bytes
            const uint64_t pos = base + 32ull * t;
            uint32_t j = lower_bound_dev<uint64_t>(a.log_off, a.n_logs + 1, pos);
            while (j <= a.n_logs && a.log_off[j] < pos + 32) {
                const uint64_t b = a.log_off[j];
                const uint32_t nlb = prefix + excl_nl + __popc(P & fs_below((uint32_t)(b - pos)));
                a.first_line[j] = nlb + j;
                if (nlb + j < a.line_cap) {
                    if (j < a.n_logs) a.line_start[nlb + j] = (uint32_t)b;
                    if (j > 0) a.line_end[nlb + j - 1] = (uint32_t)b;
                } else atomicOr(&a.ctr->overflow, 64u);
                ++j;
            }
        }
        fs_half_sync(half);       // the stage and the lists are free again
        if (t == 0) {
            const uint32_t nt = a.tile_begin + atomicAdd(a.ticket, 1u);
            hs->tile[s] = nt;
            if (nt < a.tile_end && fs_tma_ok(nt, a.len))
                fs_tma_load(stage, a.text + (uint64_t)nt * fs::TILE, fs::TILE + fs::HALO, &hs->bar[s]);
        }
    }
}
