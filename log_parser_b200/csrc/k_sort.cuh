// Tile sort + counting rank for the event list and the score ranking.
// Part of engine.cu's translation unit (included inside its anonymous namespace); not a stand-alone header.
// ------------------------------------------------------------------ sorting a few ten thousand keys
// The event list and the score ranking hold ~10^4 keys: a device-wide radix sort spends 8-10 us per
// digit pass on them (13 passes for these two sorts), all of it launch and look-back latency. Here
// tiles of 4096 keys are sorted in shared memory (cub::BlockRadixSort, stable) by one kernel, and a
// second kernel finds every key's final position by counting, in each other tile, the keys that
// precede it (binary searches; earlier tiles win ties, which keeps the sort stable).
constexpr int SS_THREADS = 512, SS_ITEMS = 8, SS_TILE = SS_THREADS * SS_ITEMS;       // 4096 keys per tile: ~6 tiles for the C2 event list,
constexpr uint32_t SS_MAX_TILES = 32;                                                // so the counting pass searches 5 other tiles, not 10
constexpr int SS_GROUP = 16;              // tiles searched in lockstep by k_tile_rank (independent probes in flight)      // beyond 131072 keys the counting pass costs more than radix passes
using SsSorter = cub::BlockRadixSort<unsigned long long, SS_THREADS, SS_ITEMS, uint32_t, 6>;
__global__ void __launch_bounds__(SS_THREADS) k_tile_sort(const unsigned long long* __restrict__ in, uint32_t n, int end_bit,
                                                          unsigned long long* __restrict__ tile_keys, uint32_t* __restrict__ tile_idx) {
    using Sorter = SsSorter;
    extern __shared__ __align__(16) unsigned char ss_smem[];      // sizeof(Sorter::TempStorage) exceeds the 48 KiB static limit
    typename Sorter::TempStorage& tmp = *reinterpret_cast<typename Sorter::TempStorage*>(ss_smem);
    unsigned long long k[SS_ITEMS];
    uint32_t v[SS_ITEMS];
    const uint32_t base = blockIdx.x * SS_TILE + threadIdx.x * SS_ITEMS;
#pragma unroll
    for (int i = 0; i < SS_ITEMS; ++i) { k[i] = base + i < n ? in[base + i] : ~0ull; v[i] = base + i; }
    Sorter(tmp).Sort(k, v, 0, end_bit);
#pragma unroll
    for (int i = 0; i < SS_ITEMS; ++i) { tile_keys[base + i] = k[i]; tile_idx[base + i] = v[i]; }
}
// out_sorted (optional): keys scattered to their final positions; out_rank (optional): final position of
// original element i, written for i < *n_valid.
__global__ void __launch_bounds__(128) k_tile_rank(const unsigned long long* __restrict__ tile_keys, const uint32_t* __restrict__ tile_idx,
                                                   uint32_t n_tiles, uint32_t n, unsigned long long* __restrict__ out_sorted,
                                                   uint32_t* __restrict__ out_rank, const unsigned int* __restrict__ n_valid) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_tiles * SS_TILE) return;
    const uint32_t t0 = j / SS_TILE;
    const unsigned long long key = tile_keys[j];
    uint32_t r = j - t0 * SS_TILE;
    // branch-free binary searches over the SS_TILE keys of each other tile, SS_GROUP tiles in lockstep so
    // that their (dependent) probes overlap: keys <= key in earlier tiles, keys < key in later ones
    for (uint32_t tb = 0; tb < n_tiles; tb += SS_GROUP) {
        uint32_t pos[SS_GROUP];
        const unsigned long long* base[SS_GROUP];
        // earlier tiles count keys <= key, i.e. keys < key + 1 (key + 1 cannot overflow there: a key equal to
        // the all-ones padding in an earlier tile is handled by the saturation below)
        unsigned long long bound[SS_GROUP];
#pragma unroll
        for (int q = 0; q < SS_GROUP; ++q) {
            const uint32_t t = min(tb + q, n_tiles - 1);
            pos[q] = 0;
            base[q] = tile_keys + (size_t)t * SS_TILE;
            bound[q] = (t < t0 && key != ~0ull) ? key + 1 : key;      // count keys < bound
        }
#pragma unroll 1
        for (uint32_t step = SS_TILE / 2; step > 0; step >>= 1) {
            unsigned long long v[SS_GROUP];
#pragma unroll
            for (int q = 0; q < SS_GROUP; ++q) v[q] = base[q][pos[q] + step - 1];
#pragma unroll
            for (int q = 0; q < SS_GROUP; ++q) pos[q] += v[q] < bound[q] ? step : 0u;
        }
        unsigned long long v[SS_GROUP];
#pragma unroll
        for (int q = 0; q < SS_GROUP; ++q) v[q] = base[q][pos[q]];
#pragma unroll
        for (int q = 0; q < SS_GROUP; ++q) {
            const uint32_t t = tb + q;
            uint32_t cnt = pos[q] + (v[q] < bound[q] ? 1u : 0u);
            if (t < t0 && key == ~0ull) cnt = SS_TILE;                  // every key of an earlier tile is <= the padding key
            if (t < n_tiles && t != t0) r += cnt;
        }
    }
    if (out_sorted && r < n) out_sorted[r] = key;
    if (out_rank) { const uint32_t i = tile_idx[j]; if (i < *n_valid) out_rank[i] = r; }
}
// Up to TR_MAX keys (one small request: C1 has ~10^2 events): final positions by counting, in one launch. Every block
// holds all keys in shared memory and ranks 64 of them, four lanes per key splitting the comparisons; stable (equal keys
// keep their order). A tile sort of 64-bit keys is eleven block-wide radix passes whatever the number of keys (33 us).
constexpr uint32_t TR_MAX = 2048;
constexpr int TR_THREADS = 256, TR_PER = 4, TR_ELEMS = TR_THREADS / TR_PER;
__global__ void __launch_bounds__(TR_THREADS) k_tiny_rank(const unsigned long long* __restrict__ in, uint32_t n, unsigned long long* __restrict__ out_sorted,
                                                          uint32_t* __restrict__ out_rank, const unsigned int* __restrict__ n_valid) {
    __shared__ unsigned long long s[TR_MAX];
    for (uint32_t j = threadIdx.x; j < n; j += TR_THREADS) s[j] = in[j];
    __syncthreads();
    const uint32_t i = blockIdx.x * TR_ELEMS + threadIdx.x / TR_PER, part = threadIdx.x % TR_PER;
    const unsigned long long key = i < n ? s[i] : 0ull;
    uint32_t cnt = 0;
    if (i < n)
        for (uint32_t j = part; j < n; j += TR_PER) {
            const unsigned long long v = s[j];
            cnt += (v < key || (v == key && j < i)) ? 1u : 0u;
        }
    cnt += __shfl_xor_sync(0xffffffffu, cnt, 1);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, 2);
    if (i < n && part == 0) {
        if (out_sorted) out_sorted[cnt] = key;
        if (out_rank && i < *n_valid) out_rank[i] = cnt;
    }
}
__global__ void k_count_per_pattern(const unsigned long long* __restrict__ keys, const Counters* ctr, DevPlan P, unsigned long long* __restrict__ counts,
                                    uint32_t core_lo, uint32_t halo_after) {
    const uint32_t n = ctr->n_unique;
    const uint32_t core_hi = ctr->n_lines_total > halo_after ? ctr->n_lines_total - halo_after : 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)min(keys[i] >> P.gb, 0xFFFFFFFFull);
        if (r >= P.n_rx) continue;
        const uint32_t g = (uint32_t)(keys[i] & ((1ull << P.gb) - 1ull));
        if (g < core_lo || g >= core_hi) continue;
        for (uint32_t j = P.rxp_begin[r]; j < P.rxp_begin[r + 1]; ++j) atomicAdd(&counts[P.rxp_pat[j]], 1ull);
    }
}
