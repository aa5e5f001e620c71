// lp_split_lines_noalloc: the line-splitting step for callers that are not allowed to allocate
// device memory.
//
// Who needs this: hosts that run with a fixed, pre-sized GPU memory plan — a JVM service that
// carves one arena at start-up and hands out slices, a plugin loaded into a process whose
// allocator must stay the only caller of cudaMalloc, or code that runs where an allocation would
// synchronise the whole device at the wrong moment. Every other entry point of this library keeps
// a grow-only scratch pool behind the lp_ctx (cudaMalloc on first use / on growth). This one
// never calls cudaMalloc, cudaFree or any allocating library routine: its scratch (per-tile
// newline counts + a small header) lives in a file-scope __device__ array that every GPU loading
// the library carries until the process exits (~128 KiB). The caller supplies input and outputs.
//
// Limits and rules a caller must know:
//   * at most LP_NOALLOC_MAX_BYTES (1 GiB) of log per call; larger inputs are rejected with
//     LP_ERR_LIMIT before anything is launched;
//   * the first call on a given device resolves the scratch array's device address
//     (cudaGetSymbolAddress), with that device current, once;
//   * every call is synchronous: it reads the overflow flag back and waits for `stream` to drain
//     before it returns (the caller's current device is left as it was);
//   * all calls on one device share the one scratch array. Two calls whose kernels may overlap on
//     the GPU (different streams, different host threads) race on it: serialise them yourself;
//   * the scratch keeps whatever the previous call left behind; the words that must start at zero
//     are cleared at the top of every call.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/logparser_b200.h"

namespace {

constexpr int NA_THREADS = 256;
constexpr int NA_ITERS = 8;
constexpr uint32_t NA_TILE = NA_THREADS * 16 * NA_ITERS;                       // 32 KiB of log per tile
constexpr unsigned long long NA_MAX_BYTES = 1ull << 30;                        // 1 GiB per call
constexpr uint32_t NA_MAX_TILES = (uint32_t)(NA_MAX_BYTES / NA_TILE) + 1;
constexpr uint32_t NA_HDR_WORDS = 64;                                          // 256-byte header
// header words: [0] newline total, [1] capacity-overflow flag, [2] 1 + index of the last non-empty segment; the rest is spare.
// then one exclusive newline count per tile.
static __device__ __align__(256) uint32_t g_split_scratch[NA_HDR_WORDS + NA_MAX_TILES];

__device__ __forceinline__ uint32_t na_zero_flags(uint32_t x) {
    uint32_t t = ((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x;
    return ~(t | 0x7F7F7F7Fu);
}
__device__ __forceinline__ uint32_t na_mask4(uint32_t w) {
    uint32_t m = na_zero_flags(w ^ 0x0A0A0A0Au) >> 7;
    return ((m * 0x00204081u) >> 21) & 0xFu;
}
__device__ __forceinline__ uint32_t na_mask16(const uint8_t* __restrict__ text, unsigned long long pos, unsigned long long len) {
    if (pos >= len) return 0;
    if (pos + 16 <= len) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(text + pos));
        return na_mask4(v.x) | (na_mask4(v.y) << 4) | (na_mask4(v.z) << 8) | (na_mask4(v.w) << 12);
    }
    uint32_t m = 0;
    for (int i = 0; i < 16 && pos + i < len; ++i) m |= (text[pos + i] == '\n' ? 1u : 0u) << i;
    return m;
}

__global__ void __launch_bounds__(NA_THREADS) na_count(const uint8_t* __restrict__ text, unsigned long long len, uint32_t n_tiles,
                                                       uint32_t* __restrict__ scratch) {
    __shared__ uint32_t warp_sum[NA_THREADS / 32];
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const unsigned long long base = (unsigned long long)tile * NA_TILE + (unsigned long long)threadIdx.x * 16;
        uint32_t cnt = 0;
#pragma unroll
        for (int it = 0; it < NA_ITERS; ++it) cnt += __popc(na_mask16(text, base + (unsigned long long)it * NA_THREADS * 16, len));
        for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = cnt;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t s = 0;
            for (int w = 0; w < NA_THREADS / 32; ++w) s += warp_sum[w];
            scratch[NA_HDR_WORDS + tile] = s;
        }
        __syncthreads();
    }
}

// one block: exclusive scan of the tile counts in place; publishes the total; writes the two
// offsets that do not belong to any newline (start of the first line, end of the last one)
__global__ void __launch_bounds__(1024) na_scan(uint32_t* __restrict__ scratch, uint32_t n_tiles, unsigned long long len,
                                                uint32_t* __restrict__ line_start, uint32_t* __restrict__ line_end, unsigned long long cap) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry;
    uint32_t* counts = scratch + NA_HDR_WORDS;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_tiles; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < n_tiles ? counts[i] : 0;
        uint32_t inc = v;
        for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if ((threadIdx.x & 31) >= o) inc += t; }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
        __syncthreads();
        if (threadIdx.x < 32) {
            const uint32_t w = warp_tot[threadIdx.x];
            uint32_t wi = w;
            for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, wi, o); if (threadIdx.x >= o) wi += t; }
            warp_tot[threadIdx.x] = wi - w;
        }
        __syncthreads();
        const uint32_t excl = carry + warp_tot[threadIdx.x >> 5] + inc - v;
        if (i < n_tiles) counts[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const uint32_t total = carry;
        scratch[0] = total;
        if ((unsigned long long)total + 1 > cap) scratch[1] = 1;
        else { line_start[0] = 0; line_end[total] = (uint32_t)len; }
    }
}

// every '\n' at p ends line g (at p, or p-1 when a '\r' precedes it) and starts line g+1 at p+1
__global__ void __launch_bounds__(NA_THREADS) na_write(const uint8_t* __restrict__ text, unsigned long long len, uint32_t n_tiles,
                                                       const uint32_t* __restrict__ scratch, uint32_t* __restrict__ line_start,
                                                       uint32_t* __restrict__ line_end, unsigned long long cap) {
    __shared__ uint32_t s_warp[2][NA_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if ((unsigned long long)scratch[0] + 1 > cap) return;            // outputs too small: nothing is written
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        uint32_t running = scratch[NA_HDR_WORDS + tile];
#pragma unroll 1
        for (int it = 0; it < NA_ITERS; ++it) {
            const unsigned long long pos = (unsigned long long)tile * NA_TILE + ((unsigned long long)it * NA_THREADS + threadIdx.x) * 16;
            const uint32_t nlm = na_mask16(text, pos, len);
            const uint32_t cnt = __popc(nlm);
            uint32_t inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
            if (lane == 31) s_warp[it & 1][warp] = inc;
            __syncthreads();
            uint32_t wbase = 0, btotal = 0;
#pragma unroll
            for (int w = 0; w < NA_THREADS / 32; ++w) { uint32_t t = s_warp[it & 1][w]; if (w < warp) wbase += t; btotal += t; }
            const uint32_t excl = running + wbase + inc - cnt;
            running += btotal;
            uint32_t m = nlm;
            while (m) {
                const int i = __ffs(m) - 1;
                m &= m - 1;
                const unsigned long long p = pos + i;
                const uint32_t g = excl + __popc(nlm & ((1u << i) - 1u));
                const bool cr = p > 0 && text[p - 1] == '\r';
                line_end[g] = (uint32_t)(p - (cr ? 1 : 0));
                line_start[g + 1] = (uint32_t)(p + 1);
            }
        }
        __syncthreads();
    }
}

// 1 + index of the last non-empty segment, found by all threads (a log may end in millions of blank lines)
__global__ void na_last_nonempty(uint32_t* __restrict__ scratch, const uint32_t* __restrict__ line_start,
                                 const uint32_t* __restrict__ line_end, unsigned long long cap) {
    const unsigned long long segments = (unsigned long long)scratch[0] + 1;
    if (segments > cap) return;
    uint32_t best = 0;
    for (unsigned long long g = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; g < segments; g += (unsigned long long)gridDim.x * blockDim.x)
        if (line_end[g] != line_start[g]) best = (uint32_t)g + 1;      // g grows along the loop: the last hit is this thread's largest
    for (int o = 16; o; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    if ((threadIdx.x & 31) == 0 && best) atomicMax(&scratch[2], best);
}

// String.split's limit-0 rule: drop trailing empty strings; an input without any separator stays [input]
__global__ void na_finish(const uint32_t* __restrict__ scratch, const uint32_t* __restrict__ line_start,
                          const uint32_t* __restrict__ line_end, unsigned long long cap, unsigned long long* __restrict__ counts) {
    const unsigned long long segments = (unsigned long long)scratch[0] + 1;
    counts[1] = segments;
    counts[2] = scratch[1];
    if (scratch[1]) { counts[0] = 0; return; }
    counts[0] = segments > 1 ? (unsigned long long)scratch[2] : segments;
}

std::mutex g_setup_mu;
constexpr int NA_MAX_DEVICES = 64;
uint32_t* g_scratch_addr[NA_MAX_DEVICES];   // device address of g_split_scratch, per device
bool g_scratch_ready[NA_MAX_DEVICES];

void say(char* err, size_t errlen, const char* msg) {
    if (err && errlen) snprintf(err, errlen, "%s", msg);
}
int cuda_fail(char* err, size_t errlen, const char* what, cudaError_t e) {
    if (err && errlen) snprintf(err, errlen, "%s: %s", what, cudaGetErrorString(e));
    return LP_ERR_CUDA;
}

}  // namespace

extern "C" int lp_split_lines_noalloc(int device, const uint8_t* d_log, size_t len, uint32_t* d_line_start, uint32_t* d_line_end,
                                      size_t capacity, uint64_t* d_counts, void* stream, char* err, size_t errlen) {
    if (err && errlen) err[0] = 0;
    if (device < 0 || device >= NA_MAX_DEVICES) { say(err, errlen, "device index out of range"); return LP_ERR_INVALID; }
    if ((len && !d_log) || !d_line_start || !d_line_end || !d_counts) { say(err, errlen, "NULL device pointer"); return LP_ERR_INVALID; }
    if ((reinterpret_cast<uintptr_t>(d_log) & 15) != 0) { say(err, errlen, "d_log must be 16-byte aligned"); return LP_ERR_INVALID; }
    if ((unsigned long long)len > NA_MAX_BYTES) {
        if (err && errlen)
            snprintf(err, errlen, "log of %zu bytes exceeds the %llu-byte limit of the allocation-free entry point "
                     "(its scratch is a fixed __device__ array); split the input or use lp_analyze*", len, NA_MAX_BYTES);
        return LP_ERR_LIMIT;
    }
    if (capacity == 0) { say(err, errlen, "capacity is 0: even an empty log has one line segment"); return LP_ERR_LIMIT; }
    // the caller's current device is put back on every way out
    struct DeviceGuard {
        int prev = -1;
        ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    } guard;
    if (cudaGetDevice(&guard.prev) != cudaSuccess) guard.prev = -1;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail(err, errlen, "cudaSetDevice", e);

    // one-time set-up per device, with that device current: the array's name is only a host-side
    // handle here, so ask the runtime for the address the kernels must be given
    uint32_t* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_setup_mu);
        if (!g_scratch_ready[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_split_scratch);
            if (e != cudaSuccess) return cuda_fail(err, errlen, "cudaGetSymbolAddress(g_split_scratch)", e);
            g_scratch_addr[device] = static_cast<uint32_t*>(p);
            g_scratch_ready[device] = true;
        }
        scratch = g_scratch_addr[device];
    }

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const uint32_t n_tiles = (uint32_t)(((unsigned long long)len + NA_TILE - 1) / NA_TILE);
    // the header (total, overflow flag) must start at zero; tile counts are fully rewritten below
    e = cudaMemsetAsync(scratch, 0, NA_HDR_WORDS * sizeof(uint32_t), st);
    if (e != cudaSuccess) return cuda_fail(err, errlen, "cudaMemsetAsync(scratch header)", e);
    const uint32_t grid = n_tiles < 148u * 8u ? (n_tiles ? n_tiles : 1u) : 148u * 8u;
    if (n_tiles) na_count<<<grid, NA_THREADS, 0, st>>>(d_log, len, n_tiles, scratch);
    na_scan<<<1, 1024, 0, st>>>(scratch, n_tiles, len, d_line_start, d_line_end, capacity);
    if (n_tiles) na_write<<<grid, NA_THREADS, 0, st>>>(d_log, len, n_tiles, scratch, d_line_start, d_line_end, capacity);
    na_last_nonempty<<<grid, NA_THREADS, 0, st>>>(scratch, d_line_start, d_line_end, capacity);
    na_finish<<<1, 1, 0, st>>>(scratch, d_line_start, d_line_end, capacity, reinterpret_cast<unsigned long long*>(d_counts));
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(err, errlen, "kernel launch", e);
    unsigned long long flag = 0;
    e = cudaMemcpyAsync(&flag, d_counts + 2, sizeof flag, cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess) return cuda_fail(err, errlen, "cudaMemcpyAsync", e);
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cuda_fail(err, errlen, "cudaStreamSynchronize", e);
    if (flag) { say(err, errlen, "capacity is smaller than the number of line segments (newlines + 1); d_counts[1] holds the required size"); return LP_ERR_LIMIT; }
    return LP_OK;
}
