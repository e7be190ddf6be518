// ws_radix_sort.cu -- hand-written onesweep LSD radix sort of (u64 key, u32 value) pairs on key
// bits [0, end_bit).  Replaces cub::DeviceRadixSort::SortPairs (DGR rasterizer_impl.cu:304-309;
// KNN simple_knn.cu:211-214).  No CUB, no Thrust.
//
//   rs_histogram   one read of the keys -> global digit histograms of every pass at once
//                  (the exclusive scan of a pass's 256 bins -> digit bases is redone by every CTA of the pass)
//   rs_onesweep    one launch per pass: every CTA takes a tile of 256 x 16 keys (12 with 64-bit keys) (ticket
//                  from an atomic counter so predecessors are always resident), ranks them in
//                  shared memory, learns the number of equal digits in all earlier tiles by a
//                  decoupled look-back over per-(tile,digit) status words, and scatters.
//                  Keys and values are read once and written once per pass.
//
// Stability (needed for bit-exact parity of point_list: equal (tile,depth) keys must keep the
// emission order = ascending Gaussian id).  Within a digit the output order is
//   tile order (look-back prefix)  ->  warp order inside the tile (exclusive scan over the
//   per-warp digit counters)  ->  item order inside the warp (warp-striped load: item i of
//   lane l is element i*32 + l of the warp's segment, items ranked i-ascending, lanes by
//   lanemask_lt).  That is exactly input order.
//
// Algorithmic HBM bytes: 8n (histogram) + passes * 24n (12 B read + 12 B written per pair).
#include "ws_common.cuh"

namespace ws {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
#ifndef WS_RS_ITEMS
#define WS_RS_ITEMS 16
#endif
#ifndef WS_RS_MINB
#define WS_RS_MINB 3
#endif
constexpr int RS_ITEMS = WS_RS_ITEMS;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS; // 3072 pairs per CTA
constexpr int RS_MAX_PASSES = 8;
#ifndef WS_RS_LOOK
#define WS_RS_LOOK 8
#endif
constexpr int RS_LOOK = WS_RS_LOOK; // predecessors read per look-back round trip

constexpr uint32_t RS_FLAG_AGG = 1u << 30; // tile aggregate available
constexpr uint32_t RS_FLAG_INC = 2u << 30; // inclusive prefix available
constexpr uint32_t RS_VAL_MASK = (1u << 30) - 1;

// Pairs per thread.  (4 for small inputs -- more CTAs per SM for the 1 M-pair depth sort -- was
// measured slower, 25 vs 20 us per pass: a single-wave pass is bound by the look-back chain, which
// grows with the number of tiles.)
constexpr int RS_ITEMS64 = 12;           // 64-bit keys (keys64 binning mode): 48 KB of static shared memory at most
constexpr int RS_ITEMS_MIN = RS_ITEMS < RS_ITEMS64 ? RS_ITEMS : RS_ITEMS64;
template <typename K> constexpr int items_of() { return sizeof(K) == 8 ? RS_ITEMS64 : RS_ITEMS; }
// tiles the scratch is sized for (the smaller tile of the two key widths)
size_t sort_tiles(size_t n)
{
    const size_t t = (size_t)RS_THREADS * RS_ITEMS_MIN;
    return (n + t - 1) / t;
}

SortScratch SortScratch::from(char *&p, size_t n)
{
    SortScratch s;
    // one contiguous block: [hist 8 x 256 | counters 64 | status passes x tiles x 256] (cleared by one memset)
    size_t nt = sort_tiles(n);
    s.hist = carve<uint32_t>(p, RS_MAX_PASSES * 256 + 64 + (size_t)RS_MAX_PASSES * (nt ? nt : 1) * 256);
    s.counters = s.hist + RS_MAX_PASSES * 256;
    s.status = s.counters + 64;
    return s;
}

template <typename K>
__global__ void __launch_bounds__(256) rs_histogram(const K *__restrict__ keys, size_t n, int end_bit,
                                                    uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[RS_MAX_PASSES * 256];
    const int npass = (end_bit + 7) / 8;
    for (int j = threadIdx.x; j < npass * 256; j += blockDim.x) sh[j] = 0;
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const K key = keys[i];
        for (int p = 0; p < npass; p++) {
            const int w = sort_digit_bits(end_bit), bits = min(w, end_bit - w * p);
            const uint32_t d = (uint32_t)(key >> (w * p)) & ((1u << bits) - 1u);
            atomicAdd(&sh[p * 256 + d], 1u);
        }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < npass * 256; j += blockDim.x) {
        const uint32_t c = sh[j];
        if (c) atomicAdd(&hist[j], c);
    }
}

// exclusive scan over 256 values held one per thread (256 threads); returns the exclusive prefix
__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t *s_w /*[8]*/)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) s_w[w] = inc;
    __syncthreads();
    uint32_t base = 0;
#pragma unroll
    for (int k = 0; k < RS_WARPS; k++)
        if (k < w) base += s_w[k];
    __syncthreads();
    return base + inc - v;
}

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t *p, uint32_t v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// One tile of one pass.  FULL = all 256*ITEMS slots of the tile hold pairs (every tile but the
// last): no bounds checks, which otherwise cost a 64-bit compare per access.
template <typename K, int ITEMS, bool FULL>
__device__ __forceinline__ void rs_tile_body(const K *__restrict__ kin, const uint32_t *__restrict__ vin,
                                             K *__restrict__ keys_out, uint32_t *__restrict__ vals_out,
                                             uint32_t nt, uint32_t tile, int shift, uint32_t mask,
                                             const uint32_t *__restrict__ bins, uint32_t *__restrict__ status,
                                             K *s_keys, uint32_t *s_vals, uint32_t (*s_whist)[256],
                                             uint32_t *s_dstart, uint32_t *s_goff, uint32_t *s_w)
{
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t o0 = (uint32_t)w * (ITEMS * 32) + lane; // tile-local slot of item 0 (warp-striped)

    K key[ITEMS];
    uint32_t val[ITEMS];
    uint32_t rank[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const uint32_t o = o0 + i * 32;
        const bool valid = FULL || o < nt;
        key[i] = valid ? kin[o] : (K)~(K)0;
        val[i] = valid ? vin[o] : 0u;
    }

    // ---- rank inside the warp, item by item (stable) ----
    // The lanes holding equal digits find each other (match.any); their first lane adds the group's size to the
    // warp's counter of that digit with a shared-memory atomic and hands the old value to the group.  The atomics of
    // one warp on one address apply in program order (item order), and nothing else orders the items: the twelve
    // match / atomic / shuffle chains overlap instead of running one after the other behind warp barriers.
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const bool valid = FULL || (o0 + i * 32) < nt;
        const uint32_t d = valid ? ((uint32_t)(key[i] >> shift) & mask) : 0x100u;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t pre = 0;
        if (valid && lane == leader) pre = atomicAdd(&s_whist[w][d], (uint32_t)__popc(peers));
        pre = __shfl_sync(0xffffffffu, pre, leader);
        rank[i] = pre + __popc(peers & lt_mask);
        __syncwarp(); // memory order between this item's atomics and the next item's (different lanes may lead)
    }
    __syncthreads();

    // ---- thread d owns digit d: exclusive scan over the warps, tile count, look-back ----
    uint32_t run = 0;
#pragma unroll
    for (int k = 0; k < RS_WARPS; k++) {
        const uint32_t c = s_whist[k][tid];
        s_whist[k][tid] = run;
        run += c;
    }
    uint32_t *my_status = status + (size_t)tile * 256 + tid;
    uint32_t excl = 0;
    if (tile == 0) {
        st_relaxed(my_status, run | RS_FLAG_INC);
    } else {
        st_relaxed(my_status, run | RS_FLAG_AGG);
        // Look-back, RS_LOOK predecessors per round trip (one dependent L2 round trip per
        // predecessor serialises the pass: the tiles just before this one started at about the
        // same time and have only their aggregate out).
        int look = (int)tile - 1;
        bool found = false;
        while (!found) {
            uint32_t v[RS_LOOK];
#pragma unroll
            for (int k = 0; k < RS_LOOK; k++)
                v[k] = (look - k >= 0) ? ld_relaxed(status + (size_t)(look - k) * 256 + tid) : RS_FLAG_INC;
            int used = 0;
#pragma unroll
            for (int k = 0; k < RS_LOOK; k++) {
                if (!found && used == k && (v[k] & (RS_FLAG_AGG | RS_FLAG_INC)) != 0) {
                    excl += v[k] & RS_VAL_MASK;
                    used = k + 1;
                    found = (v[k] & RS_FLAG_INC) != 0;
                }
            }
            look -= used;
            if (!found && used == 0) __nanosleep(20);
        }
        st_relaxed(my_status, ((excl + run) & RS_VAL_MASK) | RS_FLAG_INC);
    }
    const uint32_t dstart = block_excl_scan_256(run, s_w); // contains __syncthreads
    // digit base = exclusive prefix of the pass's global histogram, scanned here by every CTA (256 values, one block
    // scan) rather than by a kernel of its own between the histogram and the first pass
    const uint32_t base = block_excl_scan_256(bins[tid], s_w);
    s_dstart[tid] = dstart;
    s_goff[tid] = base + excl - dstart;
    __syncthreads();

    // ---- scatter into shared memory in sorted order ----
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const bool valid = FULL || (o0 + i * 32) < nt;
        if (valid) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
            const uint32_t pos = s_dstart[d] + s_whist[w][d] + rank[i];
            s_keys[pos] = key[i];
            s_vals[pos] = val[i];
        }
    }
    __syncthreads();

    // ---- write out: consecutive positions of one digit are consecutive in global memory ----
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        const uint32_t p = tid + j * RS_THREADS;
        if (FULL || p < nt) {
            const K k = s_keys[p];
            const uint32_t d = (uint32_t)(k >> shift) & mask;
            const uint32_t dst = s_goff[d] + p; // < n < 2^30
            keys_out[dst] = k;
            vals_out[dst] = s_vals[p];
        }
    }
}

template <typename K, int ITEMS>
__global__ void __launch_bounds__(RS_THREADS, sizeof(K) == 4 ? WS_RS_MINB : 3)
rs_onesweep(const K *__restrict__ keys_in, K *__restrict__ keys_out,
            const uint32_t *__restrict__ vals_in, uint32_t *__restrict__ vals_out, size_t n, int shift,
            int bits, const uint32_t *__restrict__ bins, uint32_t *__restrict__ counter,
            uint32_t *__restrict__ status)
{
    constexpr int TILE = RS_THREADS * ITEMS;
    __shared__ K s_keys[TILE];
    __shared__ uint32_t s_vals[TILE];
    __shared__ uint32_t s_whist[RS_WARPS][256]; // 8 KB
    __shared__ uint32_t s_dstart[256];
    __shared__ uint32_t s_goff[256];
    __shared__ uint32_t s_w[RS_WARPS];
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x;
    if (tid == 0) s_tile = atomicAdd(counter, 1u); // ticket: predecessors are always resident
#pragma unroll
    for (int k = 0; k < RS_WARPS; k++) s_whist[k][tid] = 0;
    // (the ticket counter and status words of this pass were cleared before the sort's first kernel: nothing above
    // depends on the previous pass, whose keys are read from here on)
    pdl_wait();
    pdl_trigger();
    __syncthreads();
    const uint32_t tile = s_tile;
    const size_t base = (size_t)tile * TILE;
    const uint32_t nt = (uint32_t)min((size_t)TILE, n - base);
    const uint32_t mask = (1u << bits) - 1u;
    if (nt == (uint32_t)TILE)
        rs_tile_body<K, ITEMS, true>(keys_in + base, vals_in + base, keys_out, vals_out, nt, tile, shift, mask, bins,
                                     status, s_keys, s_vals, s_whist, s_dstart, s_goff, s_w);
    else
        rs_tile_body<K, ITEMS, false>(keys_in + base, vals_in + base, keys_out, vals_out, nt, tile, shift, mask, bins,
                                      status, s_keys, s_vals, s_whist, s_dstart, s_goff, s_w);
}

// hist + counters + status words are one contiguous carve (SortScratch::from): one memset clears them
size_t sort_scratch_words(size_t n, int end_bit)
{
    const int npass = (end_bit + 7) / 8;
    return (size_t)RS_MAX_PASSES * 256 + 64 + (size_t)npass * sort_tiles(n) * 256;
}
int sort_clear_scratch(size_t n, int end_bit, SortScratch &ss, cudaStream_t st)
{
    WS_CUDA_CHECK(cudaMemsetAsync(ss.hist, 0, sort_scratch_words(n, end_bit) * sizeof(uint32_t), st));
    return WS_OK;
}

// have_hist: the caller cleared the scratch (sort_clear_scratch) and accumulated the per-pass digit histograms into
// ss.hist itself (emit_tiles_ordered does, while it writes the keys): no histogram pass over the keys here
template <typename K>
static int sort_pairs_impl(size_t n, K *keys_in, K *keys_out, uint32_t *vals_in,
                           uint32_t *vals_out, int end_bit, SortScratch &ss, cudaStream_t st, bool debug,
                           K **keys_final, uint32_t **vals_final, bool have_hist = false)
{
    *keys_final = keys_in;
    *vals_final = vals_in;
    if (n == 0) return WS_OK;
    if (n >= (size_t)RS_VAL_MASK) {
        set_error("radix sort: %zu pairs exceed the 2^30-1 limit of the look-back status words", n);
        return WS_E_UNSUPPORTED;
    }
    if (end_bit < 1 || end_bit > (int)sizeof(K) * 8) {
        set_error("radix sort: end_bit %d out of range", end_bit);
        return WS_E_INVALID;
    }
    const int npass = (end_bit + 7) / 8;
    constexpr int ITEMS = items_of<K>();
    const size_t nt = (n + (size_t)RS_THREADS * ITEMS - 1) / ((size_t)RS_THREADS * ITEMS);
    if (!have_hist) {
        int rc = sort_clear_scratch(n, end_bit, ss, st);
        if (rc != WS_OK) return rc;
        int hb = (int)((n + 256 * 16 - 1) / (256 * 16));
        if (hb > 148 * 8) hb = 148 * 8;
        rs_histogram<K><<<hb, 256, 0, st>>>(keys_in, n, end_bit, ss.hist);
        WS_LAUNCH_CHECK(st, debug);
    }
    K *ki = keys_in, *ko = keys_out;
    uint32_t *vi = vals_in, *vo = vals_out;
    for (int p = 0; p < npass; p++) {
        const int w = sort_digit_bits(end_bit), bits = end_bit - w * p < w ? end_bit - w * p : w;
        WS_CUDA_CHECK(launch_pdl(rs_onesweep<K, ITEMS>, dim3((unsigned)nt), dim3(RS_THREADS), 0, st, ki, ko, vi, vo, n,
                                 w * p, bits, ss.hist + p * 256, ss.counters + p, ss.status + (size_t)p * nt * 256));
        WS_LAUNCH_CHECK(st, debug);
        K *tk = ki; ki = ko; ko = tk;
        uint32_t *tv = vi; vi = vo; vo = tv;
    }
    *keys_final = ki;
    *vals_final = vi;
    return WS_OK;
}

int launch_sort_pairs(size_t n, uint64_t *keys_in, uint64_t *keys_out, uint32_t *vals_in,
                      uint32_t *vals_out, int end_bit, SortScratch &ss, cudaStream_t st, bool debug,
                      uint64_t **keys_final, uint32_t **vals_final)
{
    return sort_pairs_impl<uint64_t>(n, keys_in, keys_out, vals_in, vals_out, end_bit, ss, st, debug, keys_final,
                                     vals_final);
}

int launch_sort_pairs32(size_t n, uint32_t *keys_in, uint32_t *keys_out, uint32_t *vals_in,
                        uint32_t *vals_out, int end_bit, SortScratch &ss, cudaStream_t st, bool debug,
                        uint32_t **keys_final, uint32_t **vals_final, bool have_hist)
{
    return sort_pairs_impl<uint32_t>(n, keys_in, keys_out, vals_in, vals_out, end_bit, ss, st, debug, keys_final,
                                     vals_final, have_hist);
}

} // namespace ws

extern "C" size_t ws_sort_bytes(int64_t n)
{
    char *p = nullptr;
    ws::SortScratch::from(p, (size_t)(n > 0 ? n : 0));
    return (size_t)p + 512;
}

extern "C" int ws_sort_pairs_u64(int64_t n, uint64_t *keys_in, uint64_t *keys_out, uint32_t *vals_in,
                                 uint32_t *vals_out, int32_t end_bit, void *scratch, size_t scratch_bytes,
                                 ws_stream_t stream)
{
    if (n < 0 || (n > 0 && (!keys_in || !keys_out || !vals_in || !vals_out || !scratch))) {
        ws::set_error("ws_sort_pairs_u64: bad arguments");
        return WS_E_INVALID;
    }
    if (scratch_bytes < ws_sort_bytes(n)) {
        ws::set_error("ws_sort_pairs_u64: scratch too small (%zu < %zu)", scratch_bytes, ws_sort_bytes(n));
        return WS_E_SCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    char *p = (char *)scratch;
    ws::SortScratch ss = ws::SortScratch::from(p, (size_t)n);
    uint64_t *kf;
    uint32_t *vf;
    int rc = ws::launch_sort_pairs((size_t)n, keys_in, keys_out, vals_in, vals_out, end_bit, ss, st, false, &kf, &vf);
    if (rc != WS_OK) return rc;
    if (n > 0 && kf != keys_out) {
        WS_CUDA_CHECK(cudaMemcpyAsync(keys_out, kf, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
        WS_CUDA_CHECK(cudaMemcpyAsync(vals_out, vf, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    }
    return WS_OK;
}

extern "C" int ws_sort_pairs_u32(int64_t n, uint32_t *keys_in, uint32_t *keys_out, uint32_t *vals_in,
                                 uint32_t *vals_out, int32_t end_bit, void *scratch, size_t scratch_bytes,
                                 ws_stream_t stream)
{
    if (n < 0 || (n > 0 && (!keys_in || !keys_out || !vals_in || !vals_out || !scratch))) {
        ws::set_error("ws_sort_pairs_u32: bad arguments");
        return WS_E_INVALID;
    }
    if (scratch_bytes < ws_sort_bytes(n)) {
        ws::set_error("ws_sort_pairs_u32: scratch too small (%zu < %zu)", scratch_bytes, ws_sort_bytes(n));
        return WS_E_SCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    char *p = (char *)scratch;
    ws::SortScratch ss = ws::SortScratch::from(p, (size_t)n);
    uint32_t *kf, *vf;
    int rc = ws::launch_sort_pairs32((size_t)n, keys_in, keys_out, vals_in, vals_out, end_bit, ss, st, false, &kf, &vf);
    if (rc != WS_OK) return rc;
    if (n > 0 && kf != keys_out) {
        WS_CUDA_CHECK(cudaMemcpyAsync(keys_out, kf, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
        WS_CUDA_CHECK(cudaMemcpyAsync(vals_out, vf, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    }
    return WS_OK;
}
