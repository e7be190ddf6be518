// ws_api.cu -- the C-ABI entry points declared in include/wheatsplat.h: argument checks,
// scratch carving and stage sequencing.  No device memory is allocated here.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <thread>
#include <vector>
#include "ws_common.cuh"

namespace ws {

static thread_local char g_err[512] = "";

// WS_BINNING=keys64 selects the reference's own scheme (64-bit tile|depth keys, one sort over all
// instances); the default is depth-first binning (ws_preprocess.cu).  Both give the same list.
static int g_binning = -1; // -1: from the environment; 0: depth-first; 1: keys64
static bool binning_keys64()
{
    if (g_binning >= 0) return g_binning == 1;
    static const bool v = [] {
        const char *e = getenv("WS_BINNING");
        return e && !strcmp(e, "keys64");
    }();
    return v;
}

long long g_launch_count = 0;

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---- stage timing (ws_prof_*): CUDA events on the caller's stream around every stage ----
static bool g_prof = false;
struct ProfRec { int stage; cudaEvent_t a, b; };
static std::vector<ProfRec> g_prof_recs;
static std::vector<cudaEvent_t> g_prof_free;
static std::mutex g_prof_mu;

static cudaEvent_t prof_event()
{
    cudaEvent_t e = nullptr;
    if (!g_prof_free.empty()) { e = g_prof_free.back(); g_prof_free.pop_back(); }
    else cudaEventCreate(&e);
    return e;
}

struct ProfScope {
    int stage; cudaStream_t st; cudaEvent_t a = nullptr;
    ProfScope(int stage_, cudaStream_t st_) : stage(stage_), st(st_)
    {
        if (!g_prof) return;
        std::lock_guard<std::mutex> lk(g_prof_mu);
        a = prof_event();
        cudaEventRecord(a, st);
    }
    ~ProfScope()
    {
        if (!a) return;
        std::lock_guard<std::mutex> lk(g_prof_mu);
        cudaEvent_t b = prof_event();
        cudaEventRecord(b, st);
        g_prof_recs.push_back({stage, a, b});
    }
};

cudaEvent_t total_event()
{
    static thread_local cudaEvent_t ev[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    dev &= 15;
    if (!ev[dev]) cudaEventCreateWithFlags(&ev[dev], cudaEventDisableTiming);
    return ev[dev];
}

// pinned landing words for device -> host counts (a D2H copy into pageable memory blocks the host until the stream
// has drained, a copy into pinned memory is asynchronous).  Mapped into the device's address space as well (same
// pointer under unified addressing): words 8..9 are the mailbox the scan kernel writes num_rendered into directly.
uint32_t *total_host()
{
    static thread_local uint32_t *p = nullptr;
    if (!p) {
        if (cudaHostAlloc((void **)&p, 64, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) p = nullptr;
        else memset(p, 0, 64);
    }
    return p;
}

// per-thread, per-device non-blocking stream for the mailbox store
cudaStream_t side_stream()
{
    static thread_local cudaStream_t ss[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    dev &= 15;
    if (!ss[dev] && cudaStreamCreateWithFlags(&ss[dev], cudaStreamNonBlocking) != cudaSuccess) ss[dev] = nullptr;
    return ss[dev];
}

// DGR rasterizer_impl.cu:35-50
uint32_t get_higher_msb(uint32_t n)
{
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

// the sorted pairs end in the "a" buffers after an even number of radix passes, else in "b"
const uint32_t *sorted_vals(const BinningState &bs, int T)
{
    const int end_bit = (binning_keys64() ? 32 : 0) + (int)get_higher_msb((uint32_t)T);
    return (((end_bit + 7) / 8) & 1) ? bs.vals_b : bs.vals_a;
}

int check_view_gauss(const char *who, const ws_view *v, const ws_gaussians *g, bool need_opacity)
{
    if (!v || !g) { set_error("%s: NULL view/gaussians", who); return WS_E_INVALID; }
    if (g->P <= 0) { set_error("%s: P must be > 0 (the wrapper handles empty inputs)", who); return WS_E_INVALID; }
    if (v->image_width <= 0 || v->image_height <= 0) { set_error("%s: bad image size", who); return WS_E_INVALID; }
    if (!v->bg || !v->viewmatrix || !v->projmatrix || !v->campos) { set_error("%s: NULL view tensor", who); return WS_E_INVALID; }
    if (!g->means3D || (need_opacity && !g->opacities)) { set_error("%s: means3D and opacities are required", who); return WS_E_INVALID; }
    const bool has_sh = g->shs != nullptr && g->M > 0, has_col = g->colors_precomp != nullptr;
    if (has_sh == has_col) { // DGR rasterizer_impl.cu:243-246 / __init__.py:194-195
        set_error("%s: provide exactly one of SHs or precomputed colors", who);
        return WS_E_INVALID;
    }
    const bool has_sr = g->scales != nullptr && g->rotations != nullptr, has_cov = g->cov3D_precomp != nullptr;
    if (has_sr == has_cov || ((g->scales != nullptr) != (g->rotations != nullptr))) {
        set_error("%s: provide exactly one of scale/rotation pair or precomputed 3D covariance", who);
        return WS_E_INVALID;
    }
    if (g->shs_rest && (!has_sh || g->M < 2)) {
        set_error("%s: shs_rest needs shs (features_dc) and M >= 2 coefficients in total", who);
        return WS_E_INVALID;
    }
    if (v->sh_degree < 0 || v->sh_degree > 3) { set_error("%s: sh_degree must be 0..3", who); return WS_E_INVALID; }
    if (has_sh && (v->sh_degree + 1) * (v->sh_degree + 1) > g->M) {
        set_error("%s: sh_degree %d needs %d coefficients, sh has %d", who, v->sh_degree,
                  (v->sh_degree + 1) * (v->sh_degree + 1), g->M);
        return WS_E_INVALID;
    }
    return WS_OK;
}

static int check_flash(const char *who, const ws_flash *fl, bool forward)
{
    if (!fl) return WS_OK;
    if (fl->num_obj < 0 || !fl->gt_mask) { set_error("%s: flash needs gt_mask and num_obj >= 0", who); return WS_E_INVALID; }
    if (forward && (!fl->out_contrib_num || !fl->used_count)) {
        set_error("%s: flash forward needs out_contrib_num and used_count", who);
        return WS_E_INVALID;
    }
    return WS_OK;
}

} // namespace ws

using namespace ws;

extern "C" int ws_abi_version(void) { return WS_ABI_VERSION; }
extern "C" const char *ws_last_error(void) { return ws::g_err; }
extern "C" int64_t ws_launch_count(void) { return ws::g_launch_count; }

extern "C" int ws_set_binning_mode(int32_t mode)
{
    if (mode < -1 || mode > 1) { set_error("ws_set_binning_mode: mode must be -1, 0 or 1"); return WS_E_INVALID; }
    ws::g_binning = mode;
    return WS_OK;
}
extern "C" void ws_prof_enable(int32_t on) { ws::g_prof = on != 0; }
extern "C" const char *ws_prof_stage_name(int32_t stage)
{
    static const char *names[WS_PROF_NSTAGES] = {"preprocess", "scan", "emit_keys", "sort", "ranges_pack",
                                                 "render_fwd", "render_bwd", "preprocess_bwd"};
    return (stage >= 0 && stage < WS_PROF_NSTAGES) ? names[stage] : "?";
}
extern "C" int ws_prof_read(double *ms, int64_t *calls)
{
    if (!ms || !calls) { set_error("ws_prof_read: NULL"); return WS_E_INVALID; }
    std::lock_guard<std::mutex> lk(ws::g_prof_mu);
    for (auto &r : ws::g_prof_recs) {
        float t = 0.f;
        WS_CUDA_CHECK(cudaEventSynchronize(r.b));
        WS_CUDA_CHECK(cudaEventElapsedTime(&t, r.a, r.b));
        ms[r.stage] += (double)t;
        calls[r.stage] += 1;
        ws::g_prof_free.push_back(r.a);
        ws::g_prof_free.push_back(r.b);
    }
    ws::g_prof_recs.clear();
    return WS_OK;
}

extern "C" size_t ws_geom_bytes(int32_t P)
{
    size_t b = 0;
    GeomState::from(nullptr, (size_t)(P > 0 ? P : 0), &b);
    return b + 256;
}
extern "C" size_t ws_image_bytes(int32_t width, int32_t height)
{
    size_t b = 0;
    const size_t T = (size_t)((width + WS_BLOCK_X - 1) / WS_BLOCK_X) * ((height + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    ImageState::from(nullptr, T, (size_t)width * height, &b);
    return b + 256;
}
extern "C" size_t ws_binning_bytes(int64_t num_rendered, int32_t width, int32_t height)
{
    (void)width; (void)height;
    size_t b = 0;
    BinningState::from(nullptr, (size_t)(num_rendered > 0 ? num_rendered : 0), &b);
    return b + 256;
}

// One turn of the mailbox spin below: a pause hint for the sibling hyper-thread while the wait is short (the count
// normally arrives within ~100 us), the scheduler's yield once it is not (a long view, a debugger, a contended host),
// so a late kernel does not pin a host core.  Builds on x86-64 and aarch64 (Grace) hosts.
static inline void cpu_relax(uint32_t spins)
{
    if (spins > (1u << 16)) { std::this_thread::yield(); return; }
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#elif defined(__aarch64__)
    asm volatile("yield" ::: "memory");
#else
    std::this_thread::yield();
#endif
}

extern "C" int64_t ws_forward_prepare(const ws_view *view, const ws_gaussians *g, int32_t *radii, void *geom,
                                      size_t geom_bytes, const ws_flash *fl, ws_stream_t stream)
{
    int rc = check_view_gauss("ws_forward_prepare", view, g);
    if (rc != WS_OK) return rc;
    if (!radii || !geom) { set_error("ws_forward_prepare: NULL radii/geom"); return WS_E_INVALID; }
    if (geom_bytes < ws_geom_bytes(g->P)) {
        set_error("ws_forward_prepare: geom scratch too small (%zu < %zu)", geom_bytes, ws_geom_bytes(g->P));
        return WS_E_SCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    GeomState gs = GeomState::from((char *)geom, (size_t)g->P);
    { ProfScope ps(0, st); rc = launch_preprocess(view, g, radii, gs, fl, st); }
    if (rc != WS_OK) return rc;
    uint32_t *th = total_host();
    if (!th) { set_error("ws_forward_prepare: cudaHostAlloc failed"); return WS_E_CUDA; }
    // num_rendered reaches the host through a mailbox in mapped pinned memory: a one-thread kernel on a side stream,
    // behind the scan, stores {sequence number, total} as one 8-byte word and the host spins on it.  A D2H copy in
    // the main stream held back everything queued behind it (25 us on the copy engine next to a 38 MB input upload,
    // plus the engine hand-over); this way the depth sort, which needs nothing from the host, starts right behind
    // the scan and runs while the host learns R, sizes the binning scratch and queues the rest of the forward (the
    // reference idles the GPU there, rasterizer_impl.cu:282).
    static thread_local uint32_t seq_counter = 0;
    const uint32_t seq = ++seq_counter ? seq_counter : ++seq_counter;
    volatile unsigned long long *mb = (volatile unsigned long long *)(th + 8);
    { ProfScope ps(1, st); rc = launch_scan_blocks(g->P, gs, st, view->debug != 0); }
    if (rc != WS_OK) return rc;
    cudaEvent_t ev = total_event();
    cudaStream_t side = side_stream();
    if (!side) { set_error("ws_forward_prepare: cudaStreamCreate failed"); return WS_E_CUDA; }
    WS_CUDA_CHECK(cudaEventRecord(ev, st));
    WS_CUDA_CHECK(cudaStreamWaitEvent(side, ev, 0));
    rc = launch_post_total(gs.total, (unsigned long long *)(th + 8), seq, side);
    if (rc != WS_OK) return rc;
    WS_CUDA_CHECK(cudaEventRecord(ev, side));
    if (!binning_keys64()) {
        { ProfScope ps(3, st); rc = launch_depth_sort(g->P, gs, st, view->debug != 0); }
        if (rc != WS_OK) return rc;
    }
    unsigned long long word;
    for (uint32_t spins = 1;; spins++) {
        word = *mb;
        if ((uint32_t)(word >> 32) == seq) break;
        cpu_relax(spins);
        if ((spins & 1023u) == 0) {   // the kernel is gone (finished without our sequence number, or failed)?
            cudaError_t q = cudaEventQuery(ev);
            if (q == cudaSuccess) {
                word = *mb;
                if ((uint32_t)(word >> 32) == seq) break;
                set_error("ws_forward_prepare: the scan finished without delivering num_rendered");
                return WS_E_CUDA;
            }
            if (q != cudaErrorNotReady) {
                set_error("ws_forward_prepare: %s while waiting for num_rendered", cudaGetErrorString(q));
                return WS_E_CUDA;
            }
        }
    }
    return (int64_t)(uint32_t)word;
}

extern "C" int ws_forward_render(const ws_view *view, const ws_gaussians *g, const int32_t *radii, void *geom,
                                 void *binning, size_t binning_bytes, void *image, size_t image_bytes,
                                 int64_t num_rendered, float *out_color, float *out_depth, float *out_alpha,
                                 const ws_flash *fl, ws_stream_t stream)
{
    int rc = check_view_gauss("ws_forward_render", view, g);
    if (rc != WS_OK) return rc;
    rc = check_flash("ws_forward_render", fl, true);
    if (rc != WS_OK) return rc;
    if (!radii || !geom || !image || !out_color || !out_depth || !out_alpha || num_rendered < 0 ||
        (num_rendered > 0 && !binning)) {
        set_error("ws_forward_render: NULL buffer");
        return WS_E_INVALID;
    }
    const int W = view->image_width, H = view->image_height;
    if (image_bytes < ws_image_bytes(W, H)) { set_error("ws_forward_render: image scratch too small"); return WS_E_SCRATCH; }
    if (num_rendered > 0 && binning_bytes < ws_binning_bytes(num_rendered, W, H)) {
        set_error("ws_forward_render: binning scratch too small (%zu < %zu)", binning_bytes,
                  ws_binning_bytes(num_rendered, W, H));
        return WS_E_SCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const bool debug = view->debug != 0;
    const int gx = (W + WS_BLOCK_X - 1) / WS_BLOCK_X, gy = (H + WS_BLOCK_Y - 1) / WS_BLOCK_Y;
    const int T = gx * gy;
    const size_t R = (size_t)num_rendered;
    GeomState gs = GeomState::from((char *)geom, (size_t)g->P);
    ImageState is = ImageState::from((char *)image, (size_t)T, (size_t)W * H);
    BinningState bs = BinningState::from((char *)binning, R);
    uint64_t *keys = nullptr;
    uint32_t *tkeys = nullptr, *vals = nullptr;
    bool ranges_cleared = false;
    if (R > 0 && binning_keys64()) {
        { ProfScope ps(2, st); rc = launch_emit_keys(g->P, W, H, radii, gs, bs.keys_a, bs.vals_a, st, debug); }
        if (rc != WS_OK) return rc;
        const int end_bit = 32 + (int)get_higher_msb((uint32_t)T); // rasterizer_impl.cu:301-309
        { ProfScope ps(3, st); rc = launch_sort_pairs(R, bs.keys_a, bs.keys_b, bs.vals_a, bs.vals_b, end_bit, bs.sort, st, debug, &keys, &vals); }
        if (rc != WS_OK) return rc;
    } else if (R > 0) {
        // the emit pass counts the tile keys' digits while it writes them: the sort needs no pass of its own for that
        const int end_bit = (int)get_higher_msb((uint32_t)T);
        // (the stage's first kernel also clears the sort's scratch words and the tile ranges)
        { ProfScope ps(2, st); rc = launch_emit_tiles_ordered(g->P, W, H, radii, gs, bs.tkeys_a, bs.vals_a, st, debug, bs.sort.hist, (end_bit + 7) / 8, sort_digit_bits(end_bit),
                                                              bs.sort.hist, sort_scratch_words(R, end_bit), (uint32_t *)is.ranges, (size_t)T * 2); }
        ranges_cleared = true;
        if (rc != WS_OK) return rc;
        { ProfScope ps(3, st); rc = launch_sort_pairs32(R, bs.tkeys_a, bs.tkeys_b, bs.vals_a, bs.vals_b, end_bit, bs.sort, st, debug, &tkeys, &vals, true); }
        if (rc != WS_OK) return rc;
    }
    { ProfScope ps(4, st); rc = launch_identify_ranges(R, T, keys, tkeys, is.ranges, st, debug, ranges_cleared); }
    if (rc != WS_OK) return rc;
    { ProfScope ps(5, st); rc = launch_render_fwd(view, g->P, gs, vals, is, out_color, out_depth, out_alpha, fl, st); }
    return rc;
}

extern "C" int ws_backward(const ws_view *view, const ws_gaussians *g, const int32_t *radii, void *geom,
                           void *binning, void *image, int64_t num_rendered, const float *out_alpha,
                           const float *dL_dcolor, const float *dL_ddepth, const float *dL_dalpha,
                           const ws_grads *gr, const ws_flash *fl, ws_stream_t stream)
{
    int rc = check_view_gauss("ws_backward", view, g, false); // opacity lives in the forward's geometry state
    if (rc != WS_OK) return rc;
    rc = check_flash("ws_backward", fl, false);
    if (rc != WS_OK) return rc;
    if (!radii || !geom || !image || !out_alpha || !dL_dcolor || !dL_ddepth || !dL_dalpha || !gr ||
        (num_rendered > 0 && !binning)) {
        set_error("ws_backward: NULL buffer");
        return WS_E_INVALID;
    }
    if (!gr->dL_dmeans2D || !gr->dL_dcolors || !gr->dL_dopacity || !gr->dL_dmeans3D || !gr->dL_dcov3D ||
        (g->shs && g->M > 0 && !gr->dL_dsh) || (g->scales && (!gr->dL_dscales || !gr->dL_drotations))) {
        set_error("ws_backward: NULL gradient tensor");
        return WS_E_INVALID;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int W = view->image_width, H = view->image_height;
    const int T = ((W + WS_BLOCK_X - 1) / WS_BLOCK_X) * ((H + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    GeomState gs = GeomState::from((char *)geom, (size_t)g->P);
    ImageState is = ImageState::from((char *)image, (size_t)T, (size_t)W * H);
    BinningState bs = BinningState::from((char *)binning, (size_t)num_rendered);
    { ProfScope ps(6, st); rc = launch_render_bwd(view, g->P, gs, sorted_vals(bs, T), is, out_alpha, dL_dcolor, dL_ddepth, dL_dalpha, fl, st); }
    if (rc != WS_OK) return rc;
    { ProfScope ps(7, st); rc = launch_preprocess_bwd(view, g, radii, gs, gr, st); }
    return rc;
}

namespace ws {
__global__ void unpack_geom(int P, int what, const GRec *__restrict__ grec, float *__restrict__ dst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float4 a = grec[i].a, b = grec[i].b, c = grec[i].c;
    if (what == 0) dst[i] = b.z;                                              // depths
    else if (what == 1) { dst[2 * i] = a.x; dst[2 * i + 1] = a.y; }          // means2D
    else if (what == 2) { dst[4 * i] = a.z; dst[4 * i + 1] = a.w; dst[4 * i + 2] = b.x; dst[4 * i + 3] = b.y; }
    else { dst[3 * i] = c.x; dst[3 * i + 1] = c.y; dst[3 * i + 2] = c.z; }   // rgb
}
} // namespace ws

namespace ws {
__global__ void rebuild_keys64(size_t R, const uint32_t *__restrict__ tkeys, const uint32_t *__restrict__ vals,
                               const GRec *__restrict__ grec, uint64_t *__restrict__ keys)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < R) keys[i] = ((uint64_t)tkeys[i] << 32) | __float_as_uint(grec[vals[i] & WS_ID_MASK].b.z);
}
// point_list as the reference holds it: the Gaussian ids without the quadrant masks
__global__ void strip_masks(size_t R, const uint32_t *__restrict__ vals, uint32_t *__restrict__ dst)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < R) dst[i] = vals[i] & WS_ID_MASK;
}
// one thread per pixel: re-walks its tile's list up to the pixel's last contributor.  counts[0] = pairs walked,
// [1] = contributing pairs, [2] = (warp, record) pairs the 8x8 quadrant masks let through, [3] = instances with an
// empty mask (never fetched by the blend kernels)
__global__ void __launch_bounds__(256) count_pairs(int W, int H, const uint2 *__restrict__ ranges,
                                                   const uint32_t *__restrict__ n_contrib,
                                                   const uint32_t *__restrict__ vals, const GRec *__restrict__ grec,
                                                   unsigned long long *__restrict__ counts)
{
    const int px = blockIdx.x * WS_BLOCK_X + (threadIdx.x & 15), py = blockIdx.y * WS_BLOCK_Y + (threadIdx.x >> 4);
    unsigned long long ne = 0, nc = 0, nq = 0, nz = 0;
    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    if (px < W && py < H) {
        const uint32_t last = n_contrib[(size_t)py * W + px];
        ne = last;
        for (uint32_t k = 0; k < last; k++) {
            const GRec *rec = grec + (vals[range.x + k] & WS_ID_MASK);
            const float4 a = rec->a, q = rec->b;
            const float dx = a.x - (float)px, dy = a.y - (float)py;
            const float power = -0.5f * (a.z * dx * dx + q.x * dy * dy) - a.w * dx * dy;
            if (power > 0.0f) continue;
            if (fminf(0.99f, q.y * expf(power)) < 1.0f / 255.0f) continue;
            nc++;
        }
    }
    for (uint32_t k = threadIdx.x; k < range.y - range.x; k += 256) {
        const uint32_t m = vals[range.x + k] >> WS_ID_BITS;
        nq += __popc(m);
        nz += m == 0u;
    }
    for (int o = 16; o > 0; o >>= 1) {
        ne += __shfl_xor_sync(0xffffffffu, ne, o);
        nc += __shfl_xor_sync(0xffffffffu, nc, o);
        nq += __shfl_xor_sync(0xffffffffu, nq, o);
        nz += __shfl_xor_sync(0xffffffffu, nz, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(counts, ne); atomicAdd(counts + 1, nc); atomicAdd(counts + 2, nq); atomicAdd(counts + 3, nz);
    }
}
} // namespace ws

namespace ws {
// FP32 issue-rate probe for the blend kernels' roofline (bench.py): every thread runs 8 independent dependent-FMA
// chains, three distinct register operands per FMA as in the blend loops.  PACKED: fma.rn.f32x2 (FFMA2, what the
// blend kernels issue); otherwise scalar FFMA.
template <bool PACKED>
__global__ void __launch_bounds__(256) fma_peak(int iters, float b0, float c0, float *__restrict__ out)
{
    const float t = (float)threadIdx.x * 1e-3f;
    if (PACKED) {
        f2 a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = pk2(t + (float)k, t - (float)k);
        const f2 b = pk2(b0, b0 * 0.999f), c = pk2(c0, c0 * 1.001f);
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int u = 0; u < 4; u++) {
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = fma2(a[k], b, c);
            }
        }
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 8; k++) s += lo2(a[k]) + hi2(a[k]);
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    } else {
        float a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = t + (float)k;
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int u = 0; u < 4; u++) {
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = __fmaf_rn(a[k], b0, c0);
            }
        }
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 8; k++) s += a[k];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    }
}
} // namespace ws

extern "C" int64_t ws_debug_fma_peak(int32_t packed, int32_t iters, int32_t ctas, float *out, ws_stream_t stream)
{
    if (iters < 1 || ctas < 1 || !out) { set_error("ws_debug_fma_peak: bad arguments"); return WS_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    if (packed) ws::fma_peak<true><<<ctas, 256, 0, st>>>(iters, 0.9999f, 1e-4f, out);
    else ws::fma_peak<false><<<ctas, 256, 0, st>>>(iters, 0.9999f, 1e-4f, out);
    WS_LAUNCH_CHECK(st, false);
    // flops issued: threads x iters x 32 FMAs x 2 (x 2 when packed)
    return (int64_t)ctas * 256 * (int64_t)iters * 32 * 2 * (packed ? 2 : 1);
}

extern "C" int ws_debug_count_pairs(int32_t P, int32_t width, int32_t height, void *geom, void *binning, void *image,
                                    int64_t num_rendered, uint64_t *counts, ws_stream_t stream)
{
    if (!geom || !image || !counts || P <= 0 || width <= 0 || height <= 0 || (num_rendered > 0 && !binning)) {
        set_error("ws_debug_count_pairs: bad arguments");
        return WS_E_INVALID;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int gx = (width + WS_BLOCK_X - 1) / WS_BLOCK_X, gy = (height + WS_BLOCK_Y - 1) / WS_BLOCK_Y;
    ImageState is = ImageState::from((char *)image, (size_t)gx * gy, (size_t)width * height);
    BinningState bs = BinningState::from((char *)binning, (size_t)num_rendered);
    GeomState gs = GeomState::from((char *)geom, (size_t)P);
    WS_CUDA_CHECK(cudaMemsetAsync(counts, 0, 32, st));
    ws::count_pairs<<<dim3(gx, gy), 256, 0, st>>>(width, height, is.ranges, is.n_contrib, ws::sorted_vals(bs, gx * gy),
                                                 gs.grec, (unsigned long long *)counts);
    WS_LAUNCH_CHECK(st, false);
    return WS_OK;
}

extern "C" int64_t ws_debug_fetch(const char *name, int32_t P, int32_t width, int32_t height, int64_t num_rendered,
                                  void *geom, void *binning, void *image, void *dst, size_t dst_bytes,
                                  ws_stream_t stream)
{
    if (!name || !dst) { set_error("ws_debug_fetch: NULL"); return WS_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    const int T = ((width + WS_BLOCK_X - 1) / WS_BLOCK_X) * ((height + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    const size_t R = (size_t)num_rendered;
    GeomState gs = GeomState::from((char *)geom, (size_t)P);
    int what = -1;
    size_t need = 0;
    if (!strcmp(name, "depths")) { what = 0; need = 4 * (size_t)P; }
    else if (!strcmp(name, "means2D")) { what = 1; need = 8 * (size_t)P; }
    else if (!strcmp(name, "conic_opacity")) { what = 2; need = 16 * (size_t)P; }
    else if (!strcmp(name, "rgb")) { what = 3; need = 12 * (size_t)P; }
    if (what >= 0) {
        // only rows of visible Gaussians are meaningful (the reference leaves the rest stale too)
        if (dst_bytes < need) { set_error("ws_debug_fetch: dst too small"); return WS_E_SCRATCH; }
        WS_CUDA_CHECK(cudaMemsetAsync(dst, 0, need, st));
        ws::unpack_geom<<<(P + 255) / 256, 256, 0, st>>>(P, what, gs.grec, (float *)dst);
        WS_LAUNCH_CHECK(st, false);
        return (int64_t)need;
    }
    const void *src = nullptr;
    if (!strcmp(name, "tiles_touched")) { src = gs.tiles_touched; need = 4 * (size_t)P; }
    else if (!strcmp(name, "ranges") || !strcmp(name, "n_contrib")) {
        ImageState is = ImageState::from((char *)image, (size_t)T, (size_t)width * height);
        if (!strcmp(name, "ranges")) { src = is.ranges; need = 8 * (size_t)T; }
        else { src = is.n_contrib; need = 4 * (size_t)width * height; }
    } else if (!strcmp(name, "point_list_keys") || !strcmp(name, "point_list")) {
        if (R == 0) return 0;
        BinningState bs = BinningState::from((char *)binning, R);
        // the sorted pairs end in the "a" buffers after an even number of passes, else in "b"
        const bool k64 = ws::binning_keys64();
        const int end_bit = (k64 ? 32 : 0) + (int)ws::get_higher_msb((uint32_t)T);
        const int npass = (end_bit + 7) / 8;
        const bool in_b = (npass & 1) != 0;
        if (!strcmp(name, "point_list")) {
            need = 4 * R;
            if (dst_bytes < need) { set_error("ws_debug_fetch: dst too small"); return WS_E_SCRATCH; }
            ws::strip_masks<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(R, in_b ? bs.vals_b : bs.vals_a, (uint32_t *)dst);
            WS_LAUNCH_CHECK(st, false);
            return (int64_t)need;
        }
        else if (k64) { src = in_b ? bs.keys_b : bs.keys_a; need = 8 * R; }
        else {
            // depth-first binning sorts tile ids only: rebuild the reference's tile|depth keys
            need = 8 * R;
            if (dst_bytes < need) { set_error("ws_debug_fetch: dst too small"); return WS_E_SCRATCH; }
            ws::rebuild_keys64<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(
                R, in_b ? bs.tkeys_b : bs.tkeys_a, in_b ? bs.vals_b : bs.vals_a, gs.grec, (uint64_t *)dst);
            WS_LAUNCH_CHECK(st, false);
            return (int64_t)need;
        }
    } else {
        set_error("ws_debug_fetch: unknown array '%s'", name);
        return WS_E_INVALID;
    }
    if (dst_bytes < need) { set_error("ws_debug_fetch: dst too small"); return WS_E_SCRATCH; }
    WS_CUDA_CHECK(cudaMemcpyAsync(dst, src, need, cudaMemcpyDeviceToDevice, st));
    return (int64_t)need;
}
