// ws_objeval.cu -- batched eval_obj_labels (replaces utils/wheatgs_helper.py:245-271 =
// eval_wheatgs.py:29-64, called per orbit frame by render_360 (utils/wheatgs_helper.py:298) and per view by
// eval_wheatgs.py:76).
//
// The reference renders every object on its own -- one flashsplat_render over the Gaussians of row i of
// all_obj_labels (gaussian_renderer/__init__.py:154-187), alpha and depth copied to the CPU -- and paints the
// objects into pred_mask in ascending order: pixels with alpha > 0.5 take the object's id, except that an
// object that lies BEHIND what is already painted (mean painted depth < mean object depth over the overlap)
// only takes free pixels.  ~500 rasterizer calls and ~1000 image-sized D2H copies per frame.
//
// Here, per frame:
//   1. ONE preprocess of all P Gaussians (per-Gaussian results do not depend on the subset).
//   2. ONE binning pass whose key is (object, tile): a Gaussian that belongs to k objects (CSR lists
//      member_start/member_obj; k = 0 drops it) emits k x tiles_touched instances, in depth order, and the
//      stable sort by object*T + tile leaves each (object, tile) list in the order the subset render would see
//      (depth, then Gaussian id).  ranges has one slot per (object, tile); the non-empty slots become JOBS.
//   3. ONE launch of the forward blend kernel over the jobs (render_fwd<2>: same arithmetic, alpha and depth of
//      the job's 16x16 pixels), so every object's alpha/depth is bit-identical to its own subset render.
//   4. ONE merge kernel that walks the objects in ascending order over their jobs' pixels only.
#include "ws_common.cuh"

namespace ws {

// effective tile counts (tiles_touched x memberships) summed over blocks of 256 Gaussians, optionally taken
// in depth order
__global__ void __launch_bounds__(WS_PRE_THREADS) objeval_block_sums(int P, const uint32_t *__restrict__ order,
                                                                     const uint32_t *__restrict__ tiles_touched,
                                                                     const int32_t *__restrict__ member_start,
                                                                     uint32_t *__restrict__ block_sums)
{
    __shared__ uint32_t s_w[WS_PRE_THREADS / 32];
    const int i = blockIdx.x * WS_PRE_THREADS + threadIdx.x;
    uint32_t s = 0;
    if (i < P) {
        const uint32_t g = order ? order[i] : (uint32_t)i;
        s = tiles_touched[g] * (uint32_t)(member_start[g + 1] - member_start[g]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int k = 0; k < WS_PRE_THREADS / 32; k++) t += s_w[k];
        block_sums[blockIdx.x] = t;
    }
}

// emit_tiles_ordered with (object, tile) keys: Gaussians in depth order, for each its memberships in ascending
// object order, for each the tiles of its rectangle row-major.  Warp-flattened as emit_tiles_ordered is.
__global__ void __launch_bounds__(WS_PRE_THREADS) objeval_emit(int P, int gx, int gy, uint32_t T,
                                                               const uint32_t *__restrict__ order,
                                                               const int32_t *__restrict__ radii,
                                                               const GRec *__restrict__ grec,
                                                               const uint32_t *__restrict__ tiles_touched,
                                                               const int32_t *__restrict__ member_start,
                                                               const int32_t *__restrict__ member_obj,
                                                               const uint32_t *__restrict__ block_offsets,
                                                               uint32_t *__restrict__ tkeys, uint32_t *__restrict__ vals)
{
    __shared__ uint32_t s_w[WS_PRE_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int i = blockIdx.x * WS_PRE_THREADS + tid;
    uint32_t tt = 0, per = 1, rminx = 0, rminy = 0, width = 1, gid = 0, ms = 0;
    float4 ga = make_float4(0.f, 0.f, 0.f, 0.f), gb = ga;
    if (i < P) {
        gid = order[i];
        const int rad = radii[gid];
        ms = (uint32_t)member_start[gid];
        const uint32_t mc = (uint32_t)member_start[gid + 1] - ms;
        if (rad > 0 && mc > 0) {
            per = tiles_touched[gid];
            tt = per * mc;
            ga = grec[gid].a; gb = grec[gid].b;
            uint2 rmin, rmax;
            get_rect(ga.x, ga.y, rad, gx, gy, rmin, rmax);
            rminx = rmin.x; rminy = rmin.y; width = rmax.x - rmin.x;
            if (per == 0) { per = 1; tt = 0; }
        }
    }
    uint32_t inc = tt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    const uint32_t warp_total = __shfl_sync(0xffffffffu, inc, 31);
    if (lane == 31) s_w[w] = inc;
    __syncthreads();
    uint32_t warp_base = block_offsets[blockIdx.x];
    for (int k = 0; k < w; k++) warp_base += s_w[k];
    const uint32_t excl = inc - tt;
    const QuadTest qt = quad_test_of(ga, gb);
    for (uint32_t k0 = 0; k0 < warp_total; k0 += 32) {
        const uint32_t k = k0 + lane;
        int j = 0; // owner lane = number of lanes whose inclusive sum <= k
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) {
            uint32_t vv = __shfl_sync(0xffffffffu, inc, min(j + s - 1, 31));
            if (vv <= k) j += s;
        }
        j = min(j, 31);
        const uint32_t o_excl = __shfl_sync(0xffffffffu, excl, j);
        const uint32_t o_minx = __shfl_sync(0xffffffffu, rminx, j);
        const uint32_t o_miny = __shfl_sync(0xffffffffu, rminy, j);
        const uint32_t o_w = __shfl_sync(0xffffffffu, width, j);
        const uint32_t o_id = __shfl_sync(0xffffffffu, gid, j);
        const uint32_t o_per = __shfl_sync(0xffffffffu, per, j);
        const uint32_t o_ms = __shfl_sync(0xffffffffu, ms, j);
        const QuadTest oq = shfl_quad_test(qt, j);
        if (k < warp_total) {
            const uint32_t local = k - o_excl;
            const uint32_t memb = local / o_per, tl = local - memb * o_per;
            const uint32_t ty = tl / o_w, tx = tl - ty * o_w;
            const uint32_t obj = (uint32_t)__ldg(member_obj + o_ms + memb);
            tkeys[(size_t)warp_base + k] = obj * T + (o_miny + ty) * (uint32_t)gx + (o_minx + tx);
            const uint32_t qm = quad_mask(oq, (float)((o_minx + tx) * WS_BLOCK_X), (float)((o_miny + ty) * WS_BLOCK_Y));
            vals[(size_t)warp_base + k] = o_id | (qm << WS_ID_BITS);
        }
    }
}

// non-empty (object, tile) slots per block of 256 slots
__global__ void __launch_bounds__(256) job_block_counts(uint32_t VT, const uint2 *__restrict__ ranges,
                                                        uint32_t *__restrict__ block_sums)
{
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    bool on = false;
    if (i < VT) { const uint2 r = ranges[i]; on = r.y > r.x; }
    const int c = __syncthreads_count(on);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = (uint32_t)c;
}

// jobs[] = the non-empty slots in ascending (object, tile) order
__global__ void __launch_bounds__(256) job_emit(uint32_t VT, const uint2 *__restrict__ ranges,
                                                const uint32_t *__restrict__ block_offsets, uint32_t *__restrict__ jobs)
{
    __shared__ uint32_t s_w[8];
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    bool on = false;
    if (i < VT) { const uint2 r = ranges[i]; on = r.y > r.x; }
    const uint32_t m = __ballot_sync(0xffffffffu, on);
    if (lane == 0) s_w[w] = __popc(m);
    __syncthreads();
    uint32_t base = block_offsets[blockIdx.x];
    for (int k = 0; k < w; k++) base += s_w[k];
    if (on) jobs[base + __popc(m & ((1u << lane) - 1u))] = i;
}

__global__ void __launch_bounds__(256) objeval_fill(size_t n, float *__restrict__ pred, float *__restrict__ min_depth)
{
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) { pred[i] = 0.0f; min_depth[i] = 1.0f; }   // wheatgs_helper.py:259-261 (zeros, zeros, ones)
}

// first job of every object: obj_start[k] = lower_bound(jobs, k*T), k = 0..num_rows (jobs are ascending)
__global__ void __launch_bounds__(256) job_obj_starts(int num_rows, int n_jobs, uint32_t T,
                                                      const uint32_t *__restrict__ jobs, uint32_t *__restrict__ obj_start)
{
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k > num_rows) return;
    const uint32_t key = (uint32_t)k * T;
    int lo = 0, hi = n_jobs;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (jobs[mid] < key) lo = mid + 1; else hi = mid;
    }
    obj_start[k] = (uint32_t)lo;
}

// utils/wheatgs_helper.py:251-270 over the jobs of each object, objects in ascending order.  The objects are a
// sequential chain (each sees what the earlier ones painted), so the whole grid (one CTA per SM, cooperative
// launch: all CTAs resident) works on ONE object at a time: pass 1 finds the overlap with what is painted and the
// two depth sums, a grid barrier, every CTA adds the per-CTA partials in the same fixed order (deterministic)
// and takes the front/behind decision, pass 2 paints, a second grid barrier.
constexpr int MG_THREADS = 256;
struct MergeSync {
    unsigned int arrive;         // grid barrier counter (monotonic)
    unsigned int pad[31];
};
struct MergePartial { double smin, sren; int cnt; int pad; };

__device__ __forceinline__ void grid_barrier(unsigned int *counter, unsigned int &target)
{
    __syncthreads();
    target += gridDim.x;
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
        while (*((volatile unsigned int *)counter) < target) { }
        __threadfence();
    }
    __syncthreads();
}

__global__ void __launch_bounds__(MG_THREADS) objeval_merge(int num_rows, const uint32_t *__restrict__ obj_start,
                                                            const uint32_t *__restrict__ jobs, uint32_t T, int tiles_x,
                                                            int W, int H, const float *__restrict__ job_alpha,
                                                            const float *__restrict__ job_depth, float *pred,
                                                            float *min_depth, MergeSync *sync, MergePartial *partials)
{
    __shared__ int s_cnt[MG_THREADS / 32];
    __shared__ double s_a[MG_THREADS / 32], s_b[MG_THREADS / 32];
    __shared__ int s_behind;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int gtid = blockIdx.x * MG_THREADS + tid, gsize = gridDim.x * MG_THREADS;
    unsigned int target = 0;
    int parity = 0;
    for (int obj = 1; obj < num_rows; obj++) {
        const int j0 = (int)obj_start[obj], j1 = (int)obj_start[obj + 1];
        if (j0 == j1) continue;                                   // nothing of this object on screen
        const int npix = (j1 - j0) * 256;
        const uint32_t tbase = (uint32_t)obj * T;
        // pass 1: the overlap with what is painted (wheatgs_helper.py:263-265)
        int cnt = 0;
        double smin = 0.0, sren = 0.0;
        for (int q = gtid; q < npix; q += gsize) {
            const int j = j0 + (q >> 8), li = q & 255;
            const float al = job_alpha[(size_t)j * 256 + li];
            if (al > 0.5f) {                                      // slots outside the image hold alpha 0
                const uint32_t tile = jobs[j] - tbase;
                const int px = (int)(tile % (uint32_t)tiles_x) * WS_BLOCK_X + (li & 15);
                const int py = (int)(tile / (uint32_t)tiles_x) * WS_BLOCK_Y + (li >> 4);
                const size_t pix = (size_t)py * W + px;
                if (__ldcg(pred + pix) > 0.0f) {
                    cnt++;
                    smin += (double)__ldcg(min_depth + pix);
                    sren += (double)job_depth[(size_t)j * 256 + li];
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cnt += __shfl_down_sync(0xffffffffu, cnt, o);
            smin += __shfl_down_sync(0xffffffffu, smin, o);
            sren += __shfl_down_sync(0xffffffffu, sren, o);
        }
        if (lane == 0) { s_cnt[wid] = cnt; s_a[wid] = smin; s_b[wid] = sren; }
        __syncthreads();
        MergePartial *mine = partials + (size_t)parity * gridDim.x;
        if (tid == 0) {
            int c = 0; double a = 0.0, b = 0.0;
            for (int k = 0; k < MG_THREADS / 32; k++) { c += s_cnt[k]; a += s_a[k]; b += s_b[k]; }
            __stcg(&mine[blockIdx.x].smin, a); __stcg(&mine[blockIdx.x].sren, b); __stcg(&mine[blockIdx.x].cnt, c);
        }
        grid_barrier(&sync->arrive, target);
        if (wid == 0) {
            // warp 0 adds the per-CTA partials: lane l takes CTAs l, l+32, ... then a fixed shuffle tree -- the same
            // order in every CTA and every run (a serial loop of L2 round trips cost ~30 us per object)
            int c = 0; double a = 0.0, b = 0.0;
            for (unsigned k = lane; k < gridDim.x; k += 32) {
                c += __ldcg(&mine[k].cnt); a += __ldcg(&mine[k].smin); b += __ldcg(&mine[k].sren);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                c += __shfl_xor_sync(0xffffffffu, c, o);
                a += __shfl_xor_sync(0xffffffffu, a, o);
                b += __shfl_xor_sync(0xffffffffu, b, o);
            }
            // min_depth[overlap].mean() < render_depth[overlap].mean(), float32 means
            if (lane == 0) s_behind = (c > 0 && (float)(a / (double)c) < (float)(b / (double)c)) ? 1 : 0;
        }
        __syncthreads();
        const bool behind = s_behind != 0;
        // pass 2: paint (wheatgs_helper.py:266-269)
        for (int q = gtid; q < npix; q += gsize) {
            const int j = j0 + (q >> 8), li = q & 255;
            const float al = job_alpha[(size_t)j * 256 + li];
            if (al > 0.5f) {
                const uint32_t tile = jobs[j] - tbase;
                const int px = (int)(tile % (uint32_t)tiles_x) * WS_BLOCK_X + (li & 15);
                const int py = (int)(tile / (uint32_t)tiles_x) * WS_BLOCK_Y + (li >> 4);
                const size_t pix = (size_t)py * W + px;
                if (!behind || !(__ldcg(pred + pix) > 0.0f)) {
                    __stcg(pred + pix, (float)obj);
                    __stcg(min_depth + pix, job_depth[(size_t)j * 256 + li]);
                }
            }
        }
        grid_barrier(&sync->arrive, target);
        parity ^= 1;
    }
}

constexpr int MG_MAX_CTAS = 256;

struct ObjScratch {
    uint2 *ranges;       // [VT]
    uint32_t *jsums, *joffs; // [ceil(VT/256)]
    uint32_t *jobs;      // [min(R, VT)]
    uint32_t *total;     // [4]
    float *min_depth;    // [H*W]
    uint32_t *obj_start; // [num_rows + 1]
    MergeSync *sync;
    MergePartial *partials; // [2][MG_MAX_CTAS]
    static ObjScratch from(char *p, size_t VT, size_t R, size_t HW, size_t num_rows, size_t *bytes = nullptr)
    {
        ObjScratch s;
        char *p0 = p;
        const size_t nb = (VT + 255) / 256;
        s.ranges = carve<uint2>(p, VT);
        s.jsums = carve<uint32_t>(p, nb);
        s.joffs = carve<uint32_t>(p, nb);
        s.jobs = carve<uint32_t>(p, (R < VT ? R : VT) + 64);
        s.total = carve<uint32_t>(p, 4);
        s.min_depth = carve<float>(p, HW);
        s.obj_start = carve<uint32_t>(p, num_rows + 1);
        s.sync = carve<MergeSync>(p, 1);
        s.partials = carve<MergePartial>(p, 2 * MG_MAX_CTAS);
        if (bytes) *bytes = (size_t)(p - p0) + 256;
        return s;
    }
};

static int check_sets(const char *who, const ws_view *v, const ws_objsets *s, uint32_t *VT)
{
    if (!s || s->num_rows < 1 || !s->member_start || !s->member_obj) {
        set_error("%s: object sets need num_rows >= 1, member_start and member_obj", who);
        return WS_E_INVALID;
    }
    const uint64_t T = (uint64_t)((v->image_width + WS_BLOCK_X - 1) / WS_BLOCK_X) *
                       ((v->image_height + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    if (T * (uint64_t)s->num_rows >= (1ull << 31)) {
        set_error("%s: num_rows x tiles = %llu does not fit the 31-bit (object, tile) key", who,
                  (unsigned long long)(T * (uint64_t)s->num_rows));
        return WS_E_UNSUPPORTED;
    }
    *VT = (uint32_t)(T * (uint64_t)s->num_rows);
    return WS_OK;
}

} // namespace ws

using namespace ws;

extern "C" size_t ws_objeval_bytes(int32_t num_rows, int32_t width, int32_t height, int64_t num_instances)
{
    if (num_rows < 1 || width <= 0 || height <= 0 || num_instances < 0) return 0;
    const size_t T = (size_t)((width + WS_BLOCK_X - 1) / WS_BLOCK_X) * ((height + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    size_t b = 0;
    ObjScratch::from(nullptr, T * (size_t)num_rows, (size_t)num_instances, (size_t)width * height, (size_t)num_rows, &b);
    return b + 256;
}

extern "C" int64_t ws_objeval_prepare(const ws_view *view, const ws_gaussians *g, const ws_objsets *sets,
                                      int32_t *radii, void *geom, size_t geom_bytes, ws_stream_t stream)
{
    int rc = check_view_gauss("ws_objeval_prepare", view, g);
    if (rc != WS_OK) return rc;
    uint32_t VT = 0;
    rc = check_sets("ws_objeval_prepare", view, sets, &VT);
    if (rc != WS_OK) return rc;
    if (!radii || !geom) { set_error("ws_objeval_prepare: NULL radii/geom"); return WS_E_INVALID; }
    if (geom_bytes < ws_geom_bytes(g->P)) { set_error("ws_objeval_prepare: geom scratch too small"); return WS_E_SCRATCH; }
    cudaStream_t st = (cudaStream_t)stream;
    const bool debug = view->debug != 0;
    GeomState gs = GeomState::from((char *)geom, (size_t)g->P);
    rc = launch_preprocess(view, g, radii, gs, nullptr, st);
    if (rc != WS_OK) return rc;
    const int nb = (g->P + WS_PRE_THREADS - 1) / WS_PRE_THREADS;
    objeval_block_sums<<<nb, WS_PRE_THREADS, 0, st>>>(g->P, nullptr, gs.tiles_touched, sets->member_start, gs.block_sums);
    WS_LAUNCH_CHECK(st, debug);
    rc = launch_scan_u32(nb, gs.block_sums, gs.block_offsets, gs.total, st, debug);
    if (rc != WS_OK) return rc;
    uint32_t *th = total_host();
    if (!th) { set_error("ws_objeval_prepare: cudaHostAlloc failed"); return WS_E_CUDA; }
    WS_CUDA_CHECK(cudaMemcpyAsync(th, gs.total, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    cudaEvent_t ev = total_event();
    WS_CUDA_CHECK(cudaEventRecord(ev, st));
    rc = launch_depth_sort(g->P, gs, st, debug);   // overlaps the host's wait, as in ws_forward_prepare
    if (rc != WS_OK) return rc;
    WS_CUDA_CHECK(cudaEventSynchronize(ev));
    return (int64_t)*th;
}

extern "C" int64_t ws_objeval_bin(const ws_view *view, const ws_gaussians *g, const ws_objsets *sets,
                                  const int32_t *radii, void *geom, void *binning, size_t binning_bytes,
                                  void *scratch, size_t scratch_bytes, int64_t num_instances, ws_stream_t stream)
{
    int rc = check_view_gauss("ws_objeval_bin", view, g);
    if (rc != WS_OK) return rc;
    uint32_t VT = 0;
    rc = check_sets("ws_objeval_bin", view, sets, &VT);
    if (rc != WS_OK) return rc;
    if (!radii || !geom || !scratch || num_instances < 0 || (num_instances > 0 && !binning)) {
        set_error("ws_objeval_bin: NULL buffer");
        return WS_E_INVALID;
    }
    const int W = view->image_width, H = view->image_height;
    if (scratch_bytes < ws_objeval_bytes(sets->num_rows, W, H, num_instances)) {
        set_error("ws_objeval_bin: scratch too small");
        return WS_E_SCRATCH;
    }
    if (num_instances > 0 && binning_bytes < ws_binning_bytes(num_instances, W, H)) {
        set_error("ws_objeval_bin: binning scratch too small");
        return WS_E_SCRATCH;
    }
    if (num_instances == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const bool debug = view->debug != 0;
    const int gx = (W + WS_BLOCK_X - 1) / WS_BLOCK_X, gy = (H + WS_BLOCK_Y - 1) / WS_BLOCK_Y;
    const uint32_t T = (uint32_t)gx * (uint32_t)gy;
    const size_t R = (size_t)num_instances;
    GeomState gs = GeomState::from((char *)geom, (size_t)g->P);
    BinningState bs = BinningState::from((char *)binning, R);
    ObjScratch os = ObjScratch::from((char *)scratch, VT, R, (size_t)W * H, (size_t)sets->num_rows);
    const int nb = (g->P + WS_PRE_THREADS - 1) / WS_PRE_THREADS;
    const uint32_t *order = depth_order(gs);
    objeval_block_sums<<<nb, WS_PRE_THREADS, 0, st>>>(g->P, order, gs.tiles_touched, sets->member_start, gs.block_sums);
    WS_LAUNCH_CHECK(st, debug);
    rc = launch_scan_u32(nb, gs.block_sums, gs.block_offsets, gs.total + 1, st, debug);
    if (rc != WS_OK) return rc;
    objeval_emit<<<nb, WS_PRE_THREADS, 0, st>>>(g->P, gx, gy, T, order, radii, gs.grec, gs.tiles_touched,
                                                sets->member_start, sets->member_obj, gs.block_offsets, bs.tkeys_a,
                                                bs.vals_a);
    WS_LAUNCH_CHECK(st, debug);
    uint32_t *tkeys = nullptr, *vals = nullptr;
    rc = launch_sort_pairs32(R, bs.tkeys_a, bs.tkeys_b, bs.vals_a, bs.vals_b, (int)get_higher_msb(VT), bs.sort, st, debug,
                             &tkeys, &vals);
    if (rc != WS_OK) return rc;
    ImageState is;
    is.ranges = os.ranges; is.n_contrib = nullptr;
    (void)vals;
    rc = launch_identify_ranges(R, (int)VT, nullptr, tkeys, is.ranges, st, debug);
    if (rc != WS_OK) return rc;
    const uint32_t jb = (VT + 255u) / 256u;
    job_block_counts<<<jb, 256, 0, st>>>(VT, os.ranges, os.jsums);
    WS_LAUNCH_CHECK(st, debug);
    rc = launch_scan_u32((int)jb, os.jsums, os.joffs, os.total, st, debug);
    if (rc != WS_OK) return rc;
    job_emit<<<jb, 256, 0, st>>>(VT, os.ranges, os.joffs, os.jobs);
    WS_LAUNCH_CHECK(st, debug);
    uint32_t *th = total_host();
    if (!th) { set_error("ws_objeval_bin: cudaHostAlloc failed"); return WS_E_CUDA; }
    WS_CUDA_CHECK(cudaMemcpyAsync(th + 1, os.total, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    WS_CUDA_CHECK(cudaStreamSynchronize(st));
    return (int64_t)th[1];
}

extern "C" int ws_objeval_render(const ws_view *view, const ws_objsets *sets, int32_t P, void *geom, void *binning, void *scratch,
                                 size_t scratch_bytes, int64_t num_instances, int64_t num_jobs, float *job_out,
                                 size_t job_out_bytes, float *pred_mask, ws_stream_t stream)
{
    if (!view || view->image_width <= 0 || view->image_height <= 0 || !view->bg) {
        set_error("ws_objeval_render: bad view");
        return WS_E_INVALID;
    }
    uint32_t VT = 0;
    int rc = check_sets("ws_objeval_render", view, sets, &VT);
    if (rc != WS_OK) return rc;
    if (!scratch || !pred_mask || !geom || P <= 0 || num_instances < 0 || num_jobs < 0 || num_jobs > num_instances ||
        (num_jobs > 0 && (!binning || !job_out))) {
        set_error("ws_objeval_render: NULL buffer or bad counts");
        return WS_E_INVALID;
    }
    const int W = view->image_width, H = view->image_height;
    if (scratch_bytes < ws_objeval_bytes(sets->num_rows, W, H, num_instances)) {
        set_error("ws_objeval_render: scratch too small");
        return WS_E_SCRATCH;
    }
    if (job_out_bytes < (size_t)num_jobs * 256 * 2 * sizeof(float)) {
        set_error("ws_objeval_render: job_out too small (%zu < %zu)", job_out_bytes, (size_t)num_jobs * 2048);
        return WS_E_SCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const bool debug = view->debug != 0;
    const int gx = (W + WS_BLOCK_X - 1) / WS_BLOCK_X, gy = (H + WS_BLOCK_Y - 1) / WS_BLOCK_Y;
    const uint32_t T = (uint32_t)gx * (uint32_t)gy;
    ObjScratch os = ObjScratch::from((char *)scratch, VT, (size_t)num_instances, (size_t)W * H, (size_t)sets->num_rows);
    const size_t HW = (size_t)W * H;
    objeval_fill<<<(unsigned)((HW + 255) / 256), 256, 0, st>>>(HW, pred_mask, os.min_depth);
    WS_LAUNCH_CHECK(st, debug);
    if (num_jobs == 0) return WS_OK;
    BinningState bs = BinningState::from((char *)binning, (size_t)num_instances);
    GeomState gs = GeomState::from((char *)geom, (size_t)P);
    // the (object, tile) sort ran over get_higher_msb(VT) bits: the sorted values sit in "a" after an even pass count
    const uint32_t *vals = ((((int)get_higher_msb(VT) + 7) / 8) & 1) ? bs.vals_b : bs.vals_a;
    float *job_alpha = job_out, *job_depth = job_out + (size_t)num_jobs * 256;
    rc = launch_render_jobs(view, P, gs, vals, os.ranges, os.jobs, (int)num_jobs, (int)T, job_alpha, job_depth, st);
    if (rc != WS_OK) return rc;
    job_obj_starts<<<(sets->num_rows + 256) / 256, 256, 0, st>>>(sets->num_rows, (int)num_jobs, T, os.jobs, os.obj_start);
    WS_LAUNCH_CHECK(st, debug);
    WS_CUDA_CHECK(cudaMemsetAsync(os.sync, 0, sizeof(MergeSync), st));
    int dev = 0, sms = 0, per_sm = 0;
    WS_CUDA_CHECK(cudaGetDevice(&dev));
    WS_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    WS_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, objeval_merge, MG_THREADS, 0));
    int ctas = sms < MG_MAX_CTAS ? sms : MG_MAX_CTAS;          // one CTA per SM: 148 on B200
    if (per_sm < 1) { set_error("ws_objeval_render: merge kernel does not fit an SM"); return WS_E_CUDA; }
    int num_rows = sets->num_rows, n_tiles_x = gx, w_ = W, h_ = H;
    const uint32_t *obj_start = os.obj_start, *jobs = os.jobs;
    const float *ja = job_alpha, *jd = job_depth;
    float *pm = pred_mask, *md = os.min_depth;
    MergeSync *sy = os.sync;
    MergePartial *pa = os.partials;
    uint32_t t_ = T;
    void *args[] = {&num_rows, &obj_start, &jobs, &t_, &n_tiles_x, &w_, &h_, &ja, &jd, &pm, &md, &sy, &pa};
    WS_CUDA_CHECK(cudaLaunchCooperativeKernel((const void *)objeval_merge, dim3(ctas), dim3(MG_THREADS), args, 0, st));
    WS_LAUNCH_CHECK(st, debug);
    return WS_OK;
}
