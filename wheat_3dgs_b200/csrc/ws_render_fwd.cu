// ws_render_fwd.cu -- forward alpha blending (replaces renderCUDA, DGR forward.cu:261-381, and
// the flashsplat variant FSR forward.cu:269-418).
//
// One CTA per 16x16 tile, 128 threads (4 warps); warp w owns the 8x8 pixel block
// (x0 + (w&1)*8 .. +7, y0 + (w>>1)*8 .. +7) and every lane TWO pixels of it, (l&7, l>>3) and
// (l&7, (l>>3)+4).  The two pixels share dx, so the per-pair arithmetic runs as packed f32x2
// instructions (FADD2/FMUL2/FFMA2, sm_100: one issue slot does both pixels, scalar operands
// broadcast for free) and the per-record overhead (3 LDS.128, loop control) is paid once
// per 64 pixels.  Every packed op is the same IEEE operation, in the same order, the reference
// build performs per pixel (its FMA contraction read from its SASS, expf2 = its expf sequence), so
// colour, depth, alpha and n_contrib stay bit-identical to the reference.
//
// Staging: ws_stage.cuh.  Every warp compacts the tile's value list by its own quadrant's mask bit and brings
// the surviving Gaussians' rows in with TMA gather4 copies through a private ring of mbarrier-guarded stages:
// the four warps of a CTA never wait for each other.  A dropped record would have been skipped by every pixel's
// own test (forward.cu:337-347), so results are unchanged, including `contributor` positions: every record
// carries its position in the tile's full list.
//
// A pixel that skips a record (power > 0, alpha < 1/255, finished, wrong label) takes part in the
// packed blend with alpha = 0, which leaves its accumulators bit-for-bit unchanged.
//
// Early termination: a pixel stops when T would fall below 1e-4 (forward.cu:348-353); it is marked by
// flipping the sign of its T (no separate flag to test per record: the flipped T fails `test_T < 1e-4` again
// for whatever comes next).  A warp stops fetching when all its 64 pixels are done (forward.cu:305-308 for the
// warp's block).
#include "ws_stage.cuh"

namespace ws {

#ifndef WS_RF_NSTAGE
#define WS_RF_NSTAGE 2
#endif
#ifndef WS_RF_MINB
#define WS_RF_MINB 6
#endif
constexpr int RF_WARPS = 4;             // 8x8 pixel blocks (warps) per CTA: the tile
constexpr int RF_THREADS = 32 * RF_WARPS;
constexpr int RF_NSTAGE = WS_RF_NSTAGE;
static_assert(RF_NSTAGE * ST_CNT_BITS <= 32 && (RF_NSTAGE & (RF_NSTAGE - 1)) == 0, "ring depth: a power of two, counts in one register");
#ifndef WS_RF_UNROLL
#define WS_RF_UNROLL 4
#endif
constexpr int RF_UNROLL = WS_RF_UNROLL; // records per trip of the blend loop
using FwdRing = WarpRing<RF_NSTAGE>;

struct RenderFwdParams {
    const uint2 *ranges;
    const uint32_t *vals; // sorted values: Gaussian id | quadrant mask << 28
    int W, H, P, tiles_x;
    const float *bg;
    float *out_color, *out_depth, *out_alpha;
    uint32_t *n_contrib;
    // flashsplat
    const float *gt_mask;
    const int32_t *unique_label;
    float *used_count;
    int32_t *out_contrib_num;
    int num_obj;
    // (object, tile) jobs of the batched eval_obj_labels (ws_objeval.cu): CTA j blends the records of
    // ranges[jobs[j]] over tile jobs[j] % tiles_total and writes alpha/depth of its 256 pixels to job_*[j][16][16]
    const uint32_t *jobs;
    int tiles_total;
    float *job_alpha, *job_depth;
    ExpConsts ec;
};

// MODE 0: the DGR rasterizer; 1: flashsplat (label votes); 2: (object, tile) jobs, alpha and depth only
template <int MODE>
__global__ void __launch_bounds__(RF_THREADS, WS_RF_MINB) render_fwd(const __grid_constant__ TensorMap tmap, const RenderFwdParams p)
{
    constexpr bool FLASH = MODE == 1;
    constexpr bool JOBS = MODE == 2;
    __shared__ FwdRing rings[RF_WARPS];

    const int lane = threadIdx.x & 31;
    const uint32_t w = threadIdx.x >> 5;
    FwdRing &sm = rings[w];
    const uint32_t slot0 = blockIdx.x;
    const uint32_t slot = JOBS ? __ldg(p.jobs + slot0) : slot0;           // index into ranges
    const uint32_t tile = JOBS ? slot % (uint32_t)p.tiles_total : slot;
    const int tile_y = (int)(tile / (uint32_t)p.tiles_x), tile_x = (int)(tile - (uint32_t)tile_y * p.tiles_x);
    const int wx0 = tile_x * WS_BLOCK_X + (int)(w & 1) * 8;
    const int wy0 = tile_y * WS_BLOCK_Y + (int)(w >> 1) * 8;
    const int px = wx0 + (lane & 7), py0 = wy0 + (lane >> 3), py1 = py0 + 4;
    const bool in0 = px < p.W && py0 < p.H, in1 = px < p.W && py1 < p.H;
    const uint32_t pix0 = (uint32_t)p.W * py0 + px, pix1 = pix0 + 4u * (uint32_t)p.W;
    const float pixfx = (float)px;
    const f2 npy = pk2(-(float)py0, -(float)py1);

    sm.init(lane);
    pdl_wait();   // everything above is this CTA's own business; the tile ranges and the sorted list come from upstream
    const uint2 range = p.ranges[slot];
    const int n = (int)(range.y - range.x);
    const int nb = (n + ST_SEG - 1) / ST_SEG;
    const uint32_t *__restrict__ vals = p.vals + range.x;

    int obj0 = 0, obj1 = 0;
    if (FLASH) {
        if (in0) obj0 = __float2int_rz(p.gt_mask[pix0]);
        if (in1) obj1 = __float2int_rz(p.gt_mask[pix1]);
    }
    const bool ok0 = FLASH ? (obj0 >= 0 && obj0 <= p.num_obj) : true;
    const bool ok1 = FLASH ? (obj1 >= 0 && obj1 <= p.num_obj) : true;

    // loop-carried state as scalar halves (packed only at the point of use: ptxas copies 64-bit
    // loop-carried values through temporaries otherwise).  T < 0: the pixel is finished (or outside the image),
    // |T| is its final transmittance.
    float Ta = in0 ? 1.0f : -1.0f, Tb = in1 ? 1.0f : -1.0f, C0a = 0.f, C0b = 0.f, C1a = 0.f, C1b = 0.f, C2a = 0.f,
          C2b = 0.f, Wa = 0.f, Wb = 0.f, Da = 0.f, Db = 0.f;
    uint32_t last0 = 0, last1 = 0;
    const ExpConsts ec = p.ec;

    // ring prologue: batches 0 .. NSTAGE-2 requested, the value of batch NSTAGE-1 in flight
    uint2 nv;
    uint32_t cq = 0; // survivors per requested batch, ST_CNT_BITS bits each, oldest lowest
    {
        uint2 pv[RF_NSTAGE - 1];
#pragma unroll
        for (int k = 0; k < RF_NSTAGE - 1; k++) pv[k] = FwdRing::load(vals, n, nb, k, lane);
        nv = FwdRing::load(vals, n, nb, RF_NSTAGE - 1, lane);
#pragma unroll
        for (int k = 0; k < RF_NSTAGE - 1; k++)
            if (k < nb) cq |= sm.issue(&tmap, k, k, pv[k], w, lane) << (ST_CNT_BITS * k);
    }
    for (int b = 0; b < nb; b++) {
        const int s = b % RF_NSTAGE;
        // slot (b-1) % NSTAGE: the warp is through with it (its reads fed the arithmetic of batch b-1)
        if (b + RF_NSTAGE - 1 < nb)
            cq |= sm.issue(&tmap, (b + RF_NSTAGE - 1) % RF_NSTAGE, b + RF_NSTAGE - 1, nv, w, lane) << (ST_CNT_BITS * (RF_NSTAGE - 1));
        nv = FwdRing::load(vals, n, nb, b + RF_NSTAGE, lane);
        const int cnt = (int)(cq & ((1u << ST_CNT_BITS) - 1u));
        cq >>= ST_CNT_BITS;
        sm.wait(b);

#pragma unroll RF_UNROLL
        for (int rr = 0; rr < cnt; rr++) {
            const float4 a = sm.rec[s][rr].a;
            const float4 q = sm.rec[s][rr].b;
            const uint2 meta = sm.meta[s][rr];
            const uint32_t gid = meta.x;
            bool a0 = true, a1 = true;
            if (FLASH) {
                if (p.unique_label != nullptr) { // FSR forward.cu:357-362
                    const int uid = __ldg(p.unique_label + gid);
                    if (uid >= 0 && obj0 != uid) a0 = false;
                    if (uid >= 0 && obj1 != uid) a1 = false;
                }
            }
            // DGR forward.cu:333-353, op order from the reference SASS, two pixels at once.
            // Straight-line: a pixel that fails a test carries on with alpha = 0 (branches
            // here cost more than they skip, and their merges made ptxas copy every packed
            // accumulator).
            const float4 c = sm.rec[s][rr].c;
            const float dx = __fadd_rn(a.x, -pixfx);
            const f2 dy = add2(bc2(a.y), npy);
            const f2 t_zz = mul2(dy, mul2(dy, bc2(q.x)));
            const float t_xx = __fmul_rn(dx, a.z);
            const f2 nt_xy = mul2(dy, bc2(__fmul_rn(dx, -a.w))); // -(dy * (dx * B))
            const f2 power = fma2(fma2(bc2(dx), bc2(t_xx), t_zz), bc2(-0.5f), nt_xy);
            const f2 og = mul2(bc2(q.y), expf2(power, ec));
            const float al0 = fminf(0.99f, lo2(og)), al1 = fminf(0.99f, hi2(og));
            a0 = a0 && !(lo2(power) > 0.0f) && !(al0 < 1.0f / 255.0f);
            a1 = a1 && !(hi2(power) > 0.0f) && !(al1 < 1.0f / 255.0f);
            const f2 T = pk2(Ta, Tb);
            const f2 test_T = mul2(T, fma2(pk2(al0, al1), bc2(-ec.one), bc2(1.0f)));
            const float tt0 = lo2(test_T), tt1 = hi2(test_T);
            // a finished pixel has T < 0: test_T <= 0 ends it again, nothing is blended
            const bool e0 = a0 && tt0 < 0.0001f, e1 = a1 && tt1 < 0.0001f;
            a0 = a0 && !e0;
            a1 = a1 && !e1;
            const f2 al = pk2(a0 ? al0 : 0.0f, a1 ? al1 : 0.0f);
            const f2 C0 = fma2(T, mul2(al, bc2(c.x)), pk2(C0a, C0b));
            const f2 C1 = fma2(T, mul2(al, bc2(c.y)), pk2(C1a, C1b));
            const f2 C2 = fma2(T, mul2(al, bc2(c.z)), pk2(C2a, C2b));
            const f2 Wt = fma2(T, al, pk2(Wa, Wb));
            const f2 D = fma2(T, mul2(al, bc2(q.z)), pk2(Da, Db));
            C0a = lo2(C0); C0b = hi2(C0); C1a = lo2(C1); C1b = hi2(C1); C2a = lo2(C2); C2b = hi2(C2);
            Wa = lo2(Wt); Wb = hi2(Wt); Da = lo2(D); Db = hi2(D);
            if (a0) last0 = meta.y;
            if (a1) last1 = meta.y;
            if (FLASH) {
                // FSR forward.cu:398.  Per-(label, Gaussian) vote: one atomic per warp when the label is uniform
                const f2 uc = mul2(al, T);
                const bool cb0 = a0 && ok0, cb1 = a1 && ok1;
                const float uc0 = lo2(uc), uc1 = hi2(uc);
                const uint32_t cm = __ballot_sync(0xffffffffu, cb0 || cb1);
                if (cm) {
                    const int first = __ffs(cm) - 1;
                    const int mine = cb0 ? obj0 : obj1;
                    const int oid0 = __shfl_sync(0xffffffffu, mine, first);
                    const uint32_t diff = __ballot_sync(0xffffffffu, (cb0 && obj0 != oid0) || (cb1 && obj1 != oid0));
                    if (diff == 0) {
                        float v = (cb0 ? uc0 : 0.0f) + (cb1 ? uc1 : 0.0f);
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                        if (lane == 0) atomicAdd(p.used_count + (size_t)oid0 * p.P + gid, v);
                    } else {
                        if (cb0) atomicAdd(p.used_count + (size_t)obj0 * p.P + gid, uc0);
                        if (cb1) atomicAdd(p.used_count + (size_t)obj1 * p.P + gid, uc1);
                    }
                }
            }
            const float Tn0 = a0 ? tt0 : Ta, Tn1 = a1 ? tt1 : Tb;
            Ta = e0 ? -fabsf(Tn0) : Tn0;
            Tb = e1 ? -fabsf(Tn1) : Tn1;
        }
        if (__all_sync(0xffffffffu, Ta < 0.0f && Tb < 0.0f)) {
            // the gathers already requested may still be in flight into this warp's ring:
            // they must land before the CTA can exit and the memory is handed to another CTA.
            for (int k = b + 1; k < nb && k < b + RF_NSTAGE; k++) sm.wait(k);
            break;
        }
    }
    Ta = fabsf(Ta); Tb = fabsf(Tb);

    if (JOBS) {
        // every slot of the job's 16x16 block is written (zeros outside the image)
        const uint32_t li = (uint32_t)(py0 - tile_y * WS_BLOCK_Y) * WS_BLOCK_X + (uint32_t)(px - tile_x * WS_BLOCK_X);
        float *ja = p.job_alpha + (size_t)slot0 * (WS_BLOCK_X * WS_BLOCK_Y);
        float *jd = p.job_depth + (size_t)slot0 * (WS_BLOCK_X * WS_BLOCK_Y);
        ja[li] = in0 ? Wa : 0.0f;
        jd[li] = in0 ? Da : 0.0f;
        ja[li + 4 * WS_BLOCK_X] = in1 ? Wb : 0.0f;
        jd[li + 4 * WS_BLOCK_X] = in1 ? Db : 0.0f;
        return;
    }
    const size_t HW = (size_t)p.H * p.W;
    const float bg0 = __ldg(p.bg + 0), bg1 = __ldg(p.bg + 1), bg2 = __ldg(p.bg + 2);
    // forward.cu:376-377: C + T * bg (contracted to one FMA in the reference build)
    const f2 T = pk2(Ta, Tb), Wt = pk2(Wa, Wb), D = pk2(Da, Db);
    const f2 o0 = fma2(bc2(bg0), T, pk2(C0a, C0b)), o1 = fma2(bc2(bg1), T, pk2(C1a, C1b)),
             o2 = fma2(bc2(bg2), T, pk2(C2a, C2b));
    if (in0) {
        p.n_contrib[pix0] = last0;
        p.out_color[pix0] = lo2(o0);
        p.out_color[HW + pix0] = lo2(o1);
        p.out_color[2 * HW + pix0] = lo2(o2);
        p.out_alpha[pix0] = lo2(Wt);
        p.out_depth[pix0] = lo2(D);
        if (FLASH) p.out_contrib_num[pix0] = (int32_t)last0;
    }
    if (in1) {
        p.n_contrib[pix1] = last1;
        p.out_color[pix1] = hi2(o0);
        p.out_color[HW + pix1] = hi2(o1);
        p.out_color[2 * HW + pix1] = hi2(o2);
        p.out_alpha[pix1] = hi2(Wt);
        p.out_depth[pix1] = hi2(D);
        if (FLASH) p.out_contrib_num[pix1] = (int32_t)last1;
    }
}

static int check_ids(int P)
{
    if ((uint32_t)P > WS_ID_MASK) {
        set_error("render: %d Gaussians exceed the 2^%u - 1 ids a sorted value can hold", P, WS_ID_BITS);
        return WS_E_UNSUPPORTED;
    }
    return WS_OK;
}

int launch_render_fwd(const ws_view *v, int P, GeomState &gs, const uint32_t *vals, ImageState &is,
                      float *out_color, float *out_depth, float *out_alpha, const ws_flash *fl,
                      cudaStream_t st)
{
    int rc = check_ids(P);
    if (rc != WS_OK) return rc;
    TensorMap tm;
    rc = make_grec_tensor_map(&tm, gs.grec, P);
    if (rc != WS_OK) return rc;
    RenderFwdParams p;
    p.ec = exp_consts_host();
    p.ranges = is.ranges; p.vals = vals;
    p.W = v->image_width; p.H = v->image_height; p.P = P;
    p.bg = v->bg; // 3-float device tensor (DGR __init__.py:164), read by the kernel
    p.out_color = out_color; p.out_depth = out_depth; p.out_alpha = out_alpha;
    p.n_contrib = is.n_contrib;
    p.gt_mask = fl ? fl->gt_mask : nullptr;
    p.unique_label = fl ? fl->unique_label : nullptr;
    p.used_count = fl ? fl->used_count : nullptr;
    p.out_contrib_num = fl ? fl->out_contrib_num : nullptr;
    p.num_obj = fl ? fl->num_obj : 0;
    p.tiles_x = (p.W + WS_BLOCK_X - 1) / WS_BLOCK_X;
    const unsigned grid = (unsigned)p.tiles_x * (unsigned)((p.H + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    p.jobs = nullptr; p.tiles_total = 0; p.job_alpha = nullptr; p.job_depth = nullptr;
    if (fl)
        WS_CUDA_CHECK(launch_pdl(render_fwd<1>, dim3(grid), dim3(RF_THREADS), 0, st, tm, p));
    else
        WS_CUDA_CHECK(launch_pdl(render_fwd<0>, dim3(grid), dim3(RF_THREADS), 0, st, tm, p));
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

int launch_render_jobs(const ws_view *v, int P, GeomState &gs, const uint32_t *vals, const uint2 *ranges,
                       const uint32_t *jobs, int n_jobs, int T, float *job_alpha, float *job_depth, cudaStream_t st)
{
    if (n_jobs <= 0) return WS_OK;
    int rc = check_ids(P);
    if (rc != WS_OK) return rc;
    TensorMap tm;
    rc = make_grec_tensor_map(&tm, gs.grec, P);
    if (rc != WS_OK) return rc;
    RenderFwdParams p;
    p.ec = exp_consts_host();
    p.ranges = ranges; p.vals = vals;
    p.W = v->image_width; p.H = v->image_height; p.P = P;
    p.bg = v->bg;
    p.out_color = nullptr; p.out_depth = nullptr; p.out_alpha = nullptr; p.n_contrib = nullptr;
    p.gt_mask = nullptr; p.unique_label = nullptr; p.used_count = nullptr; p.out_contrib_num = nullptr;
    p.num_obj = 0;
    p.tiles_x = (p.W + WS_BLOCK_X - 1) / WS_BLOCK_X;
    p.jobs = jobs; p.tiles_total = T; p.job_alpha = job_alpha; p.job_depth = job_depth;
    render_fwd<2><<<(unsigned)n_jobs, RF_THREADS, 0, st>>>(tm, p);
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

} // namespace ws
