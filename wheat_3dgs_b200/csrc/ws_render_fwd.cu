// ws_render_fwd.cu -- forward alpha blending (replaces renderCUDA, DGR forward.cu:261-381, and
// the flashsplat variant FSR forward.cu:269-418).
//
// One CTA per 16x16 tile, 128 threads (WS_RF_WARPS = 4; 2 or 1 warps per CTA split the tile into
// adjacent CTAs that each stage the tile's records -- all within 3 % of each other on B200, the
// kernel is instruction-issue bound); warp w owns the 8x8 pixel block
// (x0 + (w&1)*8 .. +7, y0 + (w>>1)*8 .. +7) and every lane TWO pixels of it, (l&7, l>>3) and
// (l&7, (l>>3)+4).  The two pixels share dx, so the per-pair arithmetic runs as packed f32x2
// instructions (FADD2/FMUL2/FFMA2, sm_100: one issue slot does both pixels, scalar operands
// broadcast for free) and the per-record overhead (bit scan, 3 LDS.128, loop control) is paid once
// per 64 pixels.  Every packed op is the same IEEE operation, in the same order, the reference
// build performs per pixel (its FMA contraction read from its SASS, expf2 = its expf sequence), so
// colour, depth, alpha and n_contrib stay bit-identical to the reference.
//
// Staging: the tile's instance records were gathered into tile order by identify_ranges_pack,
// so a batch of 256 records is three contiguous 4 KB runs (q0,q1,q2).  Thread 0 brings batch
// b+1 in with three TMA bulk copies (cp.async.bulk ... mbarrier::complete_tx::bytes) into the
// other half of a double buffer while all warps blend batch b; consumers wait on the
// mbarrier's phase parity.  No thread issues a per-record global load.
//
// Sub-tile culling: before the per-pixel loop each lane tests one record of a 32-record chunk
// against the warp's 8x8 pixel rectangle: the exact minimum of the quadratic form
// q = 1/2(A dx^2 + C dy^2) + B dx dy over the rectangle (a convex minimum lies on the two edges
// facing the centre) is compared with tau = ln(255*opacity) plus a rounding margin.  Records
// that cannot reach alpha >= 1/255 anywhere in the rectangle are dropped by a ballot; the warp
// then walks only the set bits.  Dropped records would have been skipped by every pixel's own
// test (forward.cu:337-347), so results are unchanged, including `contributor` positions.
//
// A pixel that skips a record (power > 0, alpha < 1/255, finished, wrong label) takes part in the
// packed blend with alpha = 0, which leaves its accumulators bit-for-bit unchanged.
//
// Early termination: a pixel stops when T would fall below 1e-4 (forward.cu:348-353); a warp
// leaves a batch when all 64 pixels are done (warp vote); the CTA stops fetching when
// all its threads are done (forward.cu:305-308).
#include "ws_common.cuh"

namespace ws {

#ifndef WS_RF_BATCH
#define WS_RF_BATCH 256
#endif
#ifndef WS_RF_WARPS
#define WS_RF_WARPS 4
#endif
constexpr int RF_BATCH = WS_RF_BATCH; // records per stage
constexpr int RF_WARPS = WS_RF_WARPS;    // 8x8 pixel blocks (warps) per CTA: 1, 2 (a 16x8 half tile) or 4 (the tile)
constexpr int RF_THREADS = 32 * RF_WARPS;

struct RenderFwdParams {
    const uint2 *ranges;
    const float4 *q0, *q1, *q2;
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
    // (object, tile) jobs of the batched eval_obj_labels (ws_objeval.cu): CTA group j blends the records of
    // ranges[jobs[j]] over tile jobs[j] % tiles_total and writes alpha/depth of its 256 pixels to job_*[j][16][16]
    const uint32_t *jobs;
    int tiles_total;
    float *job_alpha, *job_depth;
};

// MODE 0: the DGR rasterizer; 1: flashsplat (label votes); 2: (object, tile) jobs, alpha and depth only
template <int MODE>
__global__ void __launch_bounds__(RF_THREADS) render_fwd(const RenderFwdParams p)
{
    constexpr bool FLASH = MODE == 1;
    constexpr bool JOBS = MODE == 2;
    __shared__ __align__(128) float4 s_q0[2][RF_BATCH];
    __shared__ __align__(128) float4 s_q1[2][RF_BATCH];
    __shared__ __align__(128) float4 s_q2[2][RF_BATCH];
    __shared__ __align__(8) uint64_t s_full[2];

    const int lane = threadIdx.x & 31;
    const bool leader = threadIdx.x == 0;
    // the CTAs of a tile are adjacent in launch order
    const uint32_t slot0 = blockIdx.x / (4 / RF_WARPS);
    const uint32_t slot = JOBS ? __ldg(p.jobs + slot0) : slot0;           // index into ranges
    const uint32_t tile = JOBS ? slot % (uint32_t)p.tiles_total : slot;
    const uint32_t w = (blockIdx.x % (4 / RF_WARPS)) * RF_WARPS + (threadIdx.x >> 5);
    const int tile_y = (int)(tile / (uint32_t)p.tiles_x), tile_x = (int)(tile - (uint32_t)tile_y * p.tiles_x);
    const int wx0 = tile_x * WS_BLOCK_X + (int)(w & 1) * 8;
    const int wy0 = tile_y * WS_BLOCK_Y + (int)(w >> 1) * 8;
    const int px = wx0 + (lane & 7), py0 = wy0 + (lane >> 3), py1 = py0 + 4;
    const bool in0 = px < p.W && py0 < p.H, in1 = px < p.W && py1 < p.H;
    const uint32_t pix0 = (uint32_t)p.W * py0 + px, pix1 = pix0 + 4u * (uint32_t)p.W;
    const float pixfx = (float)px;
    const f2 npy = pk2(-(float)py0, -(float)py1);
    const float fx0 = (float)wx0, fy0 = (float)wy0;

    const uint2 range = p.ranges[slot];
    const int n = (int)(range.y - range.x);
    const int nb = (n + RF_BATCH - 1) / RF_BATCH;

    if (leader) {
        mbar_init(&s_full[0], 1);
        mbar_init(&s_full[1], 1);
        mbar_fence_init();
    }
    if (RF_WARPS == 1) __syncwarp(); else __syncthreads();

    auto issue = [&](int b) {
        const int s = b & 1;
        const int cnt = min(RF_BATCH, n - b * RF_BATCH);
        const uint32_t bytes = (uint32_t)cnt * 16u;
        const size_t off = (size_t)range.x + (size_t)b * RF_BATCH;
        mbar_expect_tx(&s_full[s], 3u * bytes);
        tma_load_1d(&s_q0[s][0], p.q0 + off, bytes, &s_full[s]);
        tma_load_1d(&s_q1[s][0], p.q1 + off, bytes, &s_full[s]);
        tma_load_1d(&s_q2[s][0], p.q2 + off, bytes, &s_full[s]);
    };

    int obj0 = 0, obj1 = 0;
    if (FLASH) {
        if (in0) obj0 = __float2int_rz(p.gt_mask[pix0]);
        if (in1) obj1 = __float2int_rz(p.gt_mask[pix1]);
    }
    const bool ok0 = FLASH ? (obj0 >= 0 && obj0 <= p.num_obj) : true;
    const bool ok1 = FLASH ? (obj1 >= 0 && obj1 <= p.num_obj) : true;

    uint32_t live = (in0 ? 1u : 0u) | (in1 ? 2u : 0u); // bit i: pixel i still blending
    // loop-carried state as scalar halves (packed only at the point of use: ptxas copies 64-bit
    // loop-carried values through temporaries otherwise)
    float Ta = 1.0f, Tb = 1.0f, C0a = 0.f, C0b = 0.f, C1a = 0.f, C1b = 0.f, C2a = 0.f, C2b = 0.f, Wa = 0.f, Wb = 0.f,
          Da = 0.f, Db = 0.f;
    uint32_t last0 = 0, last1 = 0;

    if (nb > 0 && leader) issue(0);
    for (int b = 0; b < nb; b++) {
        const int s = b & 1;
        const int cnt = min(RF_BATCH, n - b * RF_BATCH);
        if (b + 1 < nb && leader) issue(b + 1); // stage (b+1)&1 was released by the vote below
        mbar_wait(&s_full[s], (uint32_t)((b >> 1) & 1));

        for (int c0 = 0; c0 < cnt; c0 += 32) {
            if (__all_sync(0xffffffffu, live == 0u)) break;
            const int r = c0 + lane;
            bool hit = false;
            if (r < cnt) hit = rect_hit(s_q0[s][r], s_q1[s][r], fx0, fy0, 7.0f, 7.0f);
            uint32_t m = __ballot_sync(0xffffffffu, hit);
            while (m) {
                const int j = __ffs(m) - 1;
                m &= m - 1;
                const int rr = c0 + j;
                const float4 a = s_q0[s][rr];
                const float4 q = s_q1[s][rr];
                bool cb0 = false, cb1 = false; // flashsplat: this pixel votes for this record
                float uc0 = 0.0f, uc1 = 0.0f;
                uint32_t gid = 0;
                bool a0 = (live & 1u) != 0u, a1 = (live & 2u) != 0u;
                if (FLASH) {
                    gid = __float_as_uint(s_q2[s][rr].w);
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
                const float4 c = s_q2[s][rr];
                const float dx = __fadd_rn(a.x, -pixfx);
                const f2 dy = add2(bc2(a.y), npy);
                const f2 t_zz = mul2(dy, mul2(dy, bc2(q.x)));
                const float t_xx = __fmul_rn(dx, a.z);
                const f2 nt_xy = mul2(dy, bc2(__fmul_rn(dx, -a.w))); // -(dy * (dx * B))
                const f2 power = fma2(fma2(bc2(dx), bc2(t_xx), t_zz), bc2(-0.5f), nt_xy);
                const f2 og = mul2(bc2(q.y), expf2(power));
                const float al0 = fminf(0.99f, lo2(og)), al1 = fminf(0.99f, hi2(og));
                a0 = a0 && !(lo2(power) > 0.0f) && !(al0 < 1.0f / 255.0f);
                a1 = a1 && !(hi2(power) > 0.0f) && !(al1 < 1.0f / 255.0f);
                const f2 T = pk2(Ta, Tb);
                const f2 test_T = mul2(T, fma2(pk2(al0, al1), bc2(-1.0f), bc2(1.0f)));
                const float tt0 = lo2(test_T), tt1 = hi2(test_T);
                if (a0 && tt0 < 0.0001f) { live &= ~1u; a0 = false; }
                if (a1 && tt1 < 0.0001f) { live &= ~2u; a1 = false; }
                const f2 al = pk2(a0 ? al0 : 0.0f, a1 ? al1 : 0.0f);
                const f2 C0 = fma2(T, mul2(al, bc2(c.x)), pk2(C0a, C0b));
                const f2 C1 = fma2(T, mul2(al, bc2(c.y)), pk2(C1a, C1b));
                const f2 C2 = fma2(T, mul2(al, bc2(c.z)), pk2(C2a, C2b));
                const f2 Wt = fma2(T, al, pk2(Wa, Wb));
                const f2 D = fma2(T, mul2(al, bc2(q.z)), pk2(Da, Db));
                C0a = lo2(C0); C0b = hi2(C0); C1a = lo2(C1); C1b = hi2(C1); C2a = lo2(C2); C2b = hi2(C2);
                Wa = lo2(Wt); Wb = hi2(Wt); Da = lo2(D); Db = hi2(D);
                if (FLASH) {
                    const f2 uc = mul2(al, T); // FSR forward.cu:398
                    cb0 = a0 && ok0; cb1 = a1 && ok1;
                    uc0 = lo2(uc); uc1 = hi2(uc);
                }
                Ta = a0 ? tt0 : Ta;
                Tb = a1 ? tt1 : Tb;
                const uint32_t pos = (uint32_t)(b * RF_BATCH + rr + 1);
                if (a0) last0 = pos;
                if (a1) last1 = pos;
                if (FLASH) {
                    // per-(label, Gaussian) vote: one atomic per warp when the label is uniform
                    const uint32_t cm = __ballot_sync(0xffffffffu, cb0 || cb1);
                    if (cm) {
                        const int first = __ffs(cm) - 1;
                        const int mine = cb0 ? obj0 : obj1;
                        const int oid0 = __shfl_sync(0xffffffffu, mine, first);
                        const uint32_t diff =
                            __ballot_sync(0xffffffffu, (cb0 && obj0 != oid0) || (cb1 && obj1 != oid0));
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
            }
        }
        // all lanes are through with stage s (the vote orders their reads before the refill)
        const bool all_done = RF_WARPS == 1 ? (__all_sync(0xffffffffu, live == 0u) != 0)
                                            : (__syncthreads_and(live == 0u) != 0);
        if (all_done) {
            // a bulk copy for batch b+1 may still be in flight into this CTA's shared memory:
            // it must land before the CTA exits and the memory is handed to another CTA.
            if (b + 1 < nb && leader) mbar_wait(&s_full[(b + 1) & 1], (uint32_t)(((b + 1) >> 1) & 1));
            break;
        }
    }

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

int launch_render_fwd(const ws_view *v, int P, GeomState &gs, BinningState &bs, ImageState &is,
                      float *out_color, float *out_depth, float *out_alpha, const ws_flash *fl,
                      cudaStream_t st)
{
    (void)gs;
    RenderFwdParams p;
    p.ranges = is.ranges; p.q0 = bs.q0; p.q1 = bs.q1; p.q2 = bs.q2;
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
    const unsigned grid = (unsigned)(4 / RF_WARPS) * (unsigned)p.tiles_x * (unsigned)((p.H + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    p.jobs = nullptr; p.tiles_total = 0; p.job_alpha = nullptr; p.job_depth = nullptr;
    if (fl)
        render_fwd<1><<<grid, RF_THREADS, 0, st>>>(p);
    else
        render_fwd<0><<<grid, RF_THREADS, 0, st>>>(p);
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

int launch_render_jobs(const ws_view *v, const uint2 *ranges, const uint32_t *jobs, int n_jobs, int T,
                       BinningState &bs, float *job_alpha, float *job_depth, cudaStream_t st)
{
    if (n_jobs <= 0) return WS_OK;
    RenderFwdParams p;
    p.ranges = ranges; p.q0 = bs.q0; p.q1 = bs.q1; p.q2 = bs.q2;
    p.W = v->image_width; p.H = v->image_height; p.P = 0;
    p.bg = v->bg;
    p.out_color = nullptr; p.out_depth = nullptr; p.out_alpha = nullptr; p.n_contrib = nullptr;
    p.gt_mask = nullptr; p.unique_label = nullptr; p.used_count = nullptr; p.out_contrib_num = nullptr;
    p.num_obj = 0;
    p.tiles_x = (p.W + WS_BLOCK_X - 1) / WS_BLOCK_X;
    p.jobs = jobs; p.tiles_total = T; p.job_alpha = job_alpha; p.job_depth = job_depth;
    render_fwd<2><<<(unsigned)(4 / RF_WARPS) * (unsigned)n_jobs, RF_THREADS, 0, st>>>(p);
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

} // namespace ws
