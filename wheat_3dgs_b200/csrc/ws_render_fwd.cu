// ws_render_fwd.cu -- forward alpha blending (replaces renderCUDA, DGR forward.cu:261-381, and
// the flashsplat variant FSR forward.cu:269-418).
//
// One CTA per 16x16 tile, 256 threads; warp w owns the 8x4 pixel block
// (x0 + (w&1)*8 .. +7, y0 + (w>>1)*4 .. +3), lane l the pixel (l&7, l>>3) of it.
//
// Staging: the tile's instance records were gathered into tile order by identify_ranges_pack,
// so a batch of 256 records is three contiguous 4 KB runs (q0,q1,q2).  Thread 0 brings batch
// b+1 in with three TMA bulk copies (cp.async.bulk ... mbarrier::complete_tx::bytes) into the
// other half of a double buffer while all warps blend batch b; consumers wait on the
// mbarrier's phase parity.  No thread issues a per-record global load.
//
// Sub-tile culling: before the per-pixel loop each lane tests one record of a 32-record chunk
// against the warp's 8x4 pixel rectangle: the exact minimum of the quadratic form
// q = 1/2(A dx^2 + C dy^2) + B dx dy over the rectangle (a convex minimum lies on the two edges
// facing the centre) is compared with tau = ln(255*opacity) plus a rounding margin.  Records
// that cannot reach alpha >= 1/255 anywhere in the rectangle are dropped by a ballot; the warp
// then walks only the set bits.  Dropped records would have been skipped by every pixel's own
// test (forward.cu:337-347), so results are unchanged; the surviving pairs are evaluated with
// exactly the reference's float operations (FMA contraction read from its SASS), in the same
// front-to-back order, including `contributor` positions, so n_contrib is the reference's.
//
// Early termination: a pixel stops when T would fall below 1e-4 (forward.cu:348-353); a warp
// leaves a batch when all 32 lanes are done (warp vote); the CTA stops fetching when
// __syncthreads_count says all 256 are done (forward.cu:305-308).
#include "ws_common.cuh"

namespace ws {

constexpr int RF_BATCH = 256;

struct RenderFwdParams {
    const uint2 *ranges;
    const float4 *q0, *q1, *q2;
    int W, H, P;
    const float *bg;
    float *out_color, *out_depth, *out_alpha;
    uint32_t *n_contrib;
    // flashsplat
    const float *gt_mask;
    const int32_t *unique_label;
    float *used_count;
    int32_t *out_contrib_num;
    int num_obj;
};

template <bool FLASH>
__global__ void __launch_bounds__(256) render_fwd(const RenderFwdParams p)
{
    __shared__ __align__(128) float4 s_q0[2][RF_BATCH];
    __shared__ __align__(128) float4 s_q1[2][RF_BATCH];
    __shared__ __align__(128) float4 s_q2[2][RF_BATCH];
    __shared__ __align__(8) uint64_t s_full[2];

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int tiles_x = gridDim.x;
    const uint32_t tile = blockIdx.y * tiles_x + blockIdx.x;
    const int wx0 = blockIdx.x * WS_BLOCK_X + (w & 1) * 8;
    const int wy0 = blockIdx.y * WS_BLOCK_Y + (w >> 1) * 4;
    const int px = wx0 + (lane & 7), py = wy0 + (lane >> 3);
    const bool inside = px < p.W && py < p.H;
    const uint32_t pix_id = (uint32_t)p.W * py + px;
    const float pixfx = (float)px, pixfy = (float)py;
    const float fx0 = (float)wx0, fy0 = (float)wy0;

    const uint2 range = p.ranges[tile];
    const int n = (int)(range.y - range.x);
    const int nb = (n + RF_BATCH - 1) / RF_BATCH;

    if (tid == 0) {
        mbar_init(&s_full[0], 1);
        mbar_init(&s_full[1], 1);
        mbar_fence_init();
    }
    __syncthreads();

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

    int obj_id = 0;
    if (FLASH) {
        if (inside) obj_id = __float2int_rz(p.gt_mask[pix_id]);
    }
    const bool obj_ok = FLASH ? (obj_id >= 0 && obj_id <= p.num_obj) : true;

    bool done = !inside;
    float T = 1.0f, C0 = 0.0f, C1 = 0.0f, C2 = 0.0f, weight = 0.0f, D = 0.0f;
    uint32_t last_contributor = 0;

    if (nb > 0 && tid == 0) issue(0);
    for (int b = 0; b < nb; b++) {
        const int s = b & 1;
        const int cnt = min(RF_BATCH, n - b * RF_BATCH);
        if (b + 1 < nb && tid == 0) issue(b + 1); // stage (b+1)&1 was released by the barrier below
        mbar_wait(&s_full[s], (uint32_t)((b >> 1) & 1));

        for (int c0 = 0; c0 < cnt; c0 += 32) {
            if (__all_sync(0xffffffffu, done)) break;
            const int r = c0 + lane;
            bool hit = false;
            if (r < cnt) hit = rect_hit(s_q0[s][r], s_q1[s][r], fx0, fy0, 7.0f, 3.0f);
            uint32_t m = __ballot_sync(0xffffffffu, hit);
            while (m) {
                const int j = __ffs(m) - 1;
                m &= m - 1;
                const int rr = c0 + j;
                const float4 a = s_q0[s][rr];
                const float4 q = s_q1[s][rr];
                bool contrib = false;
                float uc_val = 0.0f;
                uint32_t gid = 0;
                if (FLASH) gid = __float_as_uint(s_q2[s][rr].w);
                bool gate = !done;
                if (FLASH) {
                    if (p.unique_label != nullptr) { // FSR forward.cu:357-362
                        const int uid = __ldg(p.unique_label + gid);
                        if (uid >= 0 && obj_id != uid) gate = false;
                    }
                }
                if (gate) {
                    // DGR forward.cu:333-347, op order from the reference SASS
                    const float dx = __fadd_rn(a.x, -pixfx), dy = __fadd_rn(a.y, -pixfy);
                    const float t_zz = __fmul_rn(dy, __fmul_rn(dy, q.x));
                    const float t_xx = __fmul_rn(dx, a.z);
                    const float t_xy = __fmul_rn(dy, __fmul_rn(dx, a.w));
                    const float power = __fmaf_rn(__fmaf_rn(dx, t_xx, t_zz), -0.5f, -t_xy);
                    if (!(power > 0.0f)) {
                        const float alpha = fminf(0.99f, __fmul_rn(q.y, expf(power)));
                        if (!(alpha < 1.0f / 255.0f)) {
                            const float test_T = __fmul_rn(T, __fadd_rn(1.0f, -alpha));
                            if (test_T < 0.0001f) {
                                done = true;
                            } else {
                                const float4 c = s_q2[s][rr];
                                C0 = __fmaf_rn(T, __fmul_rn(alpha, c.x), C0);
                                C1 = __fmaf_rn(T, __fmul_rn(alpha, c.y), C1);
                                C2 = __fmaf_rn(T, __fmul_rn(alpha, c.z), C2);
                                weight = __fmaf_rn(T, alpha, weight);
                                D = __fmaf_rn(T, __fmul_rn(alpha, q.z), D);
                                if (FLASH) {
                                    contrib = obj_ok;
                                    uc_val = __fmul_rn(alpha, T); // FSR forward.cu:398
                                }
                                T = test_T;
                                last_contributor = (uint32_t)(b * RF_BATCH + rr + 1);
                            }
                        }
                    }
                }
                if (FLASH) {
                    // per-(label, Gaussian) vote: one atomic per warp when the label is uniform
                    const uint32_t cm = __ballot_sync(0xffffffffu, contrib);
                    if (cm) {
                        const int first = __ffs(cm) - 1;
                        const int oid0 = __shfl_sync(0xffffffffu, obj_id, first);
                        const uint32_t diff = __ballot_sync(0xffffffffu, contrib && obj_id != oid0);
                        if (diff == 0) {
                            float v = contrib ? uc_val : 0.0f;
#pragma unroll
                            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                            if (lane == 0) atomicAdd(p.used_count + (size_t)oid0 * p.P + gid, v);
                        } else if (contrib) {
                            atomicAdd(p.used_count + (size_t)obj_id * p.P + gid, uc_val);
                        }
                    }
                }
            }
        }
        const int num_done = __syncthreads_count(done);
        if (num_done == 256) {
            // a bulk copy for batch b+1 may still be in flight into this CTA's shared memory:
            // it must land before the CTA exits and the memory is handed to another CTA.
            if (b + 1 < nb && tid == 0) mbar_wait(&s_full[(b + 1) & 1], (uint32_t)(((b + 1) >> 1) & 1));
            break;
        }
    }

    if (inside) {
        const size_t HW = (size_t)p.H * p.W;
        p.n_contrib[pix_id] = last_contributor;
        // forward.cu:376-377: C + T * bg (contracted to one FMA in the reference build)
        p.out_color[pix_id] = __fmaf_rn(__ldg(p.bg + 0), T, C0);
        p.out_color[HW + pix_id] = __fmaf_rn(__ldg(p.bg + 1), T, C1);
        p.out_color[2 * HW + pix_id] = __fmaf_rn(__ldg(p.bg + 2), T, C2);
        p.out_alpha[pix_id] = weight;
        p.out_depth[pix_id] = D;
        if (FLASH) p.out_contrib_num[pix_id] = (int32_t)last_contributor;
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
    dim3 grid((p.W + WS_BLOCK_X - 1) / WS_BLOCK_X, (p.H + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    if (fl)
        render_fwd<true><<<grid, 256, 0, st>>>(p);
    else
        render_fwd<false><<<grid, 256, 0, st>>>(p);
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

} // namespace ws
