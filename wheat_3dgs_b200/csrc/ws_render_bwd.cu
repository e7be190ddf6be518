// ws_render_bwd.cu -- backward of the alpha blending (replaces renderCUDA backward,
// DGR backward.cu:415-601; flashsplat gate FSR backward.cu:455-462,529-536).
//
// Same warp / lane mapping (warp = 8x8 pixels, two pixels per lane, packed
// f32x2 arithmetic) and record staging (every warp compacts the tile's value list by its quadrant's mask bit and
// gathers the surviving Gaussians' rows with TMA gather4 copies through a private mbarrier ring, ws_stage.cuh)
// as the forward (ws_render_fwd.cu), walking the records back to front from the largest
// last-contributor position in the warp's 8x8 block (records behind it are never fetched; the reference
// walks them with `contributor >= last_contributor -> continue`).  alpha is recomputed with the
// forward's exact operations, so the set of contributing pairs is the forward's.
//
// Per-pixel recurrence.  The reference keeps, per pixel, the colour / depth / alpha accumulated
// behind the current record (accum_rec, last_color, last_alpha ..., backward.cu:538-575) and
// forms dL_dalpha = T * sum_ch (c_ch - accum_rec_ch) dL_ch + ... .  With
//   dot_i = c_i . dL_dC + depth_i dL_dD + dL_dA   and   R_i = sum_{j behind i} alpha_j T_j dot_j
// the same quantity is  dL_dalpha_i = T_i dot_i - (R_i + T_final bg . dL_dC) / (1 - alpha_i):
// one scalar R per pixel instead of ten, accumulated back to front from small to large terms (no
// cancellation).  T_i is recovered as in the reference, T_i = T_{i+1} / (1 - alpha_i).
//
// Gradient accumulation.  The reference issues 10 global float atomics per contributing
// (pixel, Gaussian) pair (backward.cu:555,563,589-598).  Here the 64 pixels of a warp look at
// the same record at the same time: a lane adds its two pixels, the 10 partial sums are reduced
// across the warp with transposing shuffle butterflies (an 8-slot tree, 9 shuffles, and a 2-slot
// tree, 5 shuffles) and 10 lanes issue ONE red.global.add.f32 each into a 48-byte per-Gaussian
// accumulator row acc[g][12] (2 sectors).  (A shared-memory transpose was tried instead of the
// shuffles: fewer instructions, but the SM-wide shared-memory pipe made it 5 % slower.)  Float
// summation order differs from the reference's (non-deterministic) atomic order: gradients are
// in the 1e-3 relative tier.
//
// What is summed.  With g = G * dL_dalpha of a pair, every geometric gradient of the Gaussian is
// a per-Gaussian constant times a pixel moment of g:  dL_dopacity = S g,
// dL_dconic = -0.5 o (S g dx^2, S g dx dy, S g dy^2),  dL_dmean2D = -0.5 o (W (A S g dx + B S g dy),
// H (C S g dy + B S g dx))  (backward.cu:583-598 with the constants factored out of the sums), so
// the kernel accumulates {dL_dcolor rgb, dL_ddepth, S g dx, S g dy, S g dx^2, S g dx dy, S g dy^2, S g}
// and preprocess_bwd applies the constants once per Gaussian.  A lane's two pixels share dx: it sums g, g dy and
// g dy^2 over them first and multiplies by dx afterwards; the colour / depth sums run on (r,g) and (b,depth)
// pairs of the pixel gradients so that one FMUL2 + one FFMA2 gives two of them summed over both pixels.
#include "ws_stage.cuh"

namespace ws {

#ifndef WS_RB_WARPS
#define WS_RB_WARPS 2
#endif
#ifndef WS_RB_NSTAGE
#define WS_RB_NSTAGE 2
#endif
// CTAs per SM the register allocation is sized for.  Shared memory allows 11 (11 x (18688 + 1024) B = 211 KB of 228 KB;
// 12 would need 236.5 KB), so 11 is what is asked for: ptxas then has 88 registers per thread to work with instead of
// 80 (it uses 79; 76 at 12).  A/B on config 2, two rounds each (profiles/r03_ab_rb_minb.txt): 12: 0.9056 / 0.9072 ms,
// 11: 0.9045 / 0.9057 ms, 10: 0.9294 / 0.9298 ms.
#ifndef WS_RB_MINB
#define WS_RB_MINB 11
#endif
constexpr int RB_WARPS = WS_RB_WARPS;   // 8x8 pixel blocks (warps) per CTA: 2 = a 16x8 half tile, 4 = the tile
constexpr int RB_THREADS = 32 * RB_WARPS;
constexpr int RB_NSTAGE = WS_RB_NSTAGE;
static_assert(RB_NSTAGE * ST_CNT_BITS <= 32 && (RB_NSTAGE & (RB_NSTAGE - 1)) == 0, "ring depth: a power of two, counts in one register");
#ifndef WS_RB_UNROLL
#define WS_RB_UNROLL 2
#endif
constexpr int RB_UNROLL = WS_RB_UNROLL; // records per trip of the blend loop
using BwdRing = WarpRing<RB_NSTAGE>;

struct RenderBwdParams {
    const uint2 *ranges;
    const uint32_t *vals; // sorted values: Gaussian id | quadrant mask << 28
    int W, H, tiles_x;
    const float *bg;
    const float *out_alpha;
    const uint32_t *n_contrib;
    const float *dL_dcolor, *dL_ddepth, *dL_dalpha;
    float *acc; // [P][12]
    const float *gt_mask;
    const int32_t *unique_label;
    ExpConsts ec;
};

template <bool FLASH>
__global__ void __launch_bounds__(RB_THREADS, WS_RB_MINB) render_bwd(const __grid_constant__ TensorMap tmap, const RenderBwdParams p)
{
    __shared__ BwdRing rings[RB_WARPS];

    const int lane = threadIdx.x & 31;
    const uint32_t wl = threadIdx.x >> 5;     // warp within the CTA
    BwdRing &sm = rings[wl];
    // the CTAs of a tile are adjacent in launch order
    const uint32_t tile = blockIdx.x / (4 / RB_WARPS);
    const uint32_t w = (blockIdx.x % (4 / RB_WARPS)) * RB_WARPS + wl; // quadrant of the tile
    const int tile_y = (int)(tile / (uint32_t)p.tiles_x), tile_x = (int)(tile - (uint32_t)tile_y * p.tiles_x);
    const int wx0 = tile_x * WS_BLOCK_X + (int)(w & 1) * 8;
    const int wy0 = tile_y * WS_BLOCK_Y + (int)(w >> 1) * 8;
    const int px = wx0 + (lane & 7), py0 = wy0 + (lane >> 3), py1 = py0 + 4;
    const bool in0 = px < p.W && py0 < p.H, in1 = px < p.W && py1 < p.H;
    const uint32_t pix0 = (uint32_t)p.W * py0 + px, pix1 = pix0 + 4u * (uint32_t)p.W;
    const float pixfx = (float)px;
    const f2 npy = pk2(-(float)py0, -(float)py1);
    const size_t HW = (size_t)p.H * p.W;

    const uint2 range = p.ranges[tile];
    // 1-based position of the last blended record (0 outside the image: nothing is active)
    const uint32_t last0 = in0 ? p.n_contrib[pix0] : 0u, last1 = in1 ? p.n_contrib[pix1] : 0u;
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, max(last0, last1));
    const int n = (int)min(wmax, range.y - range.x); // records [0, n) matter to this warp
    const int nb = (n + ST_SEG - 1) / ST_SEG;
    const uint32_t *__restrict__ vals = p.vals + range.x;
    sm.init(lane);

    // per-pixel constants, DGR backward.cu:461-496 (lo = pixel 0, hi = pixel 1)
    const float Tf0 = in0 ? (1.0f - p.out_alpha[pix0]) : 0.0f, Tf1 = in1 ? (1.0f - p.out_alpha[pix1]) : 0.0f;
    float dl0[5] = {0.f, 0.f, 0.f, 0.f, 0.f}, dl1[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (in0) {
        dl0[0] = p.dL_dcolor[pix0]; dl0[1] = p.dL_dcolor[HW + pix0]; dl0[2] = p.dL_dcolor[2 * HW + pix0];
        dl0[3] = p.dL_ddepth[pix0]; dl0[4] = p.dL_dalpha[pix0];
    }
    if (in1) {
        dl1[0] = p.dL_dcolor[pix1]; dl1[1] = p.dL_dcolor[HW + pix1]; dl1[2] = p.dL_dcolor[2 * HW + pix1];
        dl1[3] = p.dL_ddepth[pix1]; dl1[4] = p.dL_dalpha[pix1];
    }
    // packed by pixel (for c . dL) ...
    const f2 dLr = pk2(dl0[0], dl1[0]), dLg = pk2(dl0[1], dl1[1]), dLb = pk2(dl0[2], dl1[2]),
             dLd = pk2(dl0[3], dl1[3]), dLa = pk2(dl0[4], dl1[4]);
    // ... and by channel (for the colour / depth sums over the lane's two pixels), PERMUTED BY LANE so that the first two
    // butterfly stages below need no selects: lanes with bit 4 set compute (b, depth) where the others compute (r, g)
    // and vice versa -- every lane then keeps its pair "A" and sends its pair "B" -- and lanes with bit 3 set hold
    // their pairs with the halves swapped -- every lane keeps the low half and sends the high one.  Copied through
    // opaque moves: ptxas would otherwise rebuild these pairs from the ones above with moves per record.
    const bool l16 = (lane & 16) != 0, l8 = (lane & 8) != 0;
    // the same lane bits as all-ones / zero words for bitwise selects in the butterflies: a predicated select needs its
    // predicate, and with seven predicate registers ptxas rebuilt these from the thread id for every record (ten
    // instructions); a LOP3 with the mask in a register needs nothing
    // (passed through a self-shuffle: ptxas would otherwise recognise the masks as lane predicates and go back to
    // predicated selects)
    const uint32_t m16 = __shfl_sync(0xffffffffu, l16 ? 0xffffffffu : 0u, lane),
                   m8 = __shfl_sync(0xffffffffu, l8 ? 0xffffffffu : 0u, lane),
                   m4 = __shfl_sync(0xffffffffu, (lane & 4) ? 0xffffffffu : 0u, lane);
    const int ia = l16 ? 2 : 0, ib = l16 ? 0 : 2, lo_i = l8 ? 1 : 0, hi_i = l8 ? 0 : 1;
    const f2 cA0 = opaque2(pk2(dl0[ia + lo_i], dl0[ia + hi_i])), cB0 = opaque2(pk2(dl0[ib + lo_i], dl0[ib + hi_i])),
             cA1 = opaque2(pk2(dl1[ia + lo_i], dl1[ia + hi_i])), cB1 = opaque2(pk2(dl1[ib + lo_i], dl1[ib + hi_i]));
    const float bgr = __ldg(p.bg + 0), bgg = __ldg(p.bg + 1), bgb = __ldg(p.bg + 2);
    // -T_final * (bg . dL_dpixel): the background term of backward.cu:577-580 before its 1/(1-alpha)
    const f2 ntfbg = pk2(-Tf0 * (bgr * dl0[0] + bgg * dl0[1] + bgb * dl0[2]),
                         -Tf1 * (bgr * dl1[0] + bgg * dl1[1] + bgb * dl1[2]));
    int obj0 = 0, obj1 = 0;
    if (FLASH) {
        if (in0 && p.gt_mask) obj0 = __float2int_rz(p.gt_mask[pix0]);
        if (in1 && p.gt_mask) obj1 = __float2int_rz(p.gt_mask[pix1]);
    }
    // where this lane's share of the reduced sums goes: after the butterflies lane l holds, of the first eight sums
    // {r, g, b, depth, S g dx, S g dy, S g dx^2, S g dx dy}, the colour (bit 2 clear) or moment (bit 2 set) number
    // 2 * bit 4 + bit 3, and of the last two {S g dy^2, S g} number bit 4; lanes 0 mod 4 write the former, lanes 1 and
    // 17 the latter
    const bool low8 = (lane & 3) == 0;
    const bool writer = low8 || (lane & 15) == 1;
    const int my_slot = low8 ? ((lane >> 2) & 1) * 4 + ((lane >> 4) & 1) * 2 + ((lane >> 3) & 1) : 8 + (lane >> 4);
    // (the lane's accumulator base through shuffles too: ptxas otherwise rebuilds it from the parameter for every record)
    unsigned long long acc_lane = (unsigned long long)__cvta_generic_to_global(p.acc + my_slot);
    {
        const int nxt = (lane + 1) & 31;   // a round trip over two lanes: the value is the lane's own again
        unsigned lo = (unsigned)acc_lane, hi = (unsigned)(acc_lane >> 32);
        lo = __shfl_sync(0xffffffffu, __shfl_sync(0xffffffffu, lo, nxt), (lane + 31) & 31);
        hi = __shfl_sync(0xffffffffu, __shfl_sync(0xffffffffu, hi, nxt), (lane + 31) & 31);
        acc_lane = ((unsigned long long)hi << 32) | lo;
    }
    // (the exp constants through a self-shuffle as well: ptxas otherwise reloads them from the constant bank for
    // every record, two LDC per record)
    ExpConsts ec;
    ec.a = __shfl_sync(0xffffffffu, lane == 0 ? p.ec.a : 0.0f, 0);
    ec.m = __shfl_sync(0xffffffffu, lane == 0 ? p.ec.m : 0.0f, 0);
    ec.one = __shfl_sync(0xffffffffu, lane == 0 ? p.ec.one : 0.0f, 0);
    // loop-carried state as scalar halves (packed at the point of use, see ws_render_fwd.cu):
    // T = transmittance in front of the current record, R = sum over the records behind it of
    // alpha_j T_j (c_j . dL_dC + depth_j dL_dD + dL_dA)
    float Ta = Tf0, Tb = Tf1, Ra = 0.0f, Rb = 0.0f;

    // walking back to front: step `it` blends batch nb-1-it.  Ring prologue: steps 0 .. NSTAGE-2 requested,
    // the value of step NSTAGE-1 in flight
    uint2 nv;
    uint32_t cq = 0; // survivors per requested batch, ST_CNT_BITS bits each, oldest lowest
    {
        uint2 pv[RB_NSTAGE - 1];
#pragma unroll
        for (int k = 0; k < RB_NSTAGE - 1; k++) pv[k] = BwdRing::load(vals, n, nb, nb - 1 - k, lane);
        nv = BwdRing::load(vals, n, nb, nb - RB_NSTAGE, lane);
#pragma unroll
        for (int k = 0; k < RB_NSTAGE - 1; k++)
            if (k < nb) cq |= sm.issue(&tmap, k, nb - 1 - k, pv[k], w, lane) << (ST_CNT_BITS * k);
    }
    for (int it = 0; it < nb; it++) {
        const int s = it % RB_NSTAGE;
        // slot (it-1) % NSTAGE: the warp is through with it (its reads fed the arithmetic of step it-1)
        if (it + RB_NSTAGE - 1 < nb)
            cq |= sm.issue(&tmap, (it + RB_NSTAGE - 1) % RB_NSTAGE, nb - it - RB_NSTAGE, nv, w, lane) << (ST_CNT_BITS * (RB_NSTAGE - 1));
        nv = BwdRing::load(vals, n, nb, nb - 1 - it - RB_NSTAGE, lane);
        const int cnt = (int)(cq & ((1u << ST_CNT_BITS) - 1u));
        cq >>= ST_CNT_BITS;
        sm.wait(it);

#pragma unroll RB_UNROLL
        for (int rr = cnt - 1; rr >= 0; rr--) {
            const uint2 meta = sm.meta[s][rr];
            const float4 a = sm.rec[s][rr].a;
            const float4 q = sm.rec[s][rr].b;
            const float4 c = sm.rec[s][rr].c;
            const uint32_t gid = meta.x;
            bool k0 = meta.y <= last0, k1 = meta.y <= last1; // backward.cu:516-518 (meta.y: 1-based position)
            if (FLASH) {
                if (p.unique_label != nullptr) {
                    const int uid = __ldg(p.unique_label + gid);
                    if (uid >= 0 && obj0 != uid) k0 = false;
                    if (uid >= 0 && obj1 != uid) k1 = false;
                }
            }
            // alpha exactly as the forward computed it (same ops, same order)
            const float dx = __fadd_rn(a.x, -pixfx);
            const f2 dy = add2(bc2(a.y), npy);
            const f2 t_zz = mul2(dy, mul2(dy, bc2(q.x)));
            const float t_xx = __fmul_rn(dx, a.z);
            const f2 nt_xy = mul2(dy, bc2(__fmul_rn(dx, -a.w)));
            const f2 power = fma2(fma2(bc2(dx), bc2(t_xx), t_zz), bc2(-0.5f), nt_xy);
            const f2 G = expf2(power, ec);
            const f2 og = mul2(bc2(q.y), G);
            const float al0 = fminf(0.99f, lo2(og)), al1 = fminf(0.99f, hi2(og));
            k0 = k0 && !(lo2(power) > 0.0f) && !(al0 < 1.0f / 255.0f);
            k1 = k1 && !(hi2(power) > 0.0f) && !(al1 < 1.0f / 255.0f);
            // a pixel that does not contribute runs along with alpha = 0 and G = 0: every
            // sum below gets an exact zero from it and its T and R stay as they are
            const f2 al = pk2(k0 ? al0 : 0.0f, k1 ? al1 : 0.0f);
            const f2 Gm = pk2(k0 ? lo2(G) : 0.0f, k1 ? hi2(G) : 0.0f);
            const f2 om = fma2(al, bc2(-ec.one), bc2(1.0f));
            const f2 inv = pk2(rcp_approx(lo2(om)), rcp_approx(hi2(om)));
            const f2 T = mul2(pk2(Ta, Tb), inv); // backward.cu:534
            const f2 R = pk2(Ra, Rb);
            const f2 dot = fma2(dLr, bc2(c.x), fma2(dLg, bc2(c.y), fma2(dLb, bc2(c.z), fma2(dLd, bc2(q.z), dLa))));
            // dL_dalpha = T (c.dL) - (R + T_final bg.dL) / (1 - alpha)   (backward.cu:538-580)
            const f2 dopa = fma2(T, dot, mul2(fma2(R, bc2(-ec.one), ntfbg), inv));
            const f2 wgt = mul2(al, T);
            const f2 Rn = fma2(wgt, dot, R);
            Ta = k0 ? lo2(T) : Ta; Tb = k1 ? hi2(T) : Tb;
            Ra = lo2(Rn); Rb = hi2(Rn);
            // sums over the lane's two pixels
            f2 sA = fma2(bc2(hi2(wgt)), cA1, mul2(bc2(lo2(wgt)), cA0));
            const f2 sB = fma2(bc2(hi2(wgt)), cB1, mul2(bc2(lo2(wgt)), cB0));
            const f2 g = mul2(Gm, dopa);
            const f2 gy = mul2(g, dy), gyy = mul2(gy, dy);
            const float sg = lo2(g) + hi2(g), sgy = lo2(gy) + hi2(gy), sgyy = lo2(gyy) + hi2(gyy);
            const float sgx = sg * dx, sgxx = sgx * dx, sgxy = sgy * dx;
            float out;
            {
                // Transposing butterflies.  Colour / depth sums: two select-free stages (the lane-permuted constants
                // above put the pair to keep in sA and the pair to send in sB, then the half to keep low).
                sA = add2(sA, pk2(__shfl_xor_sync(0xffffffffu, lo2(sB), 16), __shfl_xor_sync(0xffffffffu, hi2(sB), 16)));
                const float col = lo2(sA) + __shfl_xor_sync(0xffffffffu, hi2(sA), 8);
                // moments {S g dx, S g dy, S g dx^2, S g dx dy}: bit 4 picks the pair, bit 3 the member
                const float mk0 = bsel(m16, sgxx, sgx), mk1 = bsel(m16, sgxy, sgy), ms0 = bsel(m16, sgx, sgxx),
                            ms1 = bsel(m16, sgy, sgxy);
                const float m0 = mk0 + __shfl_xor_sync(0xffffffffu, ms0, 16);
                const float m1 = mk1 + __shfl_xor_sync(0xffffffffu, ms1, 16);
                const float mom = bsel(m8, m1, m0) + __shfl_xor_sync(0xffffffffu, bsel(m8, m0, m1), 8);
                // the last two {S g dy^2, S g}: bit 4 picks
                float e0 = bsel(m16, sg, sgyy) + __shfl_xor_sync(0xffffffffu, bsel(m16, sgyy, sg), 16);
                e0 += __shfl_xor_sync(0xffffffffu, e0, 8);
                // bit 2: colour or moment
                float v0 = bsel(m4, mom, col) + __shfl_xor_sync(0xffffffffu, bsel(m4, col, mom), 4);
                e0 += __shfl_xor_sync(0xffffffffu, e0, 4);
                v0 += __shfl_xor_sync(0xffffffffu, v0, 2);
                e0 += __shfl_xor_sync(0xffffffffu, e0, 2);
                v0 += __shfl_xor_sync(0xffffffffu, v0, 1);
                e0 += __shfl_xor_sync(0xffffffffu, e0, 1);
                out = low8 ? v0 : e0;
            }
            if (writer)
                asm volatile("red.global.add.f32 [%0], %1;" ::"l"(acc_lane + (unsigned long long)gid * 48ull), "f"(out) : "memory");
        }
    }
}

int launch_render_bwd(const ws_view *v, int P, GeomState &gs, const uint32_t *vals, ImageState &is,
                      const float *out_alpha, const float *dL_dcolor, const float *dL_ddepth,
                      const float *dL_dalpha, const ws_flash *fl, cudaStream_t st)
{
    if ((uint32_t)P > WS_ID_MASK) {
        set_error("render: %d Gaussians exceed the 2^%u - 1 ids a sorted value can hold", P, WS_ID_BITS);
        return WS_E_UNSUPPORTED;
    }
    TensorMap tm;
    int rc = make_grec_tensor_map(&tm, gs.grec, P);
    if (rc != WS_OK) return rc;
    WS_CUDA_CHECK(cudaMemsetAsync(gs.acc, 0, (size_t)P * 12 * sizeof(float), st));
    RenderBwdParams p;
    p.ec = exp_consts_host();
    p.ranges = is.ranges; p.vals = vals;
    p.W = v->image_width; p.H = v->image_height;
    p.bg = v->bg; p.out_alpha = out_alpha; p.n_contrib = is.n_contrib;
    p.dL_dcolor = dL_dcolor; p.dL_ddepth = dL_ddepth; p.dL_dalpha = dL_dalpha;
    p.acc = gs.acc;
    p.gt_mask = fl ? fl->gt_mask : nullptr;
    p.unique_label = fl ? fl->unique_label : nullptr;
    p.tiles_x = (p.W + WS_BLOCK_X - 1) / WS_BLOCK_X;
    const unsigned grid = (unsigned)(4 / RB_WARPS) * (unsigned)p.tiles_x * (unsigned)((p.H + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    if (fl && fl->unique_label)
        render_bwd<true><<<grid, RB_THREADS, 0, st>>>(tm, p);
    else
        render_bwd<false><<<grid, RB_THREADS, 0, st>>>(tm, p);
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

} // namespace ws
