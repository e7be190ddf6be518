// ws_render_bwd.cu -- backward of the alpha blending (replaces renderCUDA backward,
// DGR backward.cu:415-601; flashsplat gate FSR backward.cu:455-462,529-536).
//
// Same tile / warp / lane mapping, record staging (TMA bulk copies, mbarrier double buffer) and
// sub-tile culling as the forward (ws_render_fwd.cu), walking the tile's records back to front
// from the largest last-contributor position in the tile (records behind it are never
// fetched; the reference walks them with `contributor >= last_contributor -> continue`).
//
// Gradient accumulation.  The reference issues 10 global float atomics per contributing
// (pixel, Gaussian) pair (backward.cu:555,563,589-598).  Here the 32 pixels of a warp look at
// the same record at the same time, so the 10 partial sums are first reduced across the warp
// with transposing shuffle butterflies (an 8-slot tree, 9 shuffles, and a 2-slot tree, 5
// shuffles: 44 instructions instead of the 62 of one 16-slot tree) and then 10 lanes issue ONE
// red.global.add.f32 each into a 48-byte per-Gaussian accumulator row acc[g][12] (2 sectors).
// (A shared-memory transpose was tried instead of the shuffles: fewer instructions, but the
// SM-wide shared-memory pipe made it 5 % slower.)  Float summation order differs from the
// reference's (non-deterministic) atomic order: gradients are in the 1e-3 relative tier.
//
// What is summed.  With g = G * dL_dalpha of a pair, every geometric gradient of the Gaussian is
// a per-Gaussian constant times a pixel moment of g:  dL_dopacity = S g,
// dL_dconic = -0.5 o (S g dx^2, S g dx dy, S g dy^2),  dL_dmean2D = -0.5 o (W (A S g dx + B S g dy),
// H (C S g dy + B S g dx))  (backward.cu:583-598 with the constants factored out of the sums), so
// the kernel accumulates {dL_dcolor rgb, dL_ddepth, S g dx, S g dy, S g dx^2, S g dx dy, S g dy^2, S g}
// and preprocess_bwd applies the constants once per Gaussian.  The colour/depth/alpha
// recurrences (backward.cu:538-575) are written as x += last_alpha * (y - x).
#include "ws_common.cuh"

namespace ws {

constexpr int RB_BATCH = 256;

struct RenderBwdParams {
    const uint2 *ranges;
    const float4 *q0, *q1, *q2;
    int W, H;
    const float *bg;
    const float *out_alpha;
    const uint32_t *n_contrib;
    const float *dL_dcolor, *dL_ddepth, *dL_dalpha;
    float *acc; // [P][12]
    const float *gt_mask;
    const int32_t *unique_label;
};

template <bool FLASH>
__global__ void __launch_bounds__(256) render_bwd(const RenderBwdParams p)
{
    __shared__ __align__(128) float4 s_q0[2][RB_BATCH];
    __shared__ __align__(128) float4 s_q1[2][RB_BATCH];
    __shared__ __align__(128) float4 s_q2[2][RB_BATCH];
    __shared__ __align__(8) uint64_t s_full[2];
    __shared__ uint32_t s_wmax[8];
                {
                    // transposing butterfly over 8 slots (lane l ends with slot l>>2 summed over the warp)
                    float v[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
                    float e0 = vc.x, e1 = vc.y; // and over the last 2 slots
                    {
                        const bool hi = (lane & 16) != 0;
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const float send = hi ? v[i] : v[i + 4];
                            const float keep = hi ? v[i + 4] : v[i];
                            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                        }
                        const float send = hi ? e0 : e1;
                        const float keep = hi ? e1 : e0;
                        e0 = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                    }
                    {
                        const bool hi = (lane & 8) != 0;
#pragma unroll
                        for (int i = 0; i < 2; i++) {
                            const float send = hi ? v[i] : v[i + 2];
                            const float keep = hi ? v[i + 2] : v[i];
                            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                        }
                        e0 += __shfl_xor_sync(0xffffffffu, e0, 8);
                    }
                    {
                        const bool hi = (lane & 4) != 0;
                        const float send = hi ? v[0] : v[1];
                        const float keep = hi ? v[1] : v[0];
                        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                        e0 += __shfl_xor_sync(0xffffffffu, e0, 4);
                    }
                    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
                    e0 += __shfl_xor_sync(0xffffffffu, e0, 2);
                    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
                    e0 += __shfl_xor_sync(0xffffffffu, e0, 1);
                    // lane l, l % 4 == 0: slot ((l>>4)&1)*4 + ((l>>3)&1)*2 + ((l>>2)&1) of v; lanes 1 / 17: slots 8 / 9
                    const int slot8 = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
                    if ((lane & 3) == 0) atomicAdd(p.acc + (size_t)gid * 12 + slot8, v[0]);
                    else if ((lane & 15) == 1) atomicAdd(p.acc + (size_t)gid * 12 + 8 + (lane >> 4), e0);
                }
            }
        }
        __syncthreads(); // everyone is done with stage s before it is refilled
    }
}

int launch_render_bwd(const ws_view *v, int P, GeomState &gs, BinningState &bs, ImageState &is,
                      const float *out_alpha, const float *dL_dcolor, const float *dL_ddepth,
                      const float *dL_dalpha, const ws_flash *fl, cudaStream_t st)
{
    WS_CUDA_CHECK(cudaMemsetAsync(gs.acc, 0, (size_t)P * 12 * sizeof(float), st));
    RenderBwdParams p;
    p.ranges = is.ranges; p.q0 = bs.q0; p.q1 = bs.q1; p.q2 = bs.q2;
    p.W = v->image_width; p.H = v->image_height;
    p.bg = v->bg; p.out_alpha = out_alpha; p.n_contrib = is.n_contrib;
    p.dL_dcolor = dL_dcolor; p.dL_ddepth = dL_ddepth; p.dL_dalpha = dL_dalpha;
    p.acc = gs.acc;
    p.gt_mask = fl ? fl->gt_mask : nullptr;
    p.unique_label = fl ? fl->unique_label : nullptr;
    dim3 grid((p.W + WS_BLOCK_X - 1) / WS_BLOCK_X, (p.H + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    if (fl && fl->unique_label)
        render_bwd<true><<<grid, 256, 0, st>>>(p);
    else
        render_bwd<false><<<grid, 256, 0, st>>>(p);
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

} // namespace ws
