// ws_binning.cu -- identify_ranges_pack: one pass over the sorted instance list that
//   (a) writes the per-tile [start,end) ranges (replaces identifyTileRanges,
//       DGR rasterizer_impl.cu:116-138, after the cudaMemset of :311), and
//   (b) gathers each instance's per-Gaussian record into tile order (q0,q1,q2), so the blend
//       kernels read contiguous 16-byte records with TMA bulk copies instead of gathering
//       through point_list inside the pixel loop (DGR forward.cu:321-324,359,361).
// Algorithmic HBM bytes: 8R read (tile key + value; 12R with 64-bit keys) + 48R written; the 48-byte gathers hit L2
// (the per-Gaussian arrays are 48 B x P = 48 MB at 1 M Gaussians, L2 is 126 MB).
#include "ws_common.cuh"

namespace ws {

__device__ __forceinline__ uint32_t tile_of(uint64_t k) { return (uint32_t)(k >> 32); }
__device__ __forceinline__ uint32_t tile_of(uint32_t k) { return k; }

template <typename K>
__global__ void __launch_bounds__(256) identify_ranges_pack(size_t R, const K *__restrict__ keys,
                                                            const uint32_t *__restrict__ vals,
                                                            const float4 *__restrict__ g0,
                                                            const float4 *__restrict__ g1,
                                                            const float4 *__restrict__ g2,
                                                            uint2 *__restrict__ ranges, float4 *__restrict__ q0,
                                                            float4 *__restrict__ q1, float4 *__restrict__ q2)
{
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= R) return;
    const uint32_t currtile = tile_of(keys[idx]);
    if (idx == 0) {
        ranges[currtile].x = 0;
    } else {
        const uint32_t prevtile = tile_of(keys[idx - 1]);
        if (currtile != prevtile) {
            ranges[prevtile].y = (uint32_t)idx;
            ranges[currtile].x = (uint32_t)idx;
        }
    }
    if (idx == R - 1) ranges[currtile].y = (uint32_t)R;
    const uint32_t g = vals[idx];
    const float4 a = __ldg(g0 + g), b = __ldg(g1 + g);
    float4 c = __ldg(g2 + g);
    c.w = __uint_as_float(g);
    q0[idx] = a;
    q1[idx] = b;
    q2[idx] = c;
}

int launch_ranges_pack(size_t R, int T, const uint64_t *keys, const uint32_t *tkeys, const uint32_t *vals,
                       GeomState &gs, BinningState &bs, ImageState &is, cudaStream_t st, bool debug)
{
    WS_CUDA_CHECK(cudaMemsetAsync(is.ranges, 0, (size_t)T * sizeof(uint2), st));
    if (R == 0) return WS_OK;
    if (tkeys)
        identify_ranges_pack<uint32_t><<<(unsigned)((R + 255) / 256), 256, 0, st>>>(
            R, tkeys, vals, gs.g0, gs.g1, gs.g2, is.ranges, bs.q0, bs.q1, bs.q2);
    else
        identify_ranges_pack<uint64_t><<<(unsigned)((R + 255) / 256), 256, 0, st>>>(
            R, keys, vals, gs.g0, gs.g1, gs.g2, is.ranges, bs.q0, bs.q1, bs.q2);
    WS_LAUNCH_CHECK(st, debug);
    return WS_OK;
}

} // namespace ws
