// ws_stage.cuh -- record staging shared by the forward and backward blend kernels (a warp = one 8x8 quadrant of a
// 16x16 tile; a CTA = the tile or half of it).
//
// The tile's slice of the sorted value list holds, per instance, the Gaussian id and a 4-bit quadrant mask (which
// of the tile's 8x8 blocks the Gaussian can reach with alpha >= 1/255; computed when the instance was emitted,
// ws_preprocess.cu).  The list is walked in batches of 32 instances per warp through a ring of NSTAGE stages:
//   load    every lane reads ONE value of a batch (coalesced 128-byte loads per warp), one batch ahead of ...
//   issue   ... the warp dropping the instances whose mask is empty (ballot + popc), writing the survivors'
//           (value, list position) pairs compacted into its segment of the stage, and bringing the survivors'
//           64-byte Gaussian rows in with TMA gather4 copies (cp.async.bulk.tensor.2d ... tile::gather4, four rows
//           per instruction, complete_tx on the stage's mbarrier), NSTAGE-1 batches ahead of ...
//   blend   ... all four warps waiting on the stage's mbarrier and walking its records.
// The reference gathers the same rows with one dependent global load chain per thread and record
// (DGR forward.cu:321-324,359,361) behind a block-wide barrier.
#pragma once
#include "ws_common.cuh"

namespace ws {

constexpr int ST_SEG = 32; // instances a warp stages per batch (one per lane)

// WARPS warps per CTA (4: the whole tile, 2: a 16x8 half tile), NSTAGE ring slots of WARPS*32 instances: rows are
// requested NSTAGE-1 batches before they are blended
template <int WARPS, int NSTAGE>
struct StageRing {
    static constexpr int BATCH = ST_SEG * WARPS;
    GRec rec[NSTAGE][BATCH];        // 64-byte rows, a gather4 lands 4 of them (256 bytes, 128-byte aligned)
    uint2 meta[NSTAGE][BATCH];      // {value = id | mask << 28, position in the tile's list}
    uint32_t cnt[NSTAGE][WARPS];    // survivors per segment
    uint64_t full[NSTAGE];          // mbarriers: WARPS arrivals + the gathers' bytes per phase

    __device__ __forceinline__ void init()
    {
        if (threadIdx.x == 0) {
#pragma unroll
            for (int s = 0; s < NSTAGE; s++) mbar_init(&full[s], WARPS);
            mbar_fence_init();
        }
        __syncthreads();
    }

    // the instance of batch b this lane stages: list position b*BATCH + 32 wl + lane (0 = nothing: empty mask)
    static __device__ __forceinline__ uint32_t load(const uint32_t *__restrict__ vals, int n, int nb, int b, uint32_t wl, int lane)
    {
        const int i = b * BATCH + (int)wl * ST_SEG + lane;
        return (b >= 0 && b < nb && i < n) ? __ldg(vals + i) : 0u;
    }

    // this warp's share of batch b into ring slot s: the instances whose mask meets `quads` (the CTA's quadrants)
    __device__ __forceinline__ void issue(const TensorMap *tmap, int s, int b, uint32_t v, uint32_t quads, uint32_t wl, int lane)
    {
        const uint32_t base = wl * ST_SEG;
        const bool keep = ((v >> WS_ID_BITS) & quads) != 0u;
        const uint32_t bal = __ballot_sync(0xffffffffu, keep);
        const uint32_t n_keep = __popc(bal);
        if (keep) meta[s][base + __popc(bal & ((1u << lane) - 1u))] = make_uint2(v, (uint32_t)(b * BATCH) + base + lane);
        if (lane == 0) cnt[s][wl] = n_keep;
        __syncwarp();
        const uint32_t nops = (n_keep + 3u) >> 2;
        if (lane == 0) {
            if (nops) mbar_expect_tx(&full[s], nops * (uint32_t)(4 * sizeof(GRec)));
            else mbar_arrive(&full[s]);
        }
        if ((uint32_t)lane < nops) {
            const uint32_t e = 4u * lane, last = n_keep - 1u;
            const int r0 = (int)(meta[s][base + e].x & WS_ID_MASK);
            const int r1 = (int)(meta[s][base + min(e + 1u, last)].x & WS_ID_MASK);
            const int r2 = (int)(meta[s][base + min(e + 2u, last)].x & WS_ID_MASK);
            const int r3 = (int)(meta[s][base + min(e + 3u, last)].x & WS_ID_MASK);
            tma_gather4(&rec[s][base + e], tmap, r0, r1, r2, r3, &full[s]);
        }
    }

    __device__ __forceinline__ void wait(int k) { mbar_wait(&full[k % NSTAGE], (uint32_t)((k / NSTAGE) & 1)); }
};

} // namespace ws
