// ws_stage.cuh -- record staging shared by the forward and backward blend kernels.
//
// A warp blends one 8x8 quadrant of a 16x16 tile and stages its own records: no block-wide barrier in the walk.
// The tile's slice of the sorted value list holds, per instance, the Gaussian id and a 4-bit quadrant mask (which
// of the tile's 8x8 blocks the Gaussian can reach with alpha >= 1/255; computed when the instance was emitted,
// ws_preprocess.cu).  The warp walks the list 64 instances at a time through a private ring of NSTAGE stages:
//   load    every lane reads two values of a batch (coalesced 128-byte loads), one batch ahead of ...
//   issue   ... the warp dropping the instances whose mask lacks its quadrant (ballot + popc), writing the
//           survivors' (value, list position + 1) pairs compacted into the stage, and bringing the survivors'
//           64-byte Gaussian rows in with TMA gather4 copies (cp.async.bulk.tensor.2d ... tile::gather4, four rows
//           per instruction, complete_tx on the stage's mbarrier), NSTAGE-1 batches ahead of ...
//   blend   ... the warp waiting on the stage's mbarrier and walking its records in order.
// The reference gathers the same rows with one dependent global load chain per thread and record
// (DGR forward.cu:321-324,359,361) behind a block-wide barrier, for all 256 pixels of the tile.
#pragma once
#include "ws_common.cuh"

namespace ws {

#ifndef WS_ST_SEG
#define WS_ST_SEG 64
#endif
constexpr int ST_SEG = WS_ST_SEG; // instances per batch: 64 (two per lane) or 32 (one per lane)
static_assert(ST_SEG == 64 || ST_SEG == 32, "a batch is one or two values per lane");

template <int NSTAGE>
struct __align__(128) WarpRing {
    GRec rec[NSTAGE][ST_SEG];   // 64-byte rows; a gather4 lands 4 of them (256 bytes, 128-byte aligned)
    uint2 meta[NSTAGE][ST_SEG]; // {Gaussian id, position in the tile's list + 1}
    uint64_t full[NSTAGE];      // mbarriers: one arrival (lane 0) + the gathers' bytes per phase

    __device__ __forceinline__ void init(int lane)
    {
        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < NSTAGE; s++) mbar_init(&full[s], 1);
            mbar_fence_init();
        }
        __syncwarp();
    }

    // the instances of batch b this lane looks at: list positions 64 b + lane and + 32 (0 = nothing: empty mask)
    static __device__ __forceinline__ uint2 load(const uint32_t *__restrict__ vals, int n, int nb, int b, int lane)
    {
        const int i = b * ST_SEG + lane;
        uint2 v = make_uint2(0u, 0u);
        if (b >= 0 && b < nb) {
            if (i < n) v.x = __ldg(vals + i);
            if (ST_SEG == 64 && i + 32 < n) v.y = __ldg(vals + i + 32);
        }
        return v;
    }

    // batch b into ring slot s: the instances whose mask has bit `quad`, in list order.  Returns how many
    // (warp-uniform).
    __device__ __forceinline__ uint32_t issue(const TensorMap *tmap, int s, int b, uint2 v, uint32_t quad, int lane)
    {
        const bool k0 = ((v.x >> (WS_ID_BITS + quad)) & 1u) != 0u, k1 = ((v.y >> (WS_ID_BITS + quad)) & 1u) != 0u;
        const uint32_t b0 = __ballot_sync(0xffffffffu, k0), b1 = __ballot_sync(0xffffffffu, k1);
        const uint32_t lt = (1u << lane) - 1u, n0 = __popc(b0), n_keep = n0 + __popc(b1);
        const uint32_t pos = (uint32_t)(b * ST_SEG + lane) + 1u;
        if (k0) meta[s][__popc(b0 & lt)] = make_uint2(v.x & WS_ID_MASK, pos);
        if (k1) meta[s][n0 + __popc(b1 & lt)] = make_uint2(v.y & WS_ID_MASK, pos + 32u);
        __syncwarp();
        const uint32_t nops = (n_keep + 3u) >> 2;
        if (lane == 0) {
            if (nops) mbar_expect_tx(&full[s], nops * (uint32_t)(4 * sizeof(GRec)));
            else mbar_arrive(&full[s]);
        }
        if ((uint32_t)lane < nops) {
            const uint32_t e = 4u * lane, last = n_keep - 1u;
            const int r0 = (int)meta[s][e].x;
            const int r1 = (int)meta[s][min(e + 1u, last)].x;
            const int r2 = (int)meta[s][min(e + 2u, last)].x;
            const int r3 = (int)meta[s][min(e + 3u, last)].x;
            tma_gather4(&rec[s][e], tmap, r0, r1, r2, r3, &full[s]);
        }
        return n_keep;
    }

    __device__ __forceinline__ void wait(int k) { mbar_wait(&full[k % NSTAGE], (uint32_t)((k / NSTAGE) & 1)); }
};

constexpr int ST_CNT_BITS = 7; // survivors per requested batch (<= 64) queue up in one register, oldest lowest

} // namespace ws
