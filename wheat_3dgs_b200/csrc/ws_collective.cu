// ws_collective.cu -- the exchange step of data-parallel training: the SUM over ranks of the
// 62*P-float gradient bucket the backward wrote, done in place over NVLink/NVSwitch by ONE kernel on the
// backward's own stream instead of a hand-off to NCCL's stream and back.
//
// The bucket lives in symmetric memory (every rank maps every peer's copy, plus one MULTICAST address that
// aliases all of them through the switch).  Rank r owns the r-th slice of the bucket:
//   multimem.ld_reduce  one 16-byte load from the multicast address returns the switch-side SUM of the W copies
//   multimem.st         one 16-byte store to the multicast address writes all W copies
// so every element crosses each GPU's links once in and once out (a ring all-reduce moves 2(W-1)/W of the
// bucket each way), and all ranks receive bit-identical sums (one reducer per element).  Without multicast
// (no NVSwitch) the same slices are reduced with peer loads in rank order and written with peer stores.
// The caller brackets the kernel with the symmetric-memory barrier (all gradients written before, all sums
// landed after).
#include "ws_common.cuh"

namespace ws {

__device__ __forceinline__ float4 mc_ld_reduce(const float4 *mc)
{
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st(float4 *mc, float4 v)
{
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

constexpr int AR_UNROLL = 4;
__global__ void __launch_bounds__(512) allreduce_multimem(float4 *mc, size_t begin4, size_t end4)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = begin4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (AR_UNROLL - 1) * stride < end4; i += AR_UNROLL * stride) {
        float4 v[AR_UNROLL];
#pragma unroll
        for (int k = 0; k < AR_UNROLL; k++) v[k] = mc_ld_reduce(mc + i + k * stride);
#pragma unroll
        for (int k = 0; k < AR_UNROLL; k++) mc_st(mc + i + k * stride, v[k]);
    }
    for (; i < end4; i += stride) mc_st(mc + i, mc_ld_reduce(mc + i));
}

constexpr int AR_MAX_WORLD = 16;
struct PeerPtrs { float4 *p[AR_MAX_WORLD]; };
__global__ void __launch_bounds__(512) allreduce_peers(PeerPtrs peers, int world, size_t begin4, size_t end4)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = begin4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end4; i += stride) {
        float4 s = peers.p[0][i];
        for (int k = 1; k < world; k++) {                 // rank order: the same sum on every rank
            const float4 v = peers.p[k][i];
            s.x = __fadd_rn(s.x, v.x); s.y = __fadd_rn(s.y, v.y); s.z = __fadd_rn(s.z, v.z); s.w = __fadd_rn(s.w, v.w);
        }
        for (int k = 0; k < world; k++) peers.p[k][i] = s;
    }
}

} // namespace ws

using namespace ws;

extern "C" int ws_allreduce_sum_f32(int32_t rank, int32_t world, int64_t numel, void *multicast_ptr,
                                    void *const *peer_ptrs, int32_t num_ctas, ws_stream_t stream)
{
    if (world < 1 || world > AR_MAX_WORLD || rank < 0 || rank >= world || numel < 0 || (!multicast_ptr && !peer_ptrs)) {
        set_error("ws_allreduce_sum_f32: bad arguments (world 1..%d)", AR_MAX_WORLD);
        return WS_E_INVALID;
    }
    if (numel % 4 != 0) { set_error("ws_allreduce_sum_f32: numel must be a multiple of 4 (16-byte vectors)"); return WS_E_INVALID; }
    if (world == 1 || numel == 0) return WS_OK;
    const size_t n4 = (size_t)numel / 4;
    const size_t per = (n4 + world - 1) / world;
    const size_t begin4 = per * rank < n4 ? per * rank : n4, end4 = per * (rank + 1) < n4 ? per * (rank + 1) : n4;
    if (begin4 == end4) return WS_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int ctas = num_ctas > 0 ? num_ctas : 148;
    if (multicast_ptr) {
        if (((uintptr_t)multicast_ptr & 15) != 0) { set_error("ws_allreduce_sum_f32: multicast pointer not 16-byte aligned"); return WS_E_INVALID; }
        allreduce_multimem<<<ctas, 512, 0, st>>>((float4 *)multicast_ptr, begin4, end4);
    } else {
        PeerPtrs pp;
        for (int k = 0; k < world; k++) {
            if (!peer_ptrs[k] || ((uintptr_t)peer_ptrs[k] & 15) != 0) { set_error("ws_allreduce_sum_f32: bad peer pointer %d", k); return WS_E_INVALID; }
            pp.p[k] = (float4 *)peer_ptrs[k];
        }
        allreduce_peers<<<ctas, 512, 0, st>>>(pp, world, begin4, end4);
    }
    WS_LAUNCH_CHECK(st, false);
    return WS_OK;
}
