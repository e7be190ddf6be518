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

// UNROLL independent 16-byte ld_reduce per thread are in flight before the first store: the switch round trip
// (several microseconds) times the link rate is the data that has to be outstanding (ws_allreduce_tune).
template <int UNROLL>
__global__ void __launch_bounds__(512) allreduce_multimem(float4 *mc, size_t begin4, size_t end4)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = begin4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (UNROLL - 1) * stride < end4; i += UNROLL * stride) {
        float4 v[UNROLL];
#pragma unroll
        for (int k = 0; k < UNROLL; k++) v[k] = mc_ld_reduce(mc + i + k * stride);
#pragma unroll
        for (int k = 0; k < UNROLL; k++) mc_st(mc + i + k * stride, v[k]);
    }
    for (; i < end4; i += stride) mc_st(mc + i, mc_ld_reduce(mc + i));
}
static int g_ar_unroll = 4, g_ar_threads = 512;
// share (per mille) of a rank's slice that goes through the switch when both paths are available; the rest is
// reduced by the SMs with peer loads and stores at the same time (ws_allreduce_split)
static int g_ar_mc_permille = 1000, g_ar_peer_ctas = 148;

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

// Both paths in one launch.  The in-switch reduction runs at about half the link rate whatever the issuing grid
// (0.57 ms for 248 MB at 8 GPUs, profiles/r02_exchange_sweep_8gpu.jsonl) and leaves link capacity idle; the peer
// load/store path is bound by the links, not by the switch's reduction units.  CTAs [0, mc_ctas) reduce the first part
// of the rank's slice through the switch, the other CTAs the rest with peer loads in rank order + peer stores.  Every
// element still has exactly one reducer, so all ranks hold identical bits.
__global__ void __launch_bounds__(512) allreduce_hybrid(float4 *mc, PeerPtrs peers, int world, size_t begin4, size_t mid4,
                                                        size_t end4, int mc_ctas)
{
    if ((int)blockIdx.x < mc_ctas) {
        const size_t stride = (size_t)mc_ctas * blockDim.x;
        size_t i = begin4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + 3 * stride < mid4; i += 4 * stride) {
            float4 v[4];
#pragma unroll
            for (int k = 0; k < 4; k++) v[k] = mc_ld_reduce(mc + i + k * stride);
#pragma unroll
            for (int k = 0; k < 4; k++) mc_st(mc + i + k * stride, v[k]);
        }
        for (; i < mid4; i += stride) mc_st(mc + i, mc_ld_reduce(mc + i));
    } else {
        const size_t stride = (size_t)(gridDim.x - mc_ctas) * blockDim.x;
        for (size_t i = mid4 + (size_t)(blockIdx.x - mc_ctas) * blockDim.x + threadIdx.x; i < end4; i += stride) {
            float4 s = peers.p[0][i];
            for (int k = 1; k < world; k++) {
                const float4 v = peers.p[k][i];
                s.x = __fadd_rn(s.x, v.x); s.y = __fadd_rn(s.y, v.y); s.z = __fadd_rn(s.z, v.z); s.w = __fadd_rn(s.w, v.w);
            }
            for (int k = 0; k < world; k++) peers.p[k][i] = s;
        }
    }
}

// ---- reduce-scatter + Adam + all-gather in ONE kernel (sharded optimiser step) ----
// Parameters are replicated, so after the exchange every rank would run the same Adam step on the same summed
// gradient (59 P elements, 28 B each).  Here rank r owns the r-th slice of the flat parameter space: it pulls the
// slice's gradient SUM through the switch (multimem.ld_reduce, or peer loads in rank order), applies Adam with its
// slice of the moments, and writes the new parameters into every replica (multimem.st, or peer stores).  Link
// traffic is that of the all-reduce (a slice in, the bucket out for the reduction; the bucket in, a slice out for the
// broadcast -- of 59 P parameters instead of 62 P gradients), but the summed gradient is never written anywhere, each
// rank touches 1/W of the optimiser state, and the separate 28 B/element Adam pass over all 59 P elements disappears.
// Element arithmetic = adam_elem, so parameters are bit-identical to exchange + ws_adam_step.
constexpr int SA_MAX_SEG = 8;
struct ShardSeg { long long begin4, goff4; float neg_step; };   // float4 units: first vector, gradient vector of it
struct ShardAdamParams {
    ShardSeg seg[SA_MAX_SEG];
    int n_seg, world, rank;
    float w1, beta2, w2, eps, bc2_sqrt;
    const float4 *grad_mc;
    float4 *param_mc;
    PeerPtrs grad_peers, param_peers;
    float4 *m, *v;
    size_t begin4, end4;
};

constexpr int SA_UNROLL = 2;   // independent vectors in flight per thread (the switch round trip is the latency to hide)
template <bool MC>
__global__ void __launch_bounds__(512) sharded_adam(const ShardAdamParams P)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const float4 *__restrict__ mine = P.param_peers.p[P.rank];
    for (size_t i0 = P.begin4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < P.end4; i0 += SA_UNROLL * stride) {
        float4 g[SA_UNROLL], p[SA_UNROLL], m[SA_UNROLL], v[SA_UNROLL];
        float ns[SA_UNROLL];
#pragma unroll
        for (int u = 0; u < SA_UNROLL; u++) {
            const size_t i = i0 + u * stride;
            if (i >= P.end4) break;
            int si = 0;
#pragma unroll
            for (int k = 1; k < SA_MAX_SEG; k++)
                if (k < P.n_seg && (long long)i >= P.seg[k].begin4) si = k;
            const size_t gi = (size_t)(P.seg[si].goff4 + ((long long)i - P.seg[si].begin4));
            ns[u] = P.seg[si].neg_step;
            if (MC) {
                g[u] = mc_ld_reduce(P.grad_mc + gi);
            } else {
                g[u] = P.grad_peers.p[0][gi];
                for (int k = 1; k < P.world; k++) {
                    const float4 x = P.grad_peers.p[k][gi];
                    g[u].x = __fadd_rn(g[u].x, x.x); g[u].y = __fadd_rn(g[u].y, x.y);
                    g[u].z = __fadd_rn(g[u].z, x.z); g[u].w = __fadd_rn(g[u].w, x.w);
                }
            }
            p[u] = mine[i]; m[u] = P.m[i]; v[u] = P.v[i];
        }
#pragma unroll
        for (int u = 0; u < SA_UNROLL; u++) {
            const size_t i = i0 + u * stride;
            if (i >= P.end4) break;
            adam_elem(p[u].x, g[u].x, m[u].x, v[u].x, P.w1, P.beta2, P.w2, P.eps, P.bc2_sqrt, ns[u]);
            adam_elem(p[u].y, g[u].y, m[u].y, v[u].y, P.w1, P.beta2, P.w2, P.eps, P.bc2_sqrt, ns[u]);
            adam_elem(p[u].z, g[u].z, m[u].z, v[u].z, P.w1, P.beta2, P.w2, P.eps, P.bc2_sqrt, ns[u]);
            adam_elem(p[u].w, g[u].w, m[u].w, v[u].w, P.w1, P.beta2, P.w2, P.eps, P.bc2_sqrt, ns[u]);
            P.m[i] = m[u];
            P.v[i] = v[u];
            if (MC) {
                mc_st(P.param_mc + i, p[u]);
            } else {
                for (int k = 0; k < P.world; k++) P.param_peers.p[k][i] = p[u];
            }
        }
    }
}

} // namespace ws

using namespace ws;

extern "C" int ws_allreduce_tune(int32_t unroll, int32_t threads)
{
    if ((unroll != 2 && unroll != 4 && unroll != 8) || threads < 32 || threads > 512 || threads % 32) {
        set_error("ws_allreduce_tune: unroll 2|4|8, threads a multiple of 32 up to 512");
        return WS_E_INVALID;
    }
    ws::g_ar_unroll = unroll; ws::g_ar_threads = threads;
    return WS_OK;
}

extern "C" int ws_allreduce_split(int32_t multicast_permille, int32_t peer_ctas)
{
    if (multicast_permille < 0 || multicast_permille > 1000 || peer_ctas < 1 || peer_ctas > 4096) {
        set_error("ws_allreduce_split: share 0..1000 per mille, 1..4096 peer CTAs");
        return WS_E_INVALID;
    }
    ws::g_ar_mc_permille = multicast_permille; ws::g_ar_peer_ctas = peer_ctas;
    return WS_OK;
}

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
    if (multicast_ptr && peer_ptrs && ws::g_ar_mc_permille < 1000) {
        if (((uintptr_t)multicast_ptr & 15) != 0) { set_error("ws_allreduce_sum_f32: multicast pointer not 16-byte aligned"); return WS_E_INVALID; }
        PeerPtrs pp;
        for (int k = 0; k < world; k++) {
            if (!peer_ptrs[k] || ((uintptr_t)peer_ptrs[k] & 15) != 0) { set_error("ws_allreduce_sum_f32: bad peer pointer %d", k); return WS_E_INVALID; }
            pp.p[k] = (float4 *)peer_ptrs[k];
        }
        const size_t mid4 = begin4 + (end4 - begin4) * (size_t)ws::g_ar_mc_permille / 1000;
        allreduce_hybrid<<<ctas + ws::g_ar_peer_ctas, 512, 0, st>>>((float4 *)multicast_ptr, pp, world, begin4, mid4, end4, ctas);
    } else if (multicast_ptr) {
        if (((uintptr_t)multicast_ptr & 15) != 0) { set_error("ws_allreduce_sum_f32: multicast pointer not 16-byte aligned"); return WS_E_INVALID; }
        const int th = ws::g_ar_threads;
        if (ws::g_ar_unroll >= 8) allreduce_multimem<8><<<ctas, th, 0, st>>>((float4 *)multicast_ptr, begin4, end4);
        else if (ws::g_ar_unroll <= 2) allreduce_multimem<2><<<ctas, th, 0, st>>>((float4 *)multicast_ptr, begin4, end4);
        else allreduce_multimem<4><<<ctas, th, 0, st>>>((float4 *)multicast_ptr, begin4, end4);
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

extern "C" int ws_sharded_adam_step(int32_t rank, int32_t world, int64_t numel, int32_t n_segments,
                                    const ws_shard_segment *segments, void *grad_multicast, void *const *grad_peers,
                                    void *param_multicast, void *const *param_peers, float *exp_avg, float *exp_avg_sq,
                                    int64_t step, double beta1, double beta2, double eps, int32_t num_ctas,
                                    ws_stream_t stream)
{
    if (world < 1 || world > AR_MAX_WORLD || rank < 0 || rank >= world || numel < 0 || !param_peers || !exp_avg || !exp_avg_sq ||
        n_segments < 1 || n_segments > SA_MAX_SEG || !segments || step < 1) {
        set_error("ws_sharded_adam_step: bad arguments (world 1..%d, 1..%d segments, step >= 1)", AR_MAX_WORLD, SA_MAX_SEG);
        return WS_E_INVALID;
    }
    const bool mc = grad_multicast != nullptr && param_multicast != nullptr;
    if (!mc && !grad_peers) { set_error("ws_sharded_adam_step: neither multicast nor peer gradient pointers"); return WS_E_INVALID; }
    if (numel % 4 != 0) { set_error("ws_sharded_adam_step: numel must be a multiple of 4"); return WS_E_INVALID; }
    if (numel == 0) return WS_OK;
    ShardAdamParams P;
    P.n_seg = n_segments; P.world = world; P.rank = rank;
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    long long expect = 0;
    for (int k = 0; k < n_segments; k++) {
        const ws_shard_segment &s = segments[k];
        if (s.begin != expect || s.end < s.begin || s.begin % 4 || s.end % 4 || s.grad_offset % 4 || s.grad_offset < 0) {
            set_error("ws_sharded_adam_step: segment %d: ranges must tile [0, numel) in order, in multiples of 4", k);
            return WS_E_INVALID;
        }
        expect = s.end;
        P.seg[k].begin4 = s.begin / 4; P.seg[k].goff4 = s.grad_offset / 4;
        P.seg[k].neg_step = (float)((s.lr / bc1) * -1.0);
    }
    if (expect != numel) { set_error("ws_sharded_adam_step: segments end at %lld, numel is %lld", expect, (long long)numel); return WS_E_INVALID; }
    P.w1 = (float)(1.0 - beta1); P.beta2 = (float)beta2; P.w2 = (float)(1.0 - beta2); P.eps = (float)eps;
    P.bc2_sqrt = (float)sqrt(bc2);
    P.grad_mc = (const float4 *)grad_multicast; P.param_mc = (float4 *)param_multicast;
    for (int k = 0; k < world; k++) {
        if (!param_peers[k] || ((uintptr_t)param_peers[k] & 15) || (!mc && (!grad_peers[k] || ((uintptr_t)grad_peers[k] & 15)))) {
            set_error("ws_sharded_adam_step: bad peer pointer %d", k);
            return WS_E_INVALID;
        }
        P.param_peers.p[k] = (float4 *)param_peers[k];
        P.grad_peers.p[k] = mc ? nullptr : (float4 *)grad_peers[k];
    }
    if ((((uintptr_t)exp_avg | (uintptr_t)exp_avg_sq | (uintptr_t)grad_multicast | (uintptr_t)param_multicast) & 15) != 0) {
        set_error("ws_sharded_adam_step: pointers must be 16-byte aligned");
        return WS_E_INVALID;
    }
    P.m = (float4 *)exp_avg; P.v = (float4 *)exp_avg_sq;
    const size_t n4 = (size_t)numel / 4, per = (n4 + world - 1) / world;
    P.begin4 = per * rank < n4 ? per * rank : n4;
    P.end4 = per * (rank + 1) < n4 ? per * (rank + 1) : n4;
    if (P.begin4 == P.end4) return WS_OK;
    cudaStream_t st = (cudaStream_t)stream;
    // multicast: switch-bound from 74 CTAs up (0.56 ms at 8 GPUs whatever the grid); peer loads want more in flight
    // (2 GPUs: 148 CTAs 0.43 ms, 592 CTAs 0.38 ms)
    const int ctas = num_ctas > 0 ? num_ctas : (mc ? 148 : 592);
    if (mc) sharded_adam<true><<<ctas, 512, 0, st>>>(P);
    else sharded_adam<false><<<ctas, 512, 0, st>>>(P);
    WS_LAUNCH_CHECK(st, false);
    return WS_OK;
}
