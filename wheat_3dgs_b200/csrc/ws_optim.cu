// ws_optim.cu -- the step either side of the rasterizer backward:
//
//   adam_groups      torch.optim.Adam(l, lr=0.0, eps=1e-15).step() over the six parameter groups of
//                    GaussianModel.training_setup (scene/gaussian_model.py:167-186: xyz, f_dc, f_rest,
//                    opacity, scaling, rotation; betas (0.9, 0.999), no weight decay, no amsgrad), called
//                    at train_vanilla_3dgs.py:110-112.  torch runs its multi-tensor path on CUDA: seven
//                    foreach kernels per step (lerp, mul, addcmul, sqrt, div, add, addcdiv), 72 bytes of
//                    HBM traffic per parameter element.  Here: ONE launch over all groups, every element
//                    read once (param, grad, exp_avg, exp_avg_sq) and written once (param, exp_avg,
//                    exp_avg_sq): 28 bytes per element, HBM-bound.  The per-element float operations are
//                    the ones torch's foreach functors perform, in their order, with the contractions
//                    nvcc applies to them (explicit fmaf), so results agree with torch to the last bit on
//                    CUDA and to ~1 ulp against torch's CPU kernels (the golden).
//   densify_stats    train_vanilla_3dgs.py:100-103 + GaussianModel.add_densification_stats
//                    (scene/gaussian_model.py:462-464): for every Gaussian with radii > 0
//                    max_radii2D = max(max_radii2D, radii), xyz_gradient_accum += |dL_dmeans2D.xy|,
//                    denom += 1.  The reference does this with 2 boolean-mask gathers, a norm, 3 masked
//                    scatters (index_put) and a nonzero() host sync per iteration; here one pass, no sync.
#include "ws_common.cuh"

namespace ws {

constexpr int AD_THREADS = 256;
constexpr int AD_CHUNK = AD_THREADS * 4;   // floats per CTA iteration
constexpr int AD_MAX_GROUPS = 8;

struct AdamGroupDev {
    float *param;
    const float *grad;
    float *exp_avg, *exp_avg_sq;
    long long numel;
    long long chunk_begin;     // first chunk of this group in the flat chunk space
    float neg_step_size;       // -(lr / (1 - beta1^step))
    float bc2_sqrt;            // sqrt(1 - beta2^step)
    int grad_vec;              // 1: grad is 16-byte aligned (float4 loads); 0: only 4-byte aligned (scalar loads)
};
struct AdamParams {
    AdamGroupDev g[AD_MAX_GROUPS];
    int n_groups;
    long long n_chunks;
    float w1;                  // 1 - beta1   (lerp weight)
    float beta2, w2;           // beta2, 1 - beta2
    float eps;
};

// adam_elem: ws_common.cuh (shared with the sharded step of ws_collective.cu)

__global__ void __launch_bounds__(AD_THREADS) adam_groups(const AdamParams P)
{
    for (long long c = blockIdx.x; c < P.n_chunks; c += gridDim.x) {
        int gi = 0;
#pragma unroll
        for (int k = 1; k < AD_MAX_GROUPS; k++)
            if (k < P.n_groups && c >= P.g[k].chunk_begin) gi = k;
        const AdamGroupDev &G = P.g[gi];
        const long long base = (c - G.chunk_begin) * AD_CHUNK + threadIdx.x * 4;
        if (base >= G.numel) continue;
        if (base + 4 <= G.numel) {
            float4 p = *reinterpret_cast<const float4 *>(G.param + base);
            // the gradients are slices of ONE flat bucket (ops._backward): at offsets 3P, 6P, ... they are 16-byte
            // aligned only when P is a multiple of 4, which it is not after most densifications
            const float4 g = G.grad_vec ? __ldcs(reinterpret_cast<const float4 *>(G.grad + base))
                                        : make_float4(__ldcs(G.grad + base), __ldcs(G.grad + base + 1),
                                                      __ldcs(G.grad + base + 2), __ldcs(G.grad + base + 3));
            float4 m = *reinterpret_cast<const float4 *>(G.exp_avg + base);
            float4 v = *reinterpret_cast<const float4 *>(G.exp_avg_sq + base);
            adam_elem(p.x, g.x, m.x, v.x, P.w1, P.beta2, P.w2, P.eps, G.bc2_sqrt, G.neg_step_size);
            adam_elem(p.y, g.y, m.y, v.y, P.w1, P.beta2, P.w2, P.eps, G.bc2_sqrt, G.neg_step_size);
            adam_elem(p.z, g.z, m.z, v.z, P.w1, P.beta2, P.w2, P.eps, G.bc2_sqrt, G.neg_step_size);
            adam_elem(p.w, g.w, m.w, v.w, P.w1, P.beta2, P.w2, P.eps, G.bc2_sqrt, G.neg_step_size);
            *reinterpret_cast<float4 *>(G.param + base) = p;
            *reinterpret_cast<float4 *>(G.exp_avg + base) = m;
            *reinterpret_cast<float4 *>(G.exp_avg_sq + base) = v;
        } else {
            for (long long i = base; i < G.numel; i++) {
                float p = G.param[i], m = G.exp_avg[i], v = G.exp_avg_sq[i];
                adam_elem(p, G.grad[i], m, v, P.w1, P.beta2, P.w2, P.eps, G.bc2_sqrt, G.neg_step_size);
                G.param[i] = p; G.exp_avg[i] = m; G.exp_avg_sq[i] = v;
            }
        }
    }
}

__global__ void __launch_bounds__(256) densify_stats(int P, const int *__restrict__ radii,
                                                     const float *__restrict__ dL_dmeans2D, float *max_radii2D,
                                                     float *xyz_gradient_accum, float *denom)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r <= 0) return;                                   // visibility_filter = radii > 0
    max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);     // torch.max(float, int32) promotes to float
    const float gx = dL_dmeans2D[3 * i], gy = dL_dmeans2D[3 * i + 1];
    // torch.norm(grad[:, :2], dim=-1): sqrt of the running sum of squares, contracted
    xyz_gradient_accum[i] = __fadd_rn(xyz_gradient_accum[i], __fsqrt_rn(__fmaf_rn(gy, gy, __fmul_rn(gx, gx))));
    denom[i] = __fadd_rn(denom[i], 1.0f);
}

} // namespace ws

using namespace ws;

extern "C" int ws_adam_step(int32_t n_groups, const ws_adam_group *groups, double beta1, double beta2, double eps,
                            ws_stream_t stream)
{
    if (n_groups < 0 || n_groups > AD_MAX_GROUPS || (n_groups && !groups)) {
        set_error("ws_adam_step: 0..%d groups", AD_MAX_GROUPS);
        return WS_E_INVALID;
    }
    AdamParams P;
    P.n_groups = 0;
    long long chunks = 0;
    for (int k = 0; k < n_groups; k++) {
        const ws_adam_group &g = groups[k];
        if (g.numel < 0 || g.step < 1) { set_error("ws_adam_step: group %d: numel < 0 or step < 1", k); return WS_E_INVALID; }
        if (g.numel == 0) continue;
        if (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq) {
            set_error("ws_adam_step: group %d has a NULL tensor", k);
            return WS_E_INVALID;
        }
        if ((((uintptr_t)g.param | (uintptr_t)g.exp_avg | (uintptr_t)g.exp_avg_sq) & 15) != 0 || ((uintptr_t)g.grad & 3) != 0) {
            set_error("ws_adam_step: group %d: param / exp_avg / exp_avg_sq must be 16-byte aligned, grad 4-byte aligned", k);
            return WS_E_INVALID;
        }
        AdamGroupDev &d = P.g[P.n_groups++];
        d.param = g.param; d.grad = g.grad; d.exp_avg = g.exp_avg; d.exp_avg_sq = g.exp_avg_sq;
        d.numel = g.numel; d.chunk_begin = chunks;
        d.grad_vec = ((uintptr_t)g.grad & 15) == 0 ? 1 : 0;
        // torch/optim/adam.py (_multi_tensor_adam): Python doubles, cast to float at the kernel boundary
        const double bc1 = 1.0 - pow(beta1, (double)g.step);
        const double bc2 = 1.0 - pow(beta2, (double)g.step);
        d.neg_step_size = (float)((g.lr / bc1) * -1.0);
        d.bc2_sqrt = (float)sqrt(bc2);
        chunks += (g.numel + AD_CHUNK - 1) / AD_CHUNK;
    }
    if (chunks == 0) return WS_OK;
    P.n_chunks = chunks;
    // Python doubles on the torch side, cast to float where the foreach kernels take their scalar
    P.w1 = (float)(1.0 - beta1);
    P.beta2 = (float)beta2;
    P.w2 = (float)(1.0 - beta2);
    P.eps = (float)eps;
    cudaStream_t st = (cudaStream_t)stream;
    const long long want = chunks < 148LL * 16 ? chunks : 148LL * 16;   // 16 resident CTAs of 256 threads per SM
    adam_groups<<<(unsigned)want, AD_THREADS, 0, st>>>(P);
    WS_LAUNCH_CHECK(st, false);
    return WS_OK;
}

extern "C" int ws_densify_stats(int32_t P, const int32_t *radii, const float *dL_dmeans2D, float *max_radii2D,
                                float *xyz_gradient_accum, float *denom, ws_stream_t stream)
{
    if (P < 0 || (P && (!radii || !dL_dmeans2D || !max_radii2D || !xyz_gradient_accum || !denom))) {
        set_error("ws_densify_stats: bad arguments");
        return WS_E_INVALID;
    }
    if (P == 0) return WS_OK;
    cudaStream_t st = (cudaStream_t)stream;
    densify_stats<<<(P + 255) / 256, 256, 0, st>>>(P, radii, dL_dmeans2D, max_radii2D, xyz_gradient_accum, denom);
    WS_LAUNCH_CHECK(st, false);
    return WS_OK;
}
