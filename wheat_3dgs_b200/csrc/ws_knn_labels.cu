// ws_knn_labels.cu
//   ws_knn_dist2           simple-knn distCUDA2 (KNN/simple_knn.cu:186-222): bbox(points U origin),
//                          30-bit Morton codes, stable sort, 1024-point boxes, pruned 3-NN search.
//                          The reference copies the bbox to the host twice (:197,:200); here it stays
//                          on the device, so the call is fully asynchronous.
//   ws_multi_instance_opt  run_3d_seg.py:54-72 fused into one pass over the vote matrix (the
//                          reference loops over labels in Python, a handful of torch kernels each).
#include <float.h>
#include <mutex>
#include "ws_common.cuh"

namespace ws {

struct MinMax { float minn[3], maxx[3]; };

__global__ void __launch_bounds__(256) knn_bbox_partial(int P, const float *__restrict__ pts, float *__restrict__ part)
{
    float mn[3] = {0.f, 0.f, 0.f}, mx[3] = {0.f, 0.f, 0.f}; // init = {0,0,0}, simple_knn.cu:192
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x)
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float v = pts[3 * (size_t)i + k];
            mn[k] = fminf(mn[k], v);
            mx[k] = fmaxf(mx[k], v);
        }
    __shared__ float s[8][6];
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 3; k++) { s[threadIdx.x >> 5][k] = mn[k]; s[threadIdx.x >> 5][3 + k] = mx[k]; }
    __syncthreads();
    if (threadIdx.x < 6) {
        float r = s[0][threadIdx.x];
        for (int w = 1; w < 8; w++) r = threadIdx.x < 3 ? fminf(r, s[w][threadIdx.x]) : fmaxf(r, s[w][threadIdx.x]);
        part[blockIdx.x * 6 + threadIdx.x] = r;
    }
}
__global__ void knn_bbox_final(int nblk, const float *__restrict__ part, float *__restrict__ mm)
{
    const int k = threadIdx.x;
    if (k >= 6) return;
    float r = 0.f;
    for (int b = 0; b < nblk; b++) r = k < 3 ? fminf(r, part[b * 6 + k]) : fmaxf(r, part[b * 6 + k]);
    mm[k] = r;
}

__device__ __forceinline__ uint32_t prep_morton(uint32_t x) // simple_knn.cu:37-44
{
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}

__global__ void knn_morton(int P, const float *__restrict__ pts, const float *__restrict__ mm,
                           uint64_t *__restrict__ keys, uint32_t *__restrict__ vals)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    uint32_t c[3];
#pragma unroll
    for (int k = 0; k < 3; k++) { // simple_knn.cu:46-62
        const float t = __fmul_rn(__fdiv_rn(__fadd_rn(pts[3 * (size_t)i + k], -mm[k]), __fadd_rn(mm[3 + k], -mm[k])), 1023.0f);
        c[k] = prep_morton(__float2uint_rz(t));
    }
    keys[i] = (uint64_t)(c[0] | (c[1] << 1) | (c[2] << 2));
    vals[i] = (uint32_t)i;
}

// simple_knn.cu:79-118, one CTA of 1024 per box
__global__ void __launch_bounds__(1024) knn_box_minmax(int P, const float *__restrict__ pts,
                                                       const uint32_t *__restrict__ idx, MinMax *__restrict__ boxes)
{
    const int i = blockIdx.x * 1024 + threadIdx.x;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        const uint32_t g = idx[i];
#pragma unroll
        for (int k = 0; k < 3; k++) mn[k] = mx[k] = pts[3 * (size_t)g + k];
    }
    __shared__ float s[32][6];
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 3; k++) { s[threadIdx.x >> 5][k] = mn[k]; s[threadIdx.x >> 5][3 + k] = mx[k]; }
    __syncthreads();
    if (threadIdx.x < 6) {
        float r = s[0][threadIdx.x];
        for (int w = 1; w < 32; w++) r = threadIdx.x < 3 ? fminf(r, s[w][threadIdx.x]) : fmaxf(r, s[w][threadIdx.x]);
        if (threadIdx.x < 3) boxes[blockIdx.x].minn[threadIdx.x] = r;
        else boxes[blockIdx.x].maxx[threadIdx.x - 3] = r;
    }
}

__device__ __forceinline__ void update3(const float px, const float py, const float pz, const float *q, float *knn)
{
    const float dx = __fadd_rn(q[0], -px), dy = __fadd_rn(q[1], -py), dz = __fadd_rn(q[2], -pz);
    float dist = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy))); // contraction as in the reference SASS
#pragma unroll
    for (int j = 0; j < 3; j++)
        if (knn[j] > dist) { const float t = knn[j]; knn[j] = dist; dist = t; }
}

// simple_knn.cu:148-184
__global__ void __launch_bounds__(256) knn_box_mean_dist(int P, const float *__restrict__ pts,
                                                         const uint32_t *__restrict__ idx,
                                                         const MinMax *__restrict__ boxes, float *__restrict__ dists)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t me = idx[i];
    const float px = pts[3 * (size_t)me], py = pts[3 * (size_t)me + 1], pz = pts[3 * (size_t)me + 2];
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    for (int j = max(0, i - 3); j <= min(P - 1, i + 3); j++) {
        if (j == i) continue;
        update3(px, py, pz, pts + 3 * (size_t)idx[j], best);
    }
    const float reject = best[2];
    best[0] = best[1] = best[2] = FLT_MAX;
    const int nb = (P + 1023) / 1024;
    for (int b = 0; b < nb; b++) {
        const MinMax box = boxes[b];
        float d[3] = {0.f, 0.f, 0.f};
        const float pp[3] = {px, py, pz};
#pragma unroll
        for (int k = 0; k < 3; k++)
            if (pp[k] < box.minn[k] || pp[k] > box.maxx[k])
                d[k] = fminf(fabsf(__fadd_rn(pp[k], -box.minn[k])), fabsf(__fadd_rn(pp[k], -box.maxx[k])));
        const float dist = __fmaf_rn(d[2], d[2], __fmaf_rn(d[0], d[0], __fmul_rn(d[1], d[1])));
        if (dist > reject || dist > best[2]) continue;
        const int end = min(P, (b + 1) * 1024);
        for (int j = b * 1024; j < end; j++) {
            if (j == i) continue;
            update3(px, py, pz, pts + 3 * (size_t)idx[j], best);
        }
    }
    dists[me] = __fdiv_rn(__fadd_rn(best[2], __fadd_rn(best[0], best[1])), 3.0f);
}

struct KnnScratch {
    float *part, *mm;
    uint64_t *keys_a, *keys_b;
    uint32_t *vals_a, *vals_b;
    MinMax *boxes;
    SortScratch sort;
    static KnnScratch from(char *p, size_t P, size_t *bytes)
    {
        KnnScratch s;
        char *p0 = p;
        s.part = carve<float>(p, 6 * 1024);
        s.mm = carve<float>(p, 8);
        s.keys_a = carve<uint64_t>(p, P + 64);
        s.keys_b = carve<uint64_t>(p, P + 64);
        s.vals_a = carve<uint32_t>(p, P + 64);
        s.vals_b = carve<uint32_t>(p, P + 64);
        s.boxes = carve<MinMax>(p, (P + 1023) / 1024 + 1);
        s.sort = SortScratch::from(p, P);
        if (bytes) *bytes = (size_t)(p - p0) + 256;
        return s;
    }
};

// multi_instance_opt: thread per Gaussian (column).  labels[i][j] = argmax over
// normalize([sum_j - A_ij, A_ij]) with gamma on row 0; ties -> 0.  The column sum keeps torch's row order
// (sequential adds), but the loads are issued MI_UNROLL at a time so every thread has that many requests in
// flight (a dependent load-add chain over ~500 rows is latency-bound: 1 TB/s; batched: HBM-bound).
// Algorithmic bytes: 2 x 4 x rows x P read (the matrix does not fit L2 at 500 x 1 M) + rows x P written.
constexpr int MI_UNROLL = 16;
__device__ __forceinline__ uint8_t mi_label(float a, float sum, float gamma)
{
    const float rest = __fadd_rn(sum, -a);
    if (gamma == 0.0f) {
        // Both quotients share the positive divisor and IEEE division is monotonic in the numerator, so
        // a <= rest gives n1 <= n0 (label 0), and a clear margin (1e-6 relative, the quotient a/nrm is in
        // (0.7, 1]) gives n1 > n0 after rounding: the sqrt and the two divisions are only needed in the band
        // between (they made the kernel instruction-bound: 2.5 ms instead of ~1 ms at 501 x 1 M).
        if (!(a > rest)) return 0;
        if (a > __fmul_rn(rest, 1.000001f) && a > 1e-15f && a < 1e15f) return 1;
    }
    float nrm = sqrtf(__fadd_rn(__fmul_rn(rest, rest), __fmul_rn(a, a)));
    nrm = fmaxf(nrm, 1e-12f); // F.normalize eps
    const float n0 = __fadd_rn(__fdiv_rn(rest, nrm), gamma), n1 = __fdiv_rn(a, nrm);
    return n1 > n0 ? 1 : 0;
}
__global__ void __launch_bounds__(256) multi_instance_opt_kernel(int rows, int P, const float *__restrict__ A,
                                                                 float gamma, uint8_t *__restrict__ labels)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= P) return;
    const float *col = A + j;
    float sum = 0.f;
    int i = 0;
    for (; i + MI_UNROLL <= rows; i += MI_UNROLL) {
        float v[MI_UNROLL];
#pragma unroll
        for (int k = 0; k < MI_UNROLL; k++) v[k] = __ldcs(col + (size_t)(i + k) * P);
#pragma unroll
        for (int k = 0; k < MI_UNROLL; k++) sum = __fadd_rn(sum, v[k]);
    }
    for (; i < rows; i++) sum = __fadd_rn(sum, col[(size_t)i * P]);
    uint8_t *out = labels + j;
    i = 0;
    for (; i + MI_UNROLL <= rows; i += MI_UNROLL) {
        float v[MI_UNROLL];
#pragma unroll
        for (int k = 0; k < MI_UNROLL; k++) v[k] = __ldcs(col + (size_t)(i + k) * P);
#pragma unroll
        for (int k = 0; k < MI_UNROLL; k++) out[(size_t)(i + k) * P] = mi_label(v[k], sum, gamma);
    }
    for (; i < rows; i++) out[(size_t)i * P] = mi_label(col[(size_t)i * P], sum, gamma);
}

// four columns per thread (P % 4 == 0, 16-byte aligned): 16-byte loads, 4-byte label stores
constexpr int MI4_UNROLL = 8;
__global__ void __launch_bounds__(256) multi_instance_opt_kernel4(int rows, int P4, const float4 *__restrict__ A,
                                                                  float gamma, uchar4 *__restrict__ labels)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= P4) return;
    const float4 *col = A + j;
    float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
    int i = 0;
    for (; i + MI4_UNROLL <= rows; i += MI4_UNROLL) {
        float4 v[MI4_UNROLL];
#pragma unroll
        for (int k = 0; k < MI4_UNROLL; k++) v[k] = __ldcs(col + (size_t)(i + k) * P4);
#pragma unroll
        for (int k = 0; k < MI4_UNROLL; k++) {
            sum.x = __fadd_rn(sum.x, v[k].x); sum.y = __fadd_rn(sum.y, v[k].y);
            sum.z = __fadd_rn(sum.z, v[k].z); sum.w = __fadd_rn(sum.w, v[k].w);
        }
    }
    for (; i < rows; i++) {
        const float4 v = col[(size_t)i * P4];
        sum.x = __fadd_rn(sum.x, v.x); sum.y = __fadd_rn(sum.y, v.y); sum.z = __fadd_rn(sum.z, v.z); sum.w = __fadd_rn(sum.w, v.w);
    }
    uchar4 *out = labels + j;
    i = 0;
    for (; i + MI4_UNROLL <= rows; i += MI4_UNROLL) {
        float4 v[MI4_UNROLL];
#pragma unroll
        for (int k = 0; k < MI4_UNROLL; k++) v[k] = __ldcs(col + (size_t)(i + k) * P4);
#pragma unroll
        for (int k = 0; k < MI4_UNROLL; k++)
            out[(size_t)(i + k) * P4] = make_uchar4(mi_label(v[k].x, sum.x, gamma), mi_label(v[k].y, sum.y, gamma),
                                                    mi_label(v[k].z, sum.z, gamma), mi_label(v[k].w, sum.w, gamma));
    }
    for (; i < rows; i++) {
        const float4 v = col[(size_t)i * P4];
        out[(size_t)i * P4] = make_uchar4(mi_label(v.x, sum.x, gamma), mi_label(v.y, sum.y, gamma),
                                          mi_label(v.z, sum.z, gamma), mi_label(v.w, sum.w, gamma));
    }
}

} // namespace ws

using namespace ws;

extern "C" size_t ws_knn_bytes(int32_t P)
{
    size_t b = 0;
    KnnScratch::from(nullptr, (size_t)(P > 0 ? P : 0), &b);
    return b + 256;
}

extern "C" int ws_knn_dist2(int32_t P, const float *points, float *mean_dists, void *scratch, size_t scratch_bytes,
                            ws_stream_t stream)
{
    if (P < 0 || (P > 0 && (!points || !mean_dists || !scratch))) { set_error("ws_knn_dist2: bad arguments"); return WS_E_INVALID; }
    if (P == 0) return WS_OK;
    if (scratch_bytes < ws_knn_bytes(P)) { set_error("ws_knn_dist2: scratch too small"); return WS_E_SCRATCH; }
    cudaStream_t st = (cudaStream_t)stream;
    KnnScratch ks = KnnScratch::from((char *)scratch, (size_t)P, nullptr);
    int nblk = (P + 256 * 8 - 1) / (256 * 8);
    if (nblk > 1024) nblk = 1024;
    knn_bbox_partial<<<nblk, 256, 0, st>>>(P, points, ks.part);
    WS_LAUNCH_CHECK(st, false);
    knn_bbox_final<<<1, 32, 0, st>>>(nblk, ks.part, ks.mm);
    WS_LAUNCH_CHECK(st, false);
    knn_morton<<<(P + 255) / 256, 256, 0, st>>>(P, points, ks.mm, ks.keys_a, ks.vals_a);
    WS_LAUNCH_CHECK(st, false);
    uint64_t *kf;
    uint32_t *vf;
    int rc = launch_sort_pairs((size_t)P, ks.keys_a, ks.keys_b, ks.vals_a, ks.vals_b, 32, ks.sort, st, false, &kf, &vf);
    if (rc != WS_OK) return rc;
    const int nb = (P + 1023) / 1024;
    knn_box_minmax<<<nb, 1024, 0, st>>>(P, points, vf, ks.boxes);
    WS_LAUNCH_CHECK(st, false);
    knn_box_mean_dist<<<(P + 255) / 256, 256, 0, st>>>(P, points, vf, ks.boxes, mean_dists);
    WS_LAUNCH_CHECK(st, false);
    return WS_OK;
}

// ---- ws_knn_dist2_static: distCUDA2 for callers that have no way to obtain device scratch -------------------------
// The scratch of the search (Morton keys and point indices in two copies, the per-1024-point boxes, the radix sort's
// histograms and look-back words) lives in the array below instead of in caller memory.  The array is part of the
// library image: every GPU that loads libwheatsplat.so sets these 8 MiB aside until the process exits, used or not.
// In host code `g_knn_static` is only a handle for the runtime (its host-side value is not a device address and must
// never reach a kernel); the device address is asked for once per device with cudaGetSymbolAddress.
constexpr size_t KNN_STATIC_BYTES = (size_t)8 << 20;
static __device__ __align__(256) unsigned char g_knn_static[KNN_STATIC_BYTES];
constexpr int KNN_STATIC_MAX_DEVICES = 64;
static void *g_knn_static_ptr[KNN_STATIC_MAX_DEVICES] = {};
static std::mutex g_knn_static_mu;

// device address of g_knn_static on the CURRENT device (looked up on the first call per device, remembered after)
static int knn_static_base(char **base)
{
    int dev = -1;
    WS_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= KNN_STATIC_MAX_DEVICES) {
        set_error("ws_knn_dist2_static: device ordinal %d not supported (at most %d devices)", dev, KNN_STATIC_MAX_DEVICES);
        return WS_E_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(g_knn_static_mu);
    if (!g_knn_static_ptr[dev]) {
        void *ptr = nullptr;
        WS_CUDA_CHECK(cudaGetSymbolAddress((void **)&ptr, g_knn_static));
        if (!ptr) { set_error("ws_knn_dist2_static: cudaGetSymbolAddress returned a null address"); return WS_E_CUDA; }
        g_knn_static_ptr[dev] = ptr;
    }
    *base = (char *)g_knn_static_ptr[dev];
    return WS_OK;
}

extern "C" int ws_knn_dist2_static(int32_t P, const float *points, float *mean_dists, ws_stream_t stream)
{
    if (P < 0 || (P > 0 && (!points || !mean_dists))) { set_error("ws_knn_dist2_static: bad arguments"); return WS_E_INVALID; }
    if (P == 0) return WS_OK;
    if (P > WS_KNN_STATIC_MAX_POINTS || ws_knn_bytes(P) > KNN_STATIC_BYTES) {
        set_error("ws_knn_dist2_static: %d points exceed the built-in scratch (limit %d points, %zu bytes); "
                  "call ws_knn_dist2 with ws_knn_bytes(P) bytes of your own scratch instead",
                  P, WS_KNN_STATIC_MAX_POINTS, KNN_STATIC_BYTES);
        return WS_E_SCRATCH;
    }
    char *base = nullptr;
    int rc = knn_static_base(&base);
    if (rc != WS_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // Same carve and same launches as ws_knn_dist2.  What the previous call left in the array does not matter: the
    // bbox partials, keys, indices and boxes are all written before they are read, and the only words that must start
    // at zero (the sort's histograms, tile tickets and look-back status) are cleared on `st` by the sort itself.
    KnnScratch ks = KnnScratch::from(base, (size_t)P, nullptr);
    int nblk = (P + 256 * 8 - 1) / (256 * 8);
    if (nblk > 1024) nblk = 1024;
    knn_bbox_partial<<<nblk, 256, 0, st>>>(P, points, ks.part);
    WS_LAUNCH_CHECK(st, false);
    knn_bbox_final<<<1, 32, 0, st>>>(nblk, ks.part, ks.mm);
    WS_LAUNCH_CHECK(st, false);
    knn_morton<<<(P + 255) / 256, 256, 0, st>>>(P, points, ks.mm, ks.keys_a, ks.vals_a);
    WS_LAUNCH_CHECK(st, false);
    uint64_t *kf;
    uint32_t *vf;
    rc = launch_sort_pairs((size_t)P, ks.keys_a, ks.keys_b, ks.vals_a, ks.vals_b, 32, ks.sort, st, false, &kf, &vf);
    if (rc != WS_OK) return rc;
    knn_box_minmax<<<(P + 1023) / 1024, 1024, 0, st>>>(P, points, vf, ks.boxes);
    WS_LAUNCH_CHECK(st, false);
    knn_box_mean_dist<<<(P + 255) / 256, 256, 0, st>>>(P, points, vf, ks.boxes, mean_dists);
    WS_LAUNCH_CHECK(st, false);
    return WS_OK;
}

extern "C" int ws_multi_instance_opt(int32_t num_rows, int32_t P, const float *all_contrib, float gamma,
                                     uint8_t *labels, ws_stream_t stream)
{
    if (num_rows < 0 || P < 0 || ((size_t)num_rows * P > 0 && (!all_contrib || !labels))) {
        set_error("ws_multi_instance_opt: bad arguments");
        return WS_E_INVALID;
    }
    if ((size_t)num_rows * P == 0) return WS_OK;
    if (P % 4 == 0 && ((uintptr_t)all_contrib & 15) == 0 && ((uintptr_t)labels & 3) == 0)
        multi_instance_opt_kernel4<<<(P / 4 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
            num_rows, P / 4, (const float4 *)all_contrib, gamma, (uchar4 *)labels);
    else
        multi_instance_opt_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(num_rows, P, all_contrib, gamma, labels);
    WS_LAUNCH_CHECK((cudaStream_t)stream, false);
    return WS_OK;
}
