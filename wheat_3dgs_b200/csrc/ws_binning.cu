// ws_binning.cu -- identify_ranges: one pass over the sorted tile keys that writes the per-tile [start,end)
// ranges (replaces identifyTileRanges, DGR rasterizer_impl.cu:116-138, after the cudaMemset of :311), plus the
// small helpers that say where the sorted list lives.
//
// Round 1 gathered every instance's 48-byte Gaussian record into tile order here (48 R bytes written, read again
// by both blend kernels).  The blend kernels now gather the rows of the instances that survive the quadrant
// cull themselves (TMA gather4 from the L2-resident per-Gaussian table), so this pass only reads the keys.
// Algorithmic HBM bytes: 4 R read (8 R with 64-bit keys) + 8 T written.
#include <cuda.h>
#include "ws_common.cuh"

namespace ws {

__device__ __forceinline__ uint32_t tile_of(uint64_t k) { return (uint32_t)(k >> 32); }
__device__ __forceinline__ uint32_t tile_of(uint32_t k) { return k; }

template <typename K>
__global__ void __launch_bounds__(256) identify_ranges(size_t R, const K *__restrict__ keys, uint2 *__restrict__ ranges)
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
}

// 32-bit tile keys, four per thread (one 16-byte load + the key in front of them)
__global__ void __launch_bounds__(256) identify_ranges4(size_t R, const uint32_t *__restrict__ keys, uint2 *__restrict__ ranges)
{
    const size_t base = 4 * ((size_t)blockIdx.x * blockDim.x + threadIdx.x);
    pdl_wait();
    pdl_trigger();
    if (base >= R) return;
    uint32_t k[4];
    if (base + 3 < R) {
        const uint4 v = *reinterpret_cast<const uint4 *>(keys + base);
        k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; j++) k[j] = base + j < R ? keys[base + j] : 0u;
    }
    uint32_t prev = base ? keys[base - 1] : 0u;
    if (base == 0) ranges[k[0]].x = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const size_t idx = base + j;
        if (idx < R) {
            if (idx != 0 && k[j] != prev) {
                ranges[prev].y = (uint32_t)idx;
                ranges[k[j]].x = (uint32_t)idx;
            }
            if (idx == R - 1) ranges[k[j]].y = (uint32_t)R;
            prev = k[j];
        }
    }
}

int launch_identify_ranges(size_t R, int T, const uint64_t *keys, const uint32_t *tkeys, uint2 *ranges,
                           cudaStream_t st, bool debug, bool ranges_cleared)
{
    if (!ranges_cleared) WS_CUDA_CHECK(cudaMemsetAsync(ranges, 0, (size_t)T * sizeof(uint2), st));
    if (R == 0) return WS_OK;
    if (tkeys && ((uintptr_t)tkeys & 15) == 0)
        WS_CUDA_CHECK(launch_pdl(identify_ranges4, dim3((unsigned)((R + 1023) / 1024)), dim3(256), 0, st, R, tkeys, ranges));
    else if (tkeys)
        identify_ranges<uint32_t><<<(unsigned)((R + 255) / 256), 256, 0, st>>>(R, tkeys, ranges);
    else
        identify_ranges<uint64_t><<<(unsigned)((R + 255) / 256), 256, 0, st>>>(R, keys, ranges);
    WS_LAUNCH_CHECK(st, debug);
    return WS_OK;
}

// TMA descriptor of the per-Gaussian table: 2-D float tensor, 16 columns x P rows, 64-byte row pitch, box 16 x 1,
// no swizzle.  tile::gather4 then brings four rows given by four row coordinates.  The driver entry point is looked
// up through the runtime (no link-time dependency on libcuda).
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_grec_tensor_map(TensorMap *tm, const GRec *grec, int P)
{
    static_assert(sizeof(TensorMap) == sizeof(CUtensorMap), "TensorMap must have the size of CUtensorMap");
    static EncodeTiledFn encode = nullptr;
    if (!encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
        if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
            set_error("cuTensorMapEncodeTiled is not available from this driver (%s)", cudaGetErrorString(e));
            return WS_E_CUDA;
        }
        encode = (EncodeTiledFn)fn;
    }
    // the descriptor depends on the table's address and row count only: the last one is kept per thread (the forward
    // and the backward of a view, and every view of a trainer whose allocator hands back the same block, reuse it)
    static thread_local struct { const GRec *grec; int P; TensorMap tm; bool ok; } last = {nullptr, 0, {}, false};
    if (last.ok && last.grec == grec && last.P == P) {
        *tm = last.tm;
        return WS_OK;
    }
    const cuuint64_t dims[2] = {16, (cuuint64_t)(P > 0 ? P : 1)};
    const cuuint64_t strides[1] = {sizeof(GRec)};
    const cuuint32_t box[2] = {16, 1};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = encode((CUtensorMap *)tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)grec, dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d) for the %d x 64-byte Gaussian table at %p", (int)r, P, (const void *)grec);
        return WS_E_CUDA;
    }
    last.grec = grec; last.P = P; last.tm = *tm; last.ok = true;
    return WS_OK;
}

} // namespace ws
