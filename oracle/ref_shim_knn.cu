/*
 * ref_shim_knn.cu -- C-ABI shim around the UNMODIFIED reference simple-knn
 * (SimpleKNN::knn, reference submodules/simple-knn/simple_knn.cu:186-222), compiled by
 * oracle/build_ref.py with the reference's simple_knn.cu into oracle/_ref/libref_knn.so.
 * TEST / BASELINE INFRASTRUCTURE ONLY.  Replaces spatial.cu (which needs libtorch).
 */
#include <cuda_runtime.h>
#include "simple_knn.h"

extern "C" void ref_knn_dist2(int P, float *points_dev, float *mean_dists_dev)
{
    SimpleKNN::knn(P, (float3 *)points_dev, mean_dists_dev);
}
