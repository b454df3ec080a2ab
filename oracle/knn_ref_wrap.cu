// TEST INFRASTRUCTURE (oracle): a C entry point in front of the reference's own SimpleKNN::knn
// (/root/reference/submodules/simple-knn/simple_knn.cu:185-220, compiled from where it lies by oracle/Makefile's
// `_ref` target into oracle/_ref/libsimple_knn_ref.so).  Only tests/, smoke() and bench.py's reference leg may load it.
#include <cuda_runtime.h>
#include "simple_knn.h"

extern "C" int ref_simple_knn(int P, const float* points_dev, float* mean_dist2_dev)
{
    SimpleKNN::knn(P, (float3*)points_dev, mean_dist2_dev);
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaGetLastError();
    return (int)e;
}
