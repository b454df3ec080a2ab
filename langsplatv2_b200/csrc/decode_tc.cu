// decode_tc.cu -- K7 (tensor-core variant) placeholder; replaced by the tcgen05 kernel.
#include "common.cuh"
namespace lsv2 {
size_t decode_tc_scratch_bytes(int levels, int Lc, int D) { return 256; }
int launch_decode_tc(int levels, int Lc, int D, int64_t HW, const float* wmap, const float* codebooks,
                     int normalize, int out_dtype, void* out, void* scratch, size_t scratch_bytes, cudaStream_t st)
{
    return LSV2_EUNSUPPORTED;
}
}  // namespace lsv2
