// Test infrastructure only: a C entry point around the reference's vendored libSGM (multiview/contrib/libSGM), built by
// oracle/Makefile (target _ref/libsgm_ref.so) from the sources where they lie.  It lets the GPU tests run the reference's own
// kernels for the CUDA_SGM strategy beside the CPU restatement (oracle/census_oracle.c) on the same inputs.
#include <cuda_runtime.h>
#include <libsgm.h>

extern "C" int sgm_ref_execute(const void* left, const void* right, void* dst, int w, int h, int disparity_size, int in_bits, int out_bits)
{
    try {
        sgm::StereoSGM sgm(w, h, disparity_size, in_bits, out_bits, sgm::EXECUTE_INOUT_HOST2HOST);
        void* out = dst;
        sgm.execute(left, right, &out);
        return cudaDeviceSynchronize() == cudaSuccess ? 0 : -2;
    } catch (...) {
        return -1;
    }
}
