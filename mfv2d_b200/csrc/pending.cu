// Entry points declared in include/mfv2d_cuda.h whose kernels are not written yet.  They fail loudly
// (MFV2D_CUDA_UNSUPPORTED with a message) - there is no CPU fallback behind any of them.
#include "internal.cuh"

#define PENDING(ctx, name) return mfv2d_fail((ctx), MFV2D_CUDA_UNSUPPORTED, name " is not implemented yet")

extern "C" {

int mfv2d_cuda_projector_batch(const mfv2d_elements *elems_out, const mfv2d_basis *, int, int, double *, int32_t *, int)
{
    PENDING(elems_out->ctx, "mfv2d_cuda_projector_batch");
}

} // extern "C"
