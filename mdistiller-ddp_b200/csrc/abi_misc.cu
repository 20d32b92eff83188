// Version / error-text entry points of the C-ABI (include/mdistiller_b200.h).
#include "common.cuh"

extern "C" int mdb200_abi_version(void) { return MDB200_ABI_VERSION; }

extern "C" const char *mdb200_error_string(int code) {
    switch (code) {
        case 0: return "ok";
        case MDB200_EINVAL: return "invalid argument (null pointer, non-positive size or unknown mode)";
        case MDB200_ESHAPE: return "leading dimension smaller than the row length";
        case MDB200_EWORKSPC: return "workspace missing or too small (see mdb200_loss_workspace_bytes)";
        case MDB200_EALIGN: return "pointer alignment too small";
        case MDB200_ESCRATCH:
            return "problem too large for the built-in scratch: the launch needs more than "
                   "MDB200_STATIC_SCRATCH_CTAS thread blocks; call mdb200_distill_loss with a workspace instead";
        default: break;
    }
    if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
    return "unknown mdistiller_b200 error code";
}
