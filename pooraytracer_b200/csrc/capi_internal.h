/*
 * capi_internal.h — what the parts of the C ABI share among themselves (capi.cu, multi.cu). Not installed.
 */
#ifndef PRT_CAPI_INTERNAL_H
#define PRT_CAPI_INTERNAL_H

#include <cuda_runtime.h>

#include <string>

struct prt_context {
    int device = -1; /* < 0: host-only */
    cudaStream_t stream = nullptr;
    int sm_count = 148;
};

namespace prt {
/* records the message for prt_last_error() on the calling thread and returns `code` */
int set_error(int code, const std::string& msg);
}  // namespace prt

#endif
