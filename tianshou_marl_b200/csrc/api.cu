// api.cu -- ABI bookkeeping entry points.
#include "common.cuh"

namespace tsm {
thread_local char g_err[512] = "";
thread_local bool g_pdl_hint = false;
}

extern "C" int tsm_abi_version(void) { return TSM_ABI_VERSION; }

extern "C" const char* tsm_last_error(void) { return tsm::g_err; }

extern "C" int tsm_device_info(int device, int* h_sm_count, int* h_cc_major, int* h_cc_minor) {
    int v = 0;
    if (h_sm_count) {
        TSM_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device));
        *h_sm_count = v;
    }
    if (h_cc_major) {
        TSM_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, device));
        *h_cc_major = v;
    }
    if (h_cc_minor) {
        TSM_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, device));
        *h_cc_minor = v;
    }
    return TSM_OK;
}
