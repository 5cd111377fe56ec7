// Context object and error plumbing of the C-ABI (include/moca_b200.h).
#include <stdlib.h>
#include "common.cuh"

namespace moca {
static thread_local char g_err[1024] = "";
int pdl_level() {
    static const int lvl = [] { const char* e = getenv("MOCA_PDL"); return e ? atoi(e) : 0; }();
    return lvl;
}

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace moca

using namespace moca;

extern "C" const char* moca_last_error(void) { return g_err; }
extern "C" int32_t moca_abi_version(void) { return 2; }
extern "C" int32_t moca_operand_dtype(void) { return MOCA_OP_IS_BF16 ? MOCA_DTYPE_BF16 : MOCA_DTYPE_F16; }

extern "C" int32_t moca_ctx_create(int32_t device, moca_ctx** out) {
    MOCA_REQUIRE(out != nullptr, MOCA_E_ARG, "moca_ctx_create: out is NULL");
    *out = nullptr;
    int ndev = 0;
    MOCA_CHECK_CUDA(cudaGetDeviceCount(&ndev));
    MOCA_REQUIRE(device >= 0 && device < ndev, MOCA_E_ARG, "moca_ctx_create: device %d of %d", device, ndev);
    cudaDeviceProp prop;
    MOCA_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
    // no fallback by design: the kernels are sm_100a SASS (tcgen05 / TMEM / TMA)
    MOCA_REQUIRE(prop.major == 10, MOCA_E_ARCH, "moca_ctx_create: device %d is sm_%d%d; this library is built for sm_100a only",
                 device, prop.major, prop.minor);
    // allocate on `device` without changing the caller's current device
    int prev_dev = 0;
    MOCA_CHECK_CUDA(cudaGetDevice(&prev_dev));
    MOCA_CHECK_CUDA(cudaSetDevice(device));
    moca_ctx* c = new moca_ctx();
    c->c.device = device;
    c->c.num_sms = prop.multiProcessorCount;
    c->c.cc_major = prop.major;
    c->c.cc_minor = prop.minor;
    cudaError_t e1 = cudaMalloc(&c->c.d_mask_sums, sizeof(double) * MOCA_MAX_MASKS);
    cudaError_t e2 = e1 == cudaSuccess ? cudaMemset(c->c.d_mask_sums, 0, sizeof(double) * MOCA_MAX_MASKS) : e1;
    cudaSetDevice(prev_dev);
    if (e2 != cudaSuccess) {
        set_error("moca_ctx_create: scratch allocation failed: %s", cudaGetErrorString(e2));
        delete c;
        return MOCA_E_CUDA;
    }
    *out = c;
    return MOCA_OK;
}

extern "C" int32_t moca_comm_destroy(moca_ctx* ctx);

extern "C" int32_t moca_ctx_destroy(moca_ctx* ctx) {
    if (!ctx) return MOCA_OK;
    moca_comm_destroy(ctx);
    cudaFree(ctx->c.d_mask_sums);
    delete ctx;
    return MOCA_OK;
}
