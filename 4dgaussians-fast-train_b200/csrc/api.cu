// Library info entry points of the fdgs C ABI (include/fdgs.h).
#include "fdgs_common.cuh"

extern "C" {

int fdgs_abi_version(void) { return FDGS_ABI_VERSION; }

int fdgs_compiled_arch(void) {
#ifdef FDGS_ARCH
    return FDGS_ARCH;
#else
    return 100;
#endif
}

const char* fdgs_error_string(int code) {
    switch (code) {
        case FDGS_OK: return "ok";
        case FDGS_E_BADARG: return "fdgs: bad argument (null pointer, inconsistent shapes or exclusive inputs)";
        case FDGS_E_WORKSPACE: return "fdgs: caller-owned workspace too small";
        case FDGS_E_UNSUPPORTED: return "fdgs: configuration not supported by the sm_100a kernels";
        case FDGS_E_SMEM: return "fdgs: configuration exceeds the shared-memory budget";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "fdgs: unknown error";
}

}  // extern "C"
