// status.cu -- version and thread-local error reporting of libb200cs
#include <stdarg.h>

#include "common.cuh"

namespace b200cs {

char* err_buf() {
    static thread_local char buf[512] = {0};
    return buf;
}

int set_err(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err_buf(), 512, fmt, ap);
    va_end(ap);
    return code;
}

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_err(B200CS_ECUDA, "%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
    return B200CS_OK;
}

}  // namespace b200cs

extern "C" int b200cs_version(void) { return B200CS_VERSION; }
extern "C" const char* b200cs_last_error(void) { return b200cs::err_buf(); }
