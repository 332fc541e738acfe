// host_stage.cu -- b200cs_host_copy: multi-threaded copy of a caller's (pageable) host array into a
// pinned staging buffer.
//
// The drop-in entry points take plain numpy arrays (external/vcm/vcm/cubedsphere/regridz.py:31-78,
// coarsen.py:183-218): pageable memory.  cudaMemcpyAsync from pageable memory is a synchronous
// staged copy through the driver's bounce buffers on ONE host thread (~10 GB/s), five times slower
// than the PCIe link; pinning the caller's arrays with cudaHostRegister costs more than the copy.
// So the host path of core.pressure_coarsen copies each array slab into a reused ring of pinned
// buffers with all host cores (this function, memory-bandwidth bound) while the DMA of the
// previous slab and the kernels of the previous tile run.
#include <string.h>

#include <thread>
#include <vector>

#include "common.cuh"

extern "C" int b200cs_host_copy(void* dst, const void* src, int64_t bytes, int nthreads) {
    B200CS_REQUIRE(bytes >= 0 && (bytes == 0 || (dst && src)), "host_copy: NULL argument");
    if (bytes == 0) return B200CS_OK;
    const int64_t kMin = 4 << 20;  // below 4 MB per thread a thread costs more than it copies
    int64_t nt = nthreads < 1 ? 1 : nthreads;
    if (nt > 64) nt = 64;
    if (nt > (bytes + kMin - 1) / kMin) nt = (bytes + kMin - 1) / kMin;
    if (nt <= 1) {
        memcpy(dst, src, (size_t)bytes);
        return B200CS_OK;
    }
    const int64_t chunk = ((bytes + nt - 1) / nt + 4095) & ~(int64_t)4095;
    std::vector<std::thread> pool;
    pool.reserve((size_t)nt - 1);
    for (int64_t t = 1; t < nt; t++) {
        const int64_t o = t * chunk;
        if (o >= bytes) break;
        const int64_t n = bytes - o < chunk ? bytes - o : chunk;
        pool.emplace_back([=] { memcpy((char*)dst + o, (const char*)src + o, (size_t)n); });
    }
    memcpy(dst, src, (size_t)(chunk < bytes ? chunk : bytes));
    for (auto& th : pool) th.join();
    return B200CS_OK;
}
