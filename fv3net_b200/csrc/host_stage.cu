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
//
// r02: (1) a persistent pool of workers (a 47 MB slab is 1 ms of copying; creating and joining a
// dozen threads per slab was a quarter of that); (2) non-temporal stores: the staging slab is read
// next by the DMA engine, not by a core, so the destination lines are neither fetched first
// (read-for-ownership) nor left in the caches -- a third less memory traffic per byte copied;
// (3) 1 MB parts handed out through an atomic counter, so a descheduled worker does not hold the
// slab back.
#include <emmintrin.h>
#include <pthread.h>
#include <string.h>

#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

// dst and src do not overlap
void copy_stream(char* dst, const char* src, size_t n) {
    if (n < 256) {
        memcpy(dst, src, n);
        return;
    }
    const size_t head = (64 - ((uintptr_t)dst & 63)) & 63;
    if (head) {
        memcpy(dst, src, head);
        dst += head; src += head; n -= head;
    }
    const size_t body = n & ~(size_t)63;
    for (size_t i = 0; i < body; i += 64) {
        const __m128i a = _mm_loadu_si128((const __m128i*)(src + i));
        const __m128i b = _mm_loadu_si128((const __m128i*)(src + i + 16));
        const __m128i c = _mm_loadu_si128((const __m128i*)(src + i + 32));
        const __m128i d = _mm_loadu_si128((const __m128i*)(src + i + 48));
        _mm_stream_si128((__m128i*)(dst + i), a);
        _mm_stream_si128((__m128i*)(dst + i + 16), b);
        _mm_stream_si128((__m128i*)(dst + i + 32), c);
        _mm_stream_si128((__m128i*)(dst + i + 48), d);
    }
    if (n > body) memcpy(dst + body, src + body, n - body);
    _mm_sfence();  // the DMA that reads the slab is queued after this function returns
}

constexpr int64_t kPart = 1 << 20;

struct Pool {
    std::mutex call;  // one copy at a time
    std::mutex m;
    std::condition_variable wake, done;
    std::vector<std::thread> workers;
    // the job
    char* dst = nullptr;
    const char* src = nullptr;
    int64_t bytes = 0, nparts = 0;
    std::atomic<int64_t> next{0};
    uint64_t generation = 0;
    int wanted = 0;    // workers that should take part in this generation
    int running = 0;   // workers that have not finished this generation yet

    void work() {
        for (;;) {
            const int64_t p = next.fetch_add(1, std::memory_order_relaxed);
            if (p >= nparts) return;
            const int64_t o = p * kPart;
            copy_stream(dst + o, src + o, (size_t)(bytes - o < kPart ? bytes - o : kPart));
        }
    }
    void loop(int id) {
        uint64_t seen = 0;
        std::unique_lock<std::mutex> lk(m);
        for (;;) {
            wake.wait(lk, [&] { return generation != seen; });
            seen = generation;
            if (id >= wanted) continue;
            lk.unlock();
            work();
            lk.lock();
            if (--running == 0) done.notify_one();
        }
    }
    void run(char* d, const char* s, int64_t n, int nt) {
        std::lock_guard<std::mutex> guard(call);
        {
            std::unique_lock<std::mutex> lk(m);
            while ((int)workers.size() < nt - 1) {
                const int id = (int)workers.size();
                workers.emplace_back([this, id] { loop(id); });
                workers.back().detach();  // live until the process exits
            }
            dst = d; src = s; bytes = n; nparts = (n + kPart - 1) / kPart;
            next.store(0, std::memory_order_relaxed);
            wanted = nt - 1; running = nt - 1;
            generation++;
        }
        wake.notify_all();
        work();  // the caller takes parts too
        std::unique_lock<std::mutex> lk(m);
        done.wait(lk, [&] { return running == 0; });
    }
};

Pool* g_pool = nullptr;  // never destroyed: detached workers may outlive static destructors
std::once_flag g_pool_once;

Pool& pool() {
    std::call_once(g_pool_once, [] {
        g_pool = new Pool;
        // a forked child has none of the parent's threads: it starts with a pool of its own
        pthread_atfork(nullptr, nullptr, [] { g_pool = new Pool; });
    });
    return *g_pool;
}

}  // namespace

extern "C" int b200cs_host_copy(void* dst, const void* src, int64_t bytes, int nthreads) {
    B200CS_REQUIRE(bytes >= 0 && (bytes == 0 || (dst && src)), "host_copy: NULL argument");
    if (bytes == 0) return B200CS_OK;
    int64_t nt = nthreads < 1 ? 1 : nthreads;
    if (nt > 64) nt = 64;
    const int64_t nparts = (bytes + kPart - 1) / kPart;
    if (nt > nparts) nt = nparts;
    if (nt <= 1) {
        copy_stream((char*)dst, (const char*)src, (size_t)bytes);
        return B200CS_OK;
    }
    pool().run((char*)dst, (const char*)src, bytes, (int)nt);
    return B200CS_OK;
}
