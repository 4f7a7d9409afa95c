// hostsim: test double of the CUDA runtime entry points the elb engine uses (see sim.hpp).
// "Device" memory is host memory tracked in a registry (every access by a simulated kernel or
// copy is bounds-checked against it, freed memory is poisoned); streams are FIFO worker threads;
// every event record is its own object (a wait captures the record that was latest when it was issued); cuStreamWaitValue32 is a stream-ordered spin on a memory word.
#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <thread>
#include <vector>

#include "sim.hpp"

namespace sim {

[[noreturn]] void fail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    std::fprintf(stderr, "[hostsim] FATAL: ");
    std::vfprintf(stderr, fmt, ap);
    std::fprintf(stderr, "\n");
    va_end(ap);
    std::fflush(stderr);
    std::_Exit(97);
}

double wait_timeout_seconds() {
    static double t = [] { const char* e = std::getenv("HOSTSIM_TIMEOUT"); return e && std::atof(e) > 0 ? std::atof(e) : 120.0; }();
    return t;
}

namespace {

using Clock = std::chrono::steady_clock;

// ------------------------------------------------------------------ memory registry
std::mutex g_mem_mutex;
std::map<uintptr_t, size_t> g_live;  // base -> bytes

void* sim_alloc(size_t bytes) {
    const size_t n = bytes ? bytes : 1;
    void* p = nullptr;
    if (posix_memalign(&p, 256, (n + 255) / 256 * 256) != 0 || !p) return nullptr;
    // fresh device memory holds garbage: signalling pattern (NaN as fp32 and fp64) for small blocks,
    // large arenas are left untouched so that they stay lazily committed
    if (n <= (size_t(64) << 20)) std::memset(p, 0xFF, n);
    std::lock_guard<std::mutex> l(g_mem_mutex);
    g_live[reinterpret_cast<uintptr_t>(p)] = n;
    return p;
}

void sim_free(void* p) {
    if (!p) return;
    size_t n = 0;
    {
        std::lock_guard<std::mutex> l(g_mem_mutex);
        auto it = g_live.find(reinterpret_cast<uintptr_t>(p));
        if (it == g_live.end()) fail("free of a pointer that is not a live device allocation: %p", p);
        n = it->second;
        g_live.erase(it);
    }
    if (n <= (size_t(64) << 20)) std::memset(p, 0xFF, n);  // use-after-free reads NaN
    std::free(p);
}

// ------------------------------------------------------------------ streams
// A hardware work queue: one worker thread executing closures in FIFO order. By default every stream has a
// queue of its own (streams are fully concurrent). HOSTSIM_HWQ=N models CUDA_DEVICE_MAX_CONNECTIONS=N: the
// streams of a rank ("device") are spread round-robin over N queues, and work of different streams that lands
// in the same queue runs in ISSUE order — so a stream-ordered wait (an event, a cuStreamWaitValue32 on a peer's
// flag) at the head of a queue also holds back the other streams behind it, the false dependency real hardware
// has. A schedule that only works because streams never share a queue deadlocks here (and is reported by the
// wait timeouts) instead of on an 8-GPU box.
struct HwQueue {
    std::mutex m;
    std::condition_variable cv;
    std::deque<std::function<void()>> q;
    bool stop = false;
    std::thread th;
    HwQueue() {
        th = std::thread([this] {
            for (;;) {
                std::function<void()> fn;
                {
                    std::unique_lock<std::mutex> l(m);
                    cv.wait(l, [this] { return stop || !q.empty(); });
                    if (q.empty()) return;
                    fn = std::move(q.front());
                    q.pop_front();
                }
                fn();
            }
        });
    }
    void push(std::function<void()> fn) {
        {
            std::lock_guard<std::mutex> l(m);
            q.push_back(std::move(fn));
        }
        cv.notify_one();
    }
    ~HwQueue() {
        {
            std::lock_guard<std::mutex> l(m);
            stop = true;
        }
        cv.notify_one();
        if (th.joinable()) th.join();
    }
};

int hwq_count() {
    static int n = [] { const char* e = std::getenv("HOSTSIM_HWQ"); return e ? std::atoi(e) : 0; }();
    return n > 0 ? n : 0;
}
// the shared queues of the calling rank (host thread); leaked at exit on purpose
thread_local std::vector<HwQueue*> t_hwq;
thread_local int t_hwq_next = 0;

struct Stream {
    std::mutex m;
    std::condition_variable idle;
    uint64_t pending = 0;  // closures pushed and not yet finished
    std::thread::id owner = std::this_thread::get_id();  // the rank (host thread) whose "device" this stream is on
    HwQueue* hw = nullptr;
    bool own_hw = false;
    Stream() {
        const int n = hwq_count();
        if (n == 0) {
            hw = new HwQueue();
            own_hw = true;
        } else {
            if (int(t_hwq.size()) < n) t_hwq.push_back(new HwQueue());
            hw = t_hwq[size_t(t_hwq_next++ % n) % t_hwq.size()];
        }
    }
    void push(std::function<void()> fn) {
        {
            std::lock_guard<std::mutex> l(m);
            ++pending;
        }
        hw->push([this, fn = std::move(fn)] {
            fn();
            {
                std::lock_guard<std::mutex> l(m);
                --pending;
                if (pending == 0) idle.notify_all();
            }
        });
    }
    void sync() {
        std::unique_lock<std::mutex> l(m);
        if (!idle.wait_for(l, std::chrono::duration<double>(wait_timeout_seconds()), [this] { return pending == 0; }))
            fail("cudaStreamSynchronize: stream %p did not drain within %.0f s (deadlock?)", (void*)this, wait_timeout_seconds());
    }
    ~Stream() {
        if (own_hw) delete hw;
    }
};

std::mutex g_streams_mutex;
std::set<Stream*> g_streams;
// the legacy default stream of the calling rank (one process per GPU in reality: one per host thread here)
Stream* default_stream() {
    thread_local Stream* s = [] {
        Stream* d = new Stream();
        std::lock_guard<std::mutex> l(g_streams_mutex);
        g_streams.insert(d);
        return d;
    }();
    return s;
}
Stream* S(cudaStream_t s) {
    if (!s) return default_stream();
    Stream* st = reinterpret_cast<Stream*>(s);
    std::lock_guard<std::mutex> l(g_streams_mutex);
    if (!g_streams.count(st)) fail("use of a destroyed or foreign stream handle %p", (void*)s);
    return st;
}

// ------------------------------------------------------------------ events
// One cudaEventRecord = one Record object. A wait (cudaStreamWaitEvent, cudaEventSynchronize) captures the
// record that was the latest when the wait was ISSUED and waits for exactly that one — re-recording the
// event afterwards, on this or another stream, neither satisfies nor delays it (CUDA semantics; a
// "highest completed generation" counter would be wrong when one event is recorded on two streams).
struct Record {
    std::mutex m;
    std::condition_variable cv;
    bool done = false;
    Clock::time_point when;
};
struct Event {
    std::mutex m;
    std::shared_ptr<Record> latest;  // null until the first record
};

void wait_record(const std::shared_ptr<Record>& r, const char* who) {
    std::unique_lock<std::mutex> l(r->m);
    if (!r->cv.wait_for(l, std::chrono::duration<double>(wait_timeout_seconds()), [&] { return r->done; }))
        fail("%s: the recorded work never completed (deadlock?)", who);
}

void copy_bytes(void* dst, const void* src, size_t n) {
    // 4-byte aligned word copies are the arena's flag writes: make them release stores so that a
    // spinning cuStreamWaitValue32 on another thread sees the payload written before them
    if (n == 4 && (reinterpret_cast<uintptr_t>(dst) & 3u) == 0 && (reinterpret_cast<uintptr_t>(src) & 3u) == 0) {
        uint32_t v;
        std::memcpy(&v, src, 4);
        __atomic_store_n(static_cast<uint32_t*>(dst), v, __ATOMIC_RELEASE);
        return;
    }
    std::memmove(dst, src, n);
}

bool dst_is_device(cudaMemcpyKind k) { return k == cudaMemcpyHostToDevice || k == cudaMemcpyDeviceToDevice; }
bool src_is_device(cudaMemcpyKind k) { return k == cudaMemcpyDeviceToHost || k == cudaMemcpyDeviceToDevice; }

CUresult sim_cuStreamWaitValue32(CUstream stream, CUdeviceptr addr, cuuint32_t value, unsigned int flags) {
    uint32_t* p = reinterpret_cast<uint32_t*>(addr);
    check_device_range(p, 4, "cuStreamWaitValue32");
    enqueue(reinterpret_cast<cudaStream_t>(stream), [p, value, flags] {
        const auto t0 = Clock::now();
        for (uint64_t spin = 0;; ++spin) {
            const uint32_t v = __atomic_load_n(p, __ATOMIC_ACQUIRE);
            const bool ok = (flags & 3u) == CU_STREAM_WAIT_VALUE_EQ ? v == value : int32_t(v - value) >= 0;
            if (ok) return;
            if (spin < 64) std::this_thread::yield();
            else std::this_thread::sleep_for(std::chrono::microseconds(50));
            if ((spin & 1023u) == 1023u &&
                std::chrono::duration<double>(Clock::now() - t0).count() > wait_timeout_seconds())
                fail("cuStreamWaitValue32: word at %p stayed %u, waiting for %u (deadlock?)", (void*)p, v, value);
        }
    });
    return CUDA_SUCCESS;
}

}  // namespace

void enqueue(cudaStream_t s, std::function<void()> fn) { S(s)->push(std::move(fn)); }

void check_device_range(const void* p, size_t bytes, const char* what) {
    if (bytes == 0) return;
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    std::lock_guard<std::mutex> l(g_mem_mutex);
    auto it = g_live.upper_bound(a);
    if (it != g_live.begin()) {
        --it;
        if (a >= it->first && a + bytes <= it->first + it->second) return;
    }
    fail("%s: [%p, +%zu) is not inside a live device allocation", what, p, bytes);
}

}  // namespace sim

using namespace sim;

extern "C" {

cudaError_t cudaSetDevice(int) { return cudaSuccess; }
cudaError_t cudaGetDevice(int* d) { if (d) *d = 0; return cudaSuccess; }
cudaError_t cudaGetLastError(void) { return cudaSuccess; }
const char* cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "hostsim error"; }
cudaError_t cudaDeviceGetDefaultMemPool(cudaMemPool_t*, int) { return cudaErrorNotSupported; }
cudaError_t cudaMemPoolSetAttribute(cudaMemPool_t, enum cudaMemPoolAttr, void*) { return cudaErrorNotSupported; }

cudaError_t cudaMalloc(void** p, size_t n) {
    *p = sim_alloc(n);
    return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
cudaError_t cudaFree(void* p) {
    // cudaFree synchronises the device before releasing
    cudaDeviceSynchronize();
    sim_free(p);
    return cudaSuccess;
}
cudaError_t cudaMallocAsync(void** p, size_t n, cudaStream_t s) {
    (void)S(s);
    *p = sim_alloc(n);
    return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
cudaError_t cudaFreeAsync(void* p, cudaStream_t s) {
    if (!p) return cudaSuccess;
    enqueue(s, [p] { sim_free(p); });
    return cudaSuccess;
}

cudaError_t cudaMemcpyAsync(void* dst, const void* src, size_t n, enum cudaMemcpyKind kind, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    if (!dst || !src) return cudaErrorInvalidValue;
    enqueue(s, [=] {
        if (dst_is_device(kind)) check_device_range(dst, n, "cudaMemcpyAsync dst");
        if (src_is_device(kind)) check_device_range(src, n, "cudaMemcpyAsync src");
        copy_bytes(dst, src, n);
    });
    return cudaSuccess;
}
cudaError_t cudaMemcpy(void* dst, const void* src, size_t n, enum cudaMemcpyKind kind) {
    cudaError_t e = cudaMemcpyAsync(dst, src, n, kind, nullptr);
    if (e != cudaSuccess) return e;
    return cudaStreamSynchronize(nullptr);
}
cudaError_t cudaMemcpy2DAsync(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t height,
                              enum cudaMemcpyKind kind, cudaStream_t s) {
    if (width == 0 || height == 0) return cudaSuccess;
    if (!dst || !src || dpitch < width || spitch < width) return cudaErrorInvalidValue;
    enqueue(s, [=] {
        if (dst_is_device(kind)) check_device_range(dst, (height - 1) * dpitch + width, "cudaMemcpy2DAsync dst");
        if (src_is_device(kind)) check_device_range(src, (height - 1) * spitch + width, "cudaMemcpy2DAsync src");
        for (size_t r = 0; r < height; ++r)
            std::memmove(static_cast<char*>(dst) + r * dpitch, static_cast<const char*>(src) + r * spitch, width);
    });
    return cudaSuccess;
}
cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    enqueue(s, [=] {
        check_device_range(p, n, "cudaMemsetAsync");
        std::memset(p, v, n);
    });
    return cudaSuccess;
}
cudaError_t cudaMemset(void* p, int v, size_t n) {
    cudaMemsetAsync(p, v, n, nullptr);
    return cudaStreamSynchronize(nullptr);
}

cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned int) {
    Stream* st = new Stream();
    {
        std::lock_guard<std::mutex> l(g_streams_mutex);
        g_streams.insert(st);
    }
    *s = reinterpret_cast<cudaStream_t>(st);
    return cudaSuccess;
}
cudaError_t cudaStreamCreate(cudaStream_t* s) { return cudaStreamCreateWithFlags(s, 0); }
cudaError_t cudaStreamSynchronize(cudaStream_t s) {
    S(s)->sync();
    return cudaSuccess;
}
cudaError_t cudaStreamDestroy(cudaStream_t s) {
    if (!s) return cudaErrorInvalidResourceHandle;
    Stream* st = S(s);
    st->sync();  // CUDA lets queued work finish before the stream goes away
    {
        std::lock_guard<std::mutex> l(g_streams_mutex);
        g_streams.erase(st);
    }
    delete st;
    return cudaSuccess;
}
cudaError_t cudaDeviceSynchronize(void) {
    // one rank = one host thread = one device: only the caller's own streams are drained (draining
    // another rank's streams could wait on a collective that rank has not joined yet)
    std::vector<Stream*> all;
    {
        std::lock_guard<std::mutex> l(g_streams_mutex);
        for (Stream* st : g_streams)
            if (st->owner == std::this_thread::get_id()) all.push_back(st);
    }
    for (Stream* st : all) {
        // a stream may be destroyed by another rank's thread meanwhile: re-validate under the lock
        bool alive;
        {
            std::lock_guard<std::mutex> l(g_streams_mutex);
            alive = g_streams.count(st) != 0;
        }
        if (alive) st->sync();
    }
    return cudaSuccess;
}

cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned int) {
    *e = reinterpret_cast<cudaEvent_t>(new Event());
    return cudaSuccess;
}
cudaError_t cudaEventCreate(cudaEvent_t* e) { return cudaEventCreateWithFlags(e, 0); }
cudaError_t cudaEventDestroy(cudaEvent_t e) {
    delete reinterpret_cast<Event*>(e);  // queued records / waits hold their Record through shared_ptr
    return cudaSuccess;
}
cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t s) {
    if (!e) return cudaErrorInvalidResourceHandle;
    Event* ev = reinterpret_cast<Event*>(e);
    auto rec = std::make_shared<Record>();
    {
        std::lock_guard<std::mutex> l(ev->m);
        ev->latest = rec;
    }
    enqueue(s, [rec] {
        {
            std::lock_guard<std::mutex> l(rec->m);
            rec->done = true;
            rec->when = Clock::now();
        }
        rec->cv.notify_all();
    });
    return cudaSuccess;
}
static std::shared_ptr<Record> latest_record(cudaEvent_t e) {
    Event* ev = reinterpret_cast<Event*>(e);
    std::lock_guard<std::mutex> l(ev->m);
    return ev->latest;
}
cudaError_t cudaStreamWaitEvent(cudaStream_t s, cudaEvent_t e, unsigned int) {
    if (!e) return cudaErrorInvalidResourceHandle;
    std::shared_ptr<Record> rec = latest_record(e);  // the most recent record at the time of THIS call
    if (!rec) return cudaSuccess;                     // never recorded: no-op
    enqueue(s, [rec] { wait_record(rec, "cudaStreamWaitEvent"); });
    return cudaSuccess;
}
cudaError_t cudaEventSynchronize(cudaEvent_t e) {
    if (!e) return cudaErrorInvalidResourceHandle;
    std::shared_ptr<Record> rec = latest_record(e);
    if (rec) wait_record(rec, "cudaEventSynchronize");
    return cudaSuccess;
}
cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t a, cudaEvent_t b) {
    if (!a || !b) return cudaErrorInvalidResourceHandle;
    auto ra = latest_record(a), rb = latest_record(b);
    if (!ra || !rb) return cudaErrorInvalidResourceHandle;
    *ms = std::chrono::duration<float, std::milli>(rb->when - ra->when).count();
    return cudaSuccess;
}

// CUDA IPC inside one process: the handle is the pointer
cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t* h, void* p) {
    check_device_range(p, 1, "cudaIpcGetMemHandle");
    std::memset(h, 0, sizeof(*h));
    std::memcpy(h->reserved, &p, sizeof(p));
    return cudaSuccess;
}
cudaError_t cudaIpcOpenMemHandle(void** p, cudaIpcMemHandle_t h, unsigned int) {
    std::memcpy(p, h.reserved, sizeof(*p));
    check_device_range(*p, 1, "cudaIpcOpenMemHandle");
    return cudaSuccess;
}
cudaError_t cudaIpcCloseMemHandle(void*) { return cudaSuccess; }

cudaError_t cudaGetDriverEntryPoint(const char* symbol, void** fn, unsigned long long,
                                    enum cudaDriverEntryPointQueryResult* status) {
    if (symbol && std::strcmp(symbol, "cuStreamWaitValue32") == 0) {
        *fn = reinterpret_cast<void*>(&sim_cuStreamWaitValue32);
        if (status) *status = cudaDriverEntryPointSuccess;
        return cudaSuccess;
    }
    *fn = nullptr;
    if (status) *status = cudaDriverEntryPointSymbolNotFound;
    return cudaErrorInvalidValue;
}

}  // extern "C"

extern "C" cudaError_t cudaGetDeviceCount(int* n) {
    if (n) *n = 1;
    return cudaSuccess;
}
