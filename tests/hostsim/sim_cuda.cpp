// hostsim: test double of the CUDA runtime entry points the elb engine uses (see sim.hpp).
// "Device" memory is host memory tracked in a registry (every access by a simulated kernel or
// copy is bounds-checked against it, freed memory is poisoned); streams are FIFO worker threads;
// every event record is its own object (a wait captures the record that was latest when it was issued); cuStreamWaitValue32 is a stream-ordered spin on a memory word.
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
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

bool vtime_on() {
    static bool v = [] { const char* e = std::getenv("HOSTSIM_VTIME"); return e && std::atoi(e) != 0; }();
    return v;
}
bool dry_run() {
    static bool v = [] { const char* e = std::getenv("HOSTSIM_DRY"); return e && std::atoi(e) != 0; }();
    return v;
}

namespace {

using Clock = std::chrono::steady_clock;

// ------------------------------------------------------------------ virtual time
std::atomic<int> g_next_dev{0};
thread_local int t_dev = -1;            // device (rank) id of the calling HOST thread, assigned on first use
thread_local double t_host_vtime = 0;   // host thread's virtual clock: advanced by synchronising calls
int host_dev() {
    if (t_dev < 0) t_dev = g_next_dev.fetch_add(1);
    return t_dev;
}
constexpr int kMaxDev = 64;
std::mutex g_res_mutex;
// Busy intervals of one exclusive resource. Worker threads reach the reservation in REAL-time order, which is not
// virtual-time order, so "free again at" is not enough: an operation takes the earliest gap that fits it.
struct Busy {
    std::map<double, double> iv;  // start -> end, disjoint
    double find(double t, double dur) const {
        auto it = iv.upper_bound(t);
        if (it != iv.begin()) {
            auto prev = std::prev(it);
            if (prev->second > t) t = prev->second;
        }
        while (it != iv.end() && it->first < t + dur - 1e-12) {
            if (it->second > t) t = it->second;
            ++it;
        }
        return t;
    }
    void add(double t, double dur) {
        auto it = iv.upper_bound(t);
        if (it != iv.begin()) {
            auto prev = std::prev(it);
            if (prev->second >= t - 1e-12) {  // touches the interval before it: extend that one
                if (prev->second < t + dur) prev->second = t + dur;
                return;
            }
        }
        iv[t] = t + dur;
    }
};
Busy g_res_busy[kMaxDev][RES_COUNT];         // tensor pipe, HBM kernels, PCIe up / down of a device
Busy g_nvl_out[kMaxDev], g_nvl_in[kMaxDev];  // NVLink egress / ingress of a device
std::mutex g_flag_mutex;
std::map<uintptr_t, double> g_flag_time;       // virtual time of the last 4-byte (flag) write to an address
struct Stream;
thread_local Stream* t_cur = nullptr;          // stream whose closure the calling WORKER thread is executing

// ------------------------------------------------------------------ memory registry
std::mutex g_mem_mutex;
std::map<uintptr_t, size_t> g_live;  // base -> bytes
std::map<uintptr_t, int> g_live_dev;  // base -> device (rank) that allocated it

void* sim_alloc(size_t bytes) {
    const size_t n = bytes ? bytes : 1;
    void* p = nullptr;
    if (posix_memalign(&p, 256, (n + 255) / 256 * 256) != 0 || !p) return nullptr;
    // fresh device memory holds garbage: signalling pattern (NaN as fp32 and fp64) for small blocks,
    // large arenas are left untouched so that they stay lazily committed
    if (n <= (size_t(64) << 20) && !dry_run()) std::memset(p, 0xFF, n);
    std::lock_guard<std::mutex> l(g_mem_mutex);
    g_live[reinterpret_cast<uintptr_t>(p)] = n;
    g_live_dev[reinterpret_cast<uintptr_t>(p)] = host_dev();
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
        g_live_dev.erase(reinterpret_cast<uintptr_t>(p));
    }
    if (n <= (size_t(64) << 20) && !dry_run()) std::memset(p, 0xFF, n);  // use-after-free reads NaN
    if (vtime_on()) {  // flag write times of this allocation die with it (the address may come back)
        std::lock_guard<std::mutex> l(g_flag_mutex);
        const uintptr_t a = reinterpret_cast<uintptr_t>(p);
        g_flag_time.erase(g_flag_time.lower_bound(a), g_flag_time.lower_bound(a + n));
    }
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
    double vtime = 0;       // virtual clock (only touched by the worker executing this stream's closures, and after sync)
    int dev = host_dev();  // device (rank) this stream belongs to
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
        const double issued = t_host_vtime;
        hw->push([this, issued, fn = std::move(fn)] {
            t_cur = this;
            if (vtime < issued) vtime = issued;  // work cannot start before the host enqueued it
            fn();
            t_cur = nullptr;
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
        if (t_host_vtime < vtime) t_host_vtime = vtime;  // the host has waited for everything on this stream
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
    double vt = 0;  // virtual time of the stream when the record executed
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

int device_of(const void* p) {  // -1: not device memory (host)
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    std::lock_guard<std::mutex> l(g_mem_mutex);
    auto it = g_live.upper_bound(a);
    if (it == g_live.begin()) return -1;
    --it;
    if (a < it->first || a >= it->first + it->second) return -1;
    return g_live_dev[it->first];
}

// virtual duration of a copy of n bytes, by where source and destination live
void account_copy(void* dst, const void* src, size_t n) {
    if (!vtime_on() || !t_cur) return;
    if (n <= 64) {  // flags and scalars: a latency, no bandwidth
        t_cur->vtime += kFlagLatency;
        if (n == 4) {
            std::lock_guard<std::mutex> l(g_flag_mutex);
            g_flag_time[reinterpret_cast<uintptr_t>(dst)] = t_cur->vtime;
        }
        return;
    }
    const int ds = device_of(src), dd = device_of(dst);
    if (ds < 0 && dd >= 0) { vt_op(RES_H2D, kLaunch + double(n) / kRatePcie); return; }
    if (ds >= 0 && dd < 0) { vt_op(RES_D2H, kLaunch + double(n) / kRatePcie); return; }
    if (ds == dd) {
        vt_trace("d2d-local", t_cur->vtime, kLaunch + double(n) / kRateLocalCopy, double(n));
        t_cur->vtime += kLaunch + double(n) / kRateLocalCopy;
        return;
    }
    // peer copy over NVLink: occupies the source's egress and the destination's ingress
    const double dur = kLaunch + double(n) / kRateNvlink;
    std::lock_guard<std::mutex> l(g_res_mutex);
    double start = t_cur->vtime;
    Busy& out = g_nvl_out[ds % kMaxDev];
    Busy& in = g_nvl_in[dd % kMaxDev];
    for (;;) {  // earliest time both the source's egress and the destination's ingress have a gap of `dur`
        const double a = out.find(start, dur), b = in.find(a, dur);
        start = b;
        if (a == b) break;
    }
    out.add(start, dur);
    in.add(start, dur);
    vt_trace("d2d-peer", start, dur, double(n));
    t_cur->vtime = start + dur;
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
    if (dry_run() && n > kDryKeep) return;  // payloads are not moved in a dry run; flags, scalars and control records are
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
            if (ok) {
                if (vtime_on()) {
                    double tw = 0;
                    {
                        std::lock_guard<std::mutex> l(g_flag_mutex);
                        auto it = g_flag_time.find(reinterpret_cast<uintptr_t>(p));
                        if (it != g_flag_time.end()) tw = it->second;
                    }
                    vt_at_least(tw);
                }
                return;
            }
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

double vt_now() { return t_cur ? t_cur->vtime : t_host_vtime; }
void vt_at_least(double t) {
    if (t_cur) { if (t_cur->vtime < t) t_cur->vtime = t; }
    else if (t_host_vtime < t) t_host_vtime = t;
}
// HOSTSIM_VTRACE=<file>: one line per modelled operation — device, stream, what, start and duration in microseconds
void vt_trace(const char* what, double start, double seconds, double bytes) {
    static FILE* f = [] { const char* e = std::getenv("HOSTSIM_VTRACE"); return e && *e ? std::fopen(e, "w") : nullptr; }();
    if (!f) return;
    static std::mutex m;
    std::lock_guard<std::mutex> l(m);
    std::fprintf(f, "dev %d stream %p %-10s start %12.2f us  dur %10.2f us  bytes %.0f\n", t_cur ? t_cur->dev : -1, (void*)t_cur, what,
                 start * 1e6, seconds * 1e6, bytes);
    std::fflush(f);
}
void vt_op(Res res, double seconds) {
    if (!vtime_on() || !t_cur) return;
    double start = t_cur->vtime;
    if (res != RES_NONE) {
        std::lock_guard<std::mutex> l(g_res_mutex);
        Busy& b = g_res_busy[t_cur->dev % kMaxDev][res];
        start = b.find(start, seconds);
        b.add(start, seconds);
    }
    static const char* names[RES_COUNT] = {"op", "tensor", "hbm", "h2d", "d2h"};
    vt_trace(names[res], start, seconds, 0);
    t_cur->vtime = start + seconds;
}

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
        account_copy(dst, src, n);
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
        account_copy(dst, src, width * height);
        if (dry_run()) return;
        for (size_t r = 0; r < height; ++r)
            std::memmove(static_cast<char*>(dst) + r * dpitch, static_cast<const char*>(src) + r * spitch, width);
    });
    return cudaSuccess;
}
cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    enqueue(s, [=] {
        check_device_range(p, n, "cudaMemsetAsync");
        vt_op(RES_HBM, kLaunch + double(n) / kRateHbm);
        if (dry_run() && n > kDryKeep) return;
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
            rec->vt = vt_now();
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
    enqueue(s, [rec] { wait_record(rec, "cudaStreamWaitEvent"); vt_at_least(rec->vt); });
    return cudaSuccess;
}
cudaError_t cudaEventSynchronize(cudaEvent_t e) {
    if (!e) return cudaErrorInvalidResourceHandle;
    std::shared_ptr<Record> rec = latest_record(e);
    if (rec) { wait_record(rec, "cudaEventSynchronize"); vt_at_least(rec->vt); }
    return cudaSuccess;
}
cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t a, cudaEvent_t b) {
    if (!a || !b) return cudaErrorInvalidResourceHandle;
    auto ra = latest_record(a), rb = latest_record(b);
    if (!ra || !rb) return cudaErrorInvalidResourceHandle;
    if (vtime_on()) *ms = float((rb->vt - ra->vt) * 1e3);  // modelled milliseconds
    else *ms = std::chrono::duration<float, std::milli>(rb->when - ra->when).count();
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
