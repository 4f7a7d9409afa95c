// hostsim — TEST INFRASTRUCTURE ONLY. A host-side test double of the pieces libelb200's C++ engine
// stands on (CUDA runtime streams/events/memory, NCCL, the elb200_* device kernels), so that the
// engine's host logic — SUMMA schedules, redistribution exchanges, the peer-arena handshake, stream
// and event ordering — can be executed and checked on a machine without a GPU, with one THREAD per
// rank. Nothing here is part of the product: it is built only by tests/hostsim/build.py into
// tests/hostsim/_build/, never by `make`/`build()`, and libelb200.so does not contain or load it.
// The product library has no CPU path.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <functional>

namespace sim {

// Append `fn` to stream `s` (nullptr = the default stream). Streams are FIFO worker threads: work
// on one stream runs in order, work on different streams runs concurrently, exactly the ordering
// guarantees CUDA gives — so a missing event dependency shows up as a wrong result or a poisoned
// read instead of being hidden by accidental serialisation.
void enqueue(cudaStream_t s, std::function<void()> fn);

// Abort with a message unless [p, p + bytes) lies inside one live "device" allocation.
void check_device_range(const void* p, size_t bytes, const char* what);

[[noreturn]] void fail(const char* fmt, ...);

// seconds a blocked wait may last before it is reported as a deadlock (HOSTSIM_TIMEOUT, default 120)
double wait_timeout_seconds();

}  // namespace sim
