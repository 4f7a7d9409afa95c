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

// ---- virtual time (HOSTSIM_VTIME=1): a timeline MODEL next to the functional simulation. Every stream carries
// a virtual clock; an operation starts when its stream, its issue on the host, the things it waits for (event
// records, flag writes, collective partners) and the resource it needs are all ready, and lasts bytes / rate or
// flops / rate with the rates measured on B200s (sim_cuda.cpp: kRate*). cudaEventElapsedTime then returns
// modelled milliseconds. HOSTSIM_DRY=1 additionally skips the payload of every operation (nothing is computed or
// copied except flags), so bench-sized schedules (n = 32768 on 2 x 4) run in seconds on a laptop.
enum Res { RES_NONE = 0, RES_TENSOR, RES_HBM, RES_H2D, RES_D2H, RES_COUNT };
bool vtime_on();
bool dry_run();
// inside a stream closure: account an operation of `seconds` on resource `res` of the stream's device
void vt_op(Res res, double seconds);
// inside a stream closure: current virtual time of the executing stream / move it forward to at least t
double vt_now();
void vt_at_least(double t);
// HOSTSIM_VTRACE=<file>: one line per modelled operation (device, stream, kind, start, duration)
void vt_trace(const char* what, double start, double seconds, double bytes);
// measured rates the operation doubles use
constexpr double kRateTensorF32 = 230e12, kRateTensorF64 = 32.5e12;   // flop/s inside El::Gemm (bench, profiles/)
constexpr double kRateHbm = 6.5e12;                                    // B/s, blas1 kernels (profiles/r01_blas1_bandwidth.json)
constexpr double kRatePcie = 55e9, kRateNvlink = 750e9, kRateLocalCopy = 2.5e12;  // B/s per direction / per copy stream
constexpr double kLaunch = 4e-6, kFlagLatency = 2e-6;
// a dry run still moves copies / collectives up to this size: flags, scalars, the arena's handle exchange
constexpr size_t kDryKeep = size_t(1) << 20;

}  // namespace sim
