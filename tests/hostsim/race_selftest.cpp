// hostsim self-test of the race detection (tests/hostsim + ThreadSanitizer): a "kernel" on one stream
// writes a buffer that a "kernel" on another stream reads.  With the event dependency between the two
// streams ThreadSanitizer must stay silent; without it (argv[1] == "race") it must report a data race —
// the same mistake in the engine (a GEMM reading a panel before its gather was ordered in front of it,
// a slot handed back before its reader finished) is reported the same way when the distributed cases
// run on the HOSTSIM_SANITIZE=thread build.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>

#include "../../include/elb200.h"

int main(int argc, char** argv) {
    const bool race = argc > 1 && !std::strcmp(argv[1], "race");
    const int64_t n = 1 << 12;
    float *a = nullptr, *b = nullptr;
    cudaMalloc(reinterpret_cast<void**>(&a), n * sizeof(float));
    cudaMalloc(reinterpret_cast<void**>(&b), n * sizeof(float));
    cudaStream_t s1, s2;
    cudaStreamCreate(&s1);
    cudaStreamCreate(&s2);
    cudaEvent_t done;
    cudaEventCreateWithFlags(&done, cudaEventDisableTiming);
    for (int rep = 0; rep < 8; ++rep) {
        elb200_fill2d_f32(n, 1, float(rep), a, n, s1);          // producer on s1
        cudaEventRecord(done, s1);
        if (!race) cudaStreamWaitEvent(s2, done, 0);            // the dependency under test
        elb200_copy2d_f32(n, 1, a, 1, n, b, 1, n, s2);          // consumer on s2
        cudaEventRecord(done, s2);
        cudaStreamWaitEvent(s1, done, 0);                       // WAR: next fill waits for this copy
    }
    cudaStreamSynchronize(s1);
    cudaStreamSynchronize(s2);
    float last = -1;
    cudaMemcpy(&last, b, sizeof(float), cudaMemcpyDeviceToHost);
    std::printf("RACE_SELFTEST_DONE %s last=%g\n", race ? "race" : "ordered", last);
    cudaFree(a);
    cudaFree(b);
    return 0;
}
