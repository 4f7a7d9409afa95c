// Fused small-message collectives over NVLink peer memory: ONE kernel per rank does a whole collective —
//   wait for the members' acks of the slot's previous use, move the data (peer stores / peer loads), publish an arrival
//   flag, wait for the other members' flags, put the result into the destination matrix, acknowledge —
// where the copy-engine forms of peer.cpp need 20-40 stream operations (flag waits, forked copies, 4-byte flag writes,
// joins, pack / unpack launches): 57-122 us per call at 1 MiB on 2 x 4 B200s, all of it launch / memop latency.
//   elb200_peer_allgather        PUSH: my block into every member's slot, then unpack my slot. Replaces pack + mpi::AllGather
//                                + unpack of copy::RowAllGather / ColAllGather (include/El/blas_like/level1/Copy/RowAllGather.hpp:53-75,
//                                ColAllGather.hpp; src/core/imports/mpi/AllGather.hpp:8-21).
//   elb200_peer_exchange         PULL: pack my messages into MY slot, then read what is addressed to me out of the members'
//                                slots straight into the destination. Replaces pack + mpi::AllToAll / SendRecv + unpack of
//                                copy::ColAllToAllDemote / Promote / Exchange (Copy/ColAllToAllDemote.hpp:17-80, Copy/Exchange.hpp).
//   elb200_peer_reduce_scatter   PULL: pack my partial sums into MY slot, then sum my portion over the members' slots in
//                                communicator order and axpy it into the destination. Replaces zero-fill + pack +
//                                mpi::ReduceScatter + InterleaveMatrixUpdate of AxpyContract (level1/AxpyContract.hpp:318-337).
// The first two are pure data movement: one instantiation per element SIZE (2, 4, 8 bytes) serves half, float and double.
//
// Protocol (same flags, same sequence numbers as PeerArena's copy-engine forms, so all forms alternate freely on a channel):
//   ack_in[i] == seq_prev  =>  member i has finished reading the previous contents of the slot generation;
//   arrive_in[i] == seq    =>  member i's data of this operation is in place (PUSH: in my slot; PULL: in its slot);
//   arrive_in[my_index]    =>  the same for my OWN CTAs' stores (a CTA reads what another CTA of this grid wrote).
// Flags are written after a __threadfence_system() by the LAST CTA to finish the preceding phase (device-wide counters) and
// read with volatile loads; slot data is read with ld.global.cv (never a stale line for memory another GPU writes).
// One CTA per SM at most, so every CTA of the grid is resident when they meet at the counters; every spin has a timeout
// (a protocol bug must trap, not hang the GPU).
#include <cuda_runtime.h>

#include <cstdint>

#include "elb200_internal.h"

namespace elb {
int check_cuda(cudaError_t e, const char* what);
namespace {

constexpr unsigned long long kSpinTimeoutNs = 60000000000ull;  // 60 s: a slow peer (host-side jitter between collectives) is not a protocol bug

__device__ __forceinline__ unsigned long long now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void spin_until(const volatile uint32_t* flag, uint32_t want) {
    if (*flag == want) return;
    const unsigned long long t0 = now_ns();
    unsigned spins = 0;
    while (*flag != want) {
        if ((++spins & 0xFFFu) == 0 && now_ns() - t0 > kSpinTimeoutNs) __trap();
    }
}

struct SegTableP {
    elb200_segment seg[ELB200_MAX_SEGMENTS];
    int n;
};
struct MemberTable {
    int m[ELB200_MAX_SEGMENTS];
};

// ---- the protocol, phase by phase (every thread of every CTA calls these) -------------------------------------------
__device__ __forceinline__ void wait_previous_acks(const elb200_peer_gather& pg) {
    const int tid = threadIdx.x;
    if (tid < pg.nmembers && tid != pg.my_index) spin_until(pg.ack_in[tid], pg.seq_prev);
    __syncthreads();
}
// my stores are done: the last CTA of the grid raises the arrival flags (the members', and my own for the other CTAs of
// this grid); then everybody waits until every member's data is in place
__device__ __forceinline__ void publish_and_wait(const elb200_peer_gather& pg) {
    const int tid = threadIdx.x;
    __shared__ unsigned last;
    __threadfence_system();
    __syncthreads();
    if (tid == 0) last = atomicAdd(&pg.counters[0], 1u) == gridDim.x - 1 ? 1u : 0u;
    __syncthreads();
    if (last) {
        __threadfence_system();
        if (tid < pg.nmembers) {
            uint32_t* f = tid == pg.my_index ? pg.arrive_in[tid] : pg.arrive_out[tid];
            *reinterpret_cast<volatile uint32_t*>(f) = pg.seq;
        }
    }
    if (tid < pg.nmembers) spin_until(pg.arrive_in[tid], pg.seq);
    __syncthreads();
    __threadfence_system();
}
// my reads are done: the last CTA acknowledges (the slots may be overwritten by the next use) and re-arms the counters
__device__ __forceinline__ void acknowledge(const elb200_peer_gather& pg) {
    const int tid = threadIdx.x;
    __shared__ unsigned last_r;
    __syncthreads();
    if (tid == 0) last_r = atomicAdd(&pg.counters[1], 1u) == gridDim.x - 1 ? 1u : 0u;
    __syncthreads();
    if (last_r) {
        if (tid < pg.nmembers && tid != pg.my_index) *reinterpret_cast<volatile uint32_t*>(pg.ack_out[tid]) = pg.seq;
        if (tid == 0) { pg.counters[0] = 0; pg.counters[1] = 0; __threadfence(); }
    }
}

// ---- lattice moves: M(row0 + i*rs, col0 + j*cs) <-> buf[buf_offset + i + j*rows] -----------------------------------
// A work item is a run of kRun consecutive rows of one column (contiguous in the buffer and, for row_step == 1, in the
// matrix: 16-byte accesses when both sides are aligned); items of all segments are dealt round-robin to the CTAs, so a
// single long column and a wide short block both spread over the grid, with one division per item, none per element.
constexpr int kRun = 4096;
enum { TO_BUF = 0, FROM_BUF = 1 };
template <typename E, int DIR>
__device__ __forceinline__ void lattice_items(const elb200_segment& sg, int64_t& item_base, E* M, int64_t ldm, E* buf) {
    constexpr int V = 16 / int(sizeof(E));
    const int tid = threadIdx.x;
    const bool vec = sg.row_step == 1 && sg.rows % V == 0 && sg.row0 % V == 0 && sg.buf_offset % V == 0 && ldm % V == 0 &&
                     (reinterpret_cast<uintptr_t>(M) & 15u) == 0 && (reinterpret_cast<uintptr_t>(buf) & 15u) == 0;
    const int64_t runs = (sg.rows + kRun - 1) / kRun, items = runs * sg.cols, G = gridDim.x;
    for (int64_t it = (int64_t(blockIdx.x) - item_base % G + G) % G; it < items; it += G) {
        const int64_t j = it / runs, i0 = (it - j * runs) * kRun;
        const int64_t n = sg.rows - i0 < kRun ? sg.rows - i0 : kRun;
        E* mcol = M + sg.row0 + i0 * sg.row_step + (sg.col0 + j * sg.col_step) * ldm;
        E* bcol = buf + sg.buf_offset + i0 + j * sg.rows;
        if (vec) {  // kRun and rows are multiples of V: so are i0 and n
            for (int64_t i = tid; i < n / V; i += blockDim.x) {
                if constexpr (DIR == TO_BUF) reinterpret_cast<uint4*>(bcol)[i] = reinterpret_cast<const uint4*>(mcol)[i];
                else reinterpret_cast<uint4*>(mcol)[i] = __ldcv(reinterpret_cast<const uint4*>(bcol) + i);
            }
        } else {
            for (int64_t i = tid; i < n; i += blockDim.x) {
                if constexpr (DIR == TO_BUF) bcol[i] = mcol[i * sg.row_step];
                else mcol[i * sg.row_step] = __ldcv(bcol + i);
            }
        }
    }
    item_base += items;
}

// ---- PUSH all-gather -------------------------------------------------------------------------------------------------
template <typename E>  // E: unsigned integer of the element size
__global__ void __launch_bounds__(256) fused_allgather_kernel(const __grid_constant__ elb200_peer_gather pg, const E* __restrict__ send,
                                                              int64_t send_elems, const __grid_constant__ SegTableP tab, E* M, int64_t ldm) {
    const int g = pg.nmembers, me = pg.my_index;
    const int tid = threadIdx.x;
    wait_previous_acks(pg);  // nobody still reads my previous block in their slot
    // my block into every member's slot (mine included), 16 bytes at a time where alignment allows
    {
        constexpr int V = 16 / int(sizeof(E));
        const bool vec = (reinterpret_cast<uintptr_t>(send) & 15u) == 0 && (size_t(pg.portion) * sizeof(E)) % 16 == 0;
        const int64_t gtid = int64_t(blockIdx.x) * blockDim.x + tid, gsz = int64_t(gridDim.x) * blockDim.x;
        const int64_t nv = vec ? send_elems / V : 0;
        const uint4* sv = reinterpret_cast<const uint4*>(send);
        for (int64_t e = gtid; e < nv; e += gsz) {
            const uint4 v = sv[e];
            for (int i = 0; i < g; ++i) reinterpret_cast<uint4*>(static_cast<E*>(pg.slots[i]) + size_t(me) * pg.portion)[e] = v;
        }
        for (int64_t e = nv * V + gtid; e < send_elems; e += gsz) {
            const E v = send[e];
            for (int i = 0; i < g; ++i) (static_cast<E*>(pg.slots[i]) + size_t(me) * pg.portion)[e] = v;
        }
    }
    publish_and_wait(pg);
    // unpack the gathered slot into the destination matrix
    {
        E* buf = static_cast<E*>(pg.slots[me]);
        int64_t base = 0;
        for (int q = 0; q < tab.n; ++q) lattice_items<E, FROM_BUF>(tab.seg[q], base, M, ldm, buf);
    }
    acknowledge(pg);
}

// ---- PULL exchange ---------------------------------------------------------------------------------------------------
// sends: PACK segments into my slot (buf_offset from MY slot base). recvs[q]: UNPACK segment whose buf_offset counts from the
// slot base of member from.m[q] — where that member packed its message for me.
template <typename E>
__global__ void __launch_bounds__(256) fused_exchange_kernel(const __grid_constant__ elb200_peer_gather pg, const __grid_constant__ SegTableP sends,
                                                             const E* A, int64_t lda, const __grid_constant__ SegTableP recvs,
                                                             const __grid_constant__ MemberTable from, E* B, int64_t ldb) {
    wait_previous_acks(pg);  // nobody still reads the previous contents of my slot
    {
        E* mine = static_cast<E*>(pg.slots[pg.my_index]);
        int64_t base = 0;
        for (int q = 0; q < sends.n; ++q) lattice_items<E, TO_BUF>(sends.seg[q], base, const_cast<E*>(A), lda, mine);
    }
    publish_and_wait(pg);
    {
        int64_t base = 0;
        for (int q = 0; q < recvs.n; ++q) lattice_items<E, FROM_BUF>(recvs.seg[q], base, B, ldb, static_cast<E*>(pg.slots[from.m[q]]));
    }
    acknowledge(pg);
}

// ---- PULL reduce-scatter + axpy --------------------------------------------------------------------------------------
// packs: PACK segments of my partial sums into my slot (member i's portion at i * pg.portion). Then
// B-lattice(un) += alpha * sum_{i in communicator order} slot_i[my_index * portion + ...]  — the same arithmetic, in the same
// order, as peer_reduce_axpy_kernel (blas1.cu), so the fused and the copy-engine forms agree bit for bit.
template <typename T> struct Vec16;
template <> struct Vec16<float> { using type = float4; static constexpr int N = 4; };
template <> struct Vec16<double> { using type = double2; static constexpr int N = 2; };
template <typename T>
__global__ void __launch_bounds__(256) fused_reduce_scatter_kernel(const __grid_constant__ elb200_peer_gather pg, const __grid_constant__ SegTableP packs,
                                                                   const T* A, int64_t lda, const elb200_segment un, T* B, int64_t ldb, T alpha) {
    using V = typename Vec16<T>::type;
    constexpr int N = Vec16<T>::N;
    const int tid = threadIdx.x, g = pg.nmembers;
    wait_previous_acks(pg);
    {
        T* mine = static_cast<T*>(pg.slots[pg.my_index]);
        int64_t base = 0;
        for (int q = 0; q < packs.n; ++q) lattice_items<T, TO_BUF>(packs.seg[q], base, const_cast<T*>(A), lda, mine);
    }
    publish_and_wait(pg);
    {
        const int64_t off = int64_t(pg.my_index) * pg.portion;
        bool vec = un.row_step == 1 && un.rows % N == 0 && un.row0 % N == 0 && off % N == 0 && ldb % N == 0 && (reinterpret_cast<uintptr_t>(B) & 15u) == 0;
        for (int s = 0; s < g; ++s) vec = vec && (reinterpret_cast<uintptr_t>(pg.slots[s]) & 15u) == 0;
        const int64_t runs = (un.rows + kRun - 1) / kRun, items = runs * un.cols;
        for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
            const int64_t j = it / runs, i0 = (it - j * runs) * kRun;
            const int64_t n = un.rows - i0 < kRun ? un.rows - i0 : kRun;
            const int64_t e0 = off + i0 + j * un.rows;
            T* mcol = B + un.row0 + i0 * un.row_step + (un.col0 + j * un.col_step) * ldb;
            if (vec) {
                for (int64_t i = tid; i < n / N; i += blockDim.x) {
                    T acc[N];
#pragma unroll
                    for (int t = 0; t < N; ++t) acc[t] = T(0);
                    for (int s = 0; s < g; ++s) {
                        const V v = __ldcv(reinterpret_cast<const V*>(static_cast<const T*>(pg.slots[s]) + e0) + i);
                        const T* vp = reinterpret_cast<const T*>(&v);
#pragma unroll
                        for (int t = 0; t < N; ++t) acc[t] += vp[t];
                    }
                    V m = reinterpret_cast<V*>(mcol)[i];
                    T* mp = reinterpret_cast<T*>(&m);
#pragma unroll
                    for (int t = 0; t < N; ++t) mp[t] = fma(alpha, acc[t], mp[t]);
                    reinterpret_cast<V*>(mcol)[i] = m;
                }
            } else {
                for (int64_t i = tid; i < n; i += blockDim.x) {
                    T sum = T(0);
                    for (int s = 0; s < g; ++s) sum += __ldcv(static_cast<const T*>(pg.slots[s]) + e0 + i);
                    mcol[i * un.row_step] = fma(alpha, sum, mcol[i * un.row_step]);
                }
            }
        }
    }
    acknowledge(pg);
}

// ---- host side -------------------------------------------------------------------------------------------------------
bool segment_ok(const elb200_segment& sg) {
    return sg.rows >= 0 && sg.cols >= 0 && sg.row_step >= 1 && sg.col_step >= 1 && sg.row0 >= 0 && sg.col0 >= 0 && sg.buf_offset >= 0;
}
// non-empty segments into a table; returns the bytes they move, -1 for a bad segment
int64_t fill_table(SegTableP& tab, int nseg, const elb200_segment* segs, int elem_bytes, MemberTable* mt = nullptr, const int* members = nullptr) {
    tab.n = 0;
    int64_t elems = 0;
    for (int q = 0; q < nseg; ++q) {
        if (!segment_ok(segs[q])) return -1;
        if (segs[q].rows == 0 || segs[q].cols == 0) continue;
        if (mt) mt->m[tab.n] = members[q];
        tab.seg[tab.n++] = segs[q];
        elems += segs[q].rows * segs[q].cols;
    }
    return elems * elem_bytes;
}
// enough CTAs to move the data, never more than one per SM: every CTA of the grid must be resident at once (they meet at
// device-wide counters), which one small CTA per SM always is
unsigned grid_for(int64_t work_bytes) {
    int64_t ctas = (work_bytes + (16 << 10) - 1) / (16 << 10);
    const int sms = device_sm_count();
    if (ctas < 1) ctas = 1;
    if (ctas > sms) ctas = sms;
    return unsigned(ctas);
}
bool descriptor_ok(const elb200_peer_gather* pg) {
    return pg && pg->nmembers >= 2 && pg->nmembers <= ELB200_MAX_SEGMENTS && pg->my_index >= 0 && pg->my_index < pg->nmembers && pg->counters;
}

template <typename E>
int launch_allgather(const elb200_peer_gather& pg, const void* send, int64_t send_elems, int nseg, const elb200_segment* segs, void* M,
                     int64_t ldm, cudaStream_t s) {
    SegTableP tab;
    const int64_t out_bytes = fill_table(tab, nseg, segs, int(sizeof(E)));
    if (out_bytes < 0) return ELB200_ERR_ARG;
    const int64_t push_bytes = send_elems * pg.nmembers * int64_t(sizeof(E));
    count_launch();
    fused_allgather_kernel<E><<<grid_for(push_bytes > out_bytes ? push_bytes : out_bytes), 256, 0, s>>>(pg, static_cast<const E*>(send), send_elems, tab,
                                                                                                      static_cast<E*>(M), ldm);
    return check_cuda(cudaGetLastError(), "fused_allgather_kernel");
}

template <typename E>
int launch_exchange(const elb200_peer_gather& pg, int nsend, const elb200_segment* sends, const void* A, int64_t lda, int nrecv,
                    const elb200_segment* recvs, const int* recv_member, void* B, int64_t ldb, cudaStream_t s) {
    SegTableP st, rt;
    MemberTable mt;
    const int64_t sb = fill_table(st, nsend, sends, int(sizeof(E)));
    const int64_t rb = fill_table(rt, nrecv, recvs, int(sizeof(E)), &mt, recv_member);
    if (sb < 0 || rb < 0) return ELB200_ERR_ARG;
    for (int q = 0; q < rt.n; ++q)
        if (mt.m[q] < 0 || mt.m[q] >= pg.nmembers) return ELB200_ERR_ARG;
    count_launch();
    fused_exchange_kernel<E><<<grid_for(sb > rb ? sb : rb), 256, 0, s>>>(pg, st, static_cast<const E*>(A), lda, rt, mt, static_cast<E*>(B), ldb);
    return check_cuda(cudaGetLastError(), "fused_exchange_kernel");
}

template <typename T>
int launch_reduce_scatter(const elb200_peer_gather* pg, int npack, const elb200_segment* packs, const T* A, int64_t lda,
                          const elb200_segment* unpack, T* B, int64_t ldb, T alpha, void* stream) {
    if (!descriptor_ok(pg) || npack < 0 || npack > ELB200_MAX_SEGMENTS || (npack > 0 && (!packs || !A)) || !unpack || !segment_ok(*unpack) ||
        pg->portion < 0 || unpack->rows * unpack->cols > pg->portion)
        return ELB200_ERR_ARG;
    if (unpack->rows * unpack->cols > 0 && !B) return ELB200_ERR_ARG;
    SegTableP pt;
    const int64_t pb = fill_table(pt, npack, packs, int(sizeof(T)));
    if (pb < 0) return ELB200_ERR_ARG;
    const int64_t rb = unpack->rows * unpack->cols * pg->nmembers * int64_t(sizeof(T));
    count_launch();
    fused_reduce_scatter_kernel<T><<<grid_for(pb > rb ? pb : rb), 256, 0, static_cast<cudaStream_t>(stream)>>>(*pg, pt, A, lda, *unpack, B, ldb, alpha);
    return check_cuda(cudaGetLastError(), "fused_reduce_scatter_kernel");
}

}  // namespace
}  // namespace elb

using namespace elb;

extern "C" int elb200_peer_allgather(const elb200_peer_gather* pg, int elem_bytes, const void* send, int64_t send_elems, int nseg,
                                     const elb200_segment* segs, void* M, int64_t ldm, void* stream) {
    if (!descriptor_ok(pg) || send_elems < 0 || send_elems > pg->portion || nseg < 0 || nseg > ELB200_MAX_SEGMENTS || (nseg > 0 && !segs))
        return ELB200_ERR_ARG;
    if ((send_elems > 0 && !send) || (nseg > 0 && !M)) return ELB200_ERR_ARG;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (elem_bytes) {
        case 2: return launch_allgather<uint16_t>(*pg, send, send_elems, nseg, segs, M, ldm, s);
        case 4: return launch_allgather<uint32_t>(*pg, send, send_elems, nseg, segs, M, ldm, s);
        case 8: return launch_allgather<unsigned long long>(*pg, send, send_elems, nseg, segs, M, ldm, s);
        default: return ELB200_ERR_ARG;
    }
}

extern "C" int elb200_peer_exchange(const elb200_peer_gather* pg, int elem_bytes, int nsend, const elb200_segment* sends, const void* A,
                                    int64_t lda, int nrecv, const elb200_segment* recvs, const int* recv_member, void* B, int64_t ldb,
                                    void* stream) {
    if (!descriptor_ok(pg) || nsend < 0 || nsend > ELB200_MAX_SEGMENTS || nrecv < 0 || nrecv > ELB200_MAX_SEGMENTS || (nsend > 0 && (!sends || !A)) ||
        (nrecv > 0 && (!recvs || !recv_member || !B)))
        return ELB200_ERR_ARG;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (elem_bytes) {
        case 2: return launch_exchange<uint16_t>(*pg, nsend, sends, A, lda, nrecv, recvs, recv_member, B, ldb, s);
        case 4: return launch_exchange<uint32_t>(*pg, nsend, sends, A, lda, nrecv, recvs, recv_member, B, ldb, s);
        case 8: return launch_exchange<unsigned long long>(*pg, nsend, sends, A, lda, nrecv, recvs, recv_member, B, ldb, s);
        default: return ELB200_ERR_ARG;
    }
}

extern "C" int elb200_peer_reduce_scatter_f32(const elb200_peer_gather* pg, int npack, const elb200_segment* packs, const float* A, int64_t lda,
                                              const elb200_segment* unpack, float* B, int64_t ldb, float alpha, void* stream) {
    return launch_reduce_scatter<float>(pg, npack, packs, A, lda, unpack, B, ldb, alpha, stream);
}
extern "C" int elb200_peer_reduce_scatter_f64(const elb200_peer_gather* pg, int npack, const elb200_segment* packs, const double* A, int64_t lda,
                                              const elb200_segment* unpack, double* B, int64_t ldb, double alpha, void* stream) {
    return launch_reduce_scatter<double>(pg, npack, packs, A, lda, unpack, B, ldb, alpha, stream);
}
