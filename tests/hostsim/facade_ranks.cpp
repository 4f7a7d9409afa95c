// hostsim only: the C++ facade checks of tools/el_level3_check.cpp and tools/el_api_check.cpp on an r x c
// grid whose ranks are THREADS of this process (tests/hostsim/sim.hpp) — include/El.hpp driven the way a
// multi-rank Hydrogen program drives it (El::Grid from a shared id, collective calls in the same order on
// every rank), without GPUs.     usage: facade_ranks <gridHeight> <gridWidth>
#define main el_level3_check_main
#include "../../tools/el_level3_check.cpp"
#undef main
#define main el_comm_check_main
#include "../../tools/el_comm_check.cpp"
#undef main

#include <atomic>
#include <cstdlib>
#include <thread>
#include <vector>

int main(int argc, char** argv) {
    const int r = argc > 1 ? std::atoi(argv[1]) : 2, c = argc > 2 ? std::atoi(argv[2]) : 2, p = r * c;
    const Grid::Id id = Grid::UniqueId(), world_id = Grid::UniqueId();
    std::atomic<int> bad{0};
    std::vector<std::thread> ranks;
    for (int q = 0; q < p; ++q)
        ranks.emplace_back([&, q] {
            try {
                Grid g(id, q, p, r);
                if (g.Height() != r || g.Width() != c || g.VCRank() != q || g.Row() != q % r || g.Col() != q / r) ++bad;
                bad += exercise<float>(g);
                bad += exercise<double>(g);
                g.Synchronize();
                // the communicator side (tools/el_comm_check.cpp): a world communicator per thread-rank, Grid(comm, r),
                // the grid's communicators, mpi:: collectives over the arena and over NCCL, proxies, MultiSync
                elb200_comm_t wh = nullptr;
                if (elb200_comm_create(world_id.data(), q, p, &wh) != ELB200_OK) throw std::runtime_error(elb200_last_error());
                {
                    mpi::Comm world(wh);
                    bad += comm_check::exercise<float>(world, r);
                    bad += comm_check::exercise<double>(world, r);
                    mpi::Free(world);
                }
            } catch (const std::exception& e) {
                std::fprintf(stderr, "rank %d: %s\n", q, e.what());
                std::_Exit(5);  // the other ranks would wait for this one forever
            }
        });
    for (auto& t : ranks) t.join();
    if (bad) std::printf("FACADE_RANKS_FAIL %d\n", bad.load());
    else std::printf("FACADE_RANKS_PASS %dx%d\n", r, c);
    return bad ? 1 : 0;
}
