// el_gemm_driver — counterpart of the reference's tests/blas_like/Gemm.cpp written against the
// El.hpp facade of this repo: for every orientation pair and every SUMMA variant it times
// El::Gemm on [MC,MR] DistMatrices (2mnk / t GFlop/s, Gemm.cpp:120) and runs the same
// associativity check, (alpha op(A) op(B) + beta C) X == alpha op(A) (op(B) X) + beta C X
// (Gemm.cpp:15-49), asserting the residual instead of only printing it.
//
//   el_gemm_driver [--m M] [--n N] [--k K] [--nb NB] [--alpha a] [--beta b] [--f64] [--reps R]
// Multi-process (one per GPU): set RANK, WORLD_SIZE, LOCAL_RANK and ELB_ID_FILE (a path all ranks
// can reach; rank 0 writes the NCCL id there), optionally --gridHeight r.
#include <El.hpp>

#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <string>

using namespace El;

static int env_int(const char* name, int dflt) {
    const char* v = std::getenv(name);
    return v ? std::atoi(v) : dflt;
}

static std::unique_ptr<Grid> make_grid(int gridHeight) {
    const int rank = env_int("RANK", 0), size = env_int("WORLD_SIZE", 1);
    cudaSetDevice(env_int("LOCAL_RANK", rank));
    if (size == 1) return std::unique_ptr<Grid>(new Grid());
    const char* path = std::getenv("ELB_ID_FILE");
    if (!path) throw std::runtime_error("WORLD_SIZE > 1 needs ELB_ID_FILE for the NCCL id rendezvous");
    Grid::Id id(ELB200_UNIQUE_ID_BYTES);
    if (rank == 0) {
        id = Grid::UniqueId();
        const std::string tmp = std::string(path) + ".tmp";
        std::ofstream(tmp, std::ios::binary).write(reinterpret_cast<const char*>(id.data()), id.size());
        std::rename(tmp.c_str(), path);  // atomic publish
    } else {
        for (int tries = 0;; ++tries) {
            std::ifstream f(path, std::ios::binary);
            if (f && f.read(reinterpret_cast<char*>(id.data()), id.size())) break;
            if (tries > 600) throw std::runtime_error("timed out waiting for the NCCL id file");
            usleep(100000);
        }
    }
    return std::unique_ptr<Grid>(new Grid(id, rank, size, gridHeight));
}

template <typename T>
static double associativity_residual(Orientation oa, Orientation ob, T alpha, const DistMatrix<T>& A,
                                     const DistMatrix<T>& B, T beta, const DistMatrix<T>& COrig,
                                     const DistMatrix<T>& CFinal) {
    const Grid& g = A.Grid();
    const Int n = COrig.Width(), numRHS = 100;
    DistMatrix<T> X(g), Y(g), Z(g);
    Uniform(X, n, numRHS, T(0), T(1), 77);
    Gemm(ob, NORMAL, T(1), B, X, Z);
    Gemm(oa, NORMAL, alpha, A, Z, Y);
    Gemm(NORMAL, NORMAL, beta, COrig, X, T(1), Y);
    const double ynorm = FrobeniusNorm(Y);
    Gemm(NORMAL, NORMAL, T(-1), CFinal, X, T(1), Y);
    return FrobeniusNorm(Y) / ynorm;
}

template <typename T>
static int run(const Grid& g, Int m, Int n, Int k, T alpha, T beta, int reps) {
    const bool root = g.Rank() == 0;
    const char* tname = sizeof(T) == 4 ? "float" : "double";
    // the reference prints the residual without a bound; here it must stay below a few ulps * sqrt(k)
    const double tol = (sizeof(T) == 4 ? 1e-5 : 1e-13) * std::sqrt(double(k)) * 4;
    int failures = 0;
    const Orientation os[2] = {NORMAL, TRANSPOSE};
    const struct { GemmAlgorithm alg; const char* name; } algs[] = {
        {GEMM_SUMMA_A, "Stationary A"}, {GEMM_SUMMA_B, "Stationary B"}, {GEMM_SUMMA_C, "Stationary C"},
        {GEMM_SUMMA_DOT, "Dot"}, {GEMM_DEFAULT, "Default"}};
    for (Orientation oa : os)
        for (Orientation ob : os) {
            DistMatrix<T> A(g), B(g), COrig(g), C(g);
            Uniform(A, oa == NORMAL ? m : k, oa == NORMAL ? k : m, T(0), T(1), 1);
            Uniform(B, ob == NORMAL ? k : n, ob == NORMAL ? n : k, T(0), T(1), 2);
            Uniform(COrig, m, n, T(0), T(1), 3);
            if (root) std::printf("Testing %s  orientA=%d orientB=%d  m=%d n=%d k=%d nb=%d grid=%dx%d\n", tname, int(oa),
                                  int(ob), m, n, k, Blocksize(), g.Height(), g.Width());
            for (const auto& a : algs) {
                double best = 1e30;
                for (int it = 0; it < reps + 1; ++it) {  // first iteration is the warm-up
                    C = COrig;
                    g.Synchronize();
                    const auto t0 = std::chrono::steady_clock::now();
                    Gemm(oa, ob, alpha, A, B, beta, C, a.alg);
                    g.Synchronize();
                    const double t = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                    if (it > 0 && t < best) best = t;
                }
                const double res = associativity_residual(oa, ob, alpha, A, B, beta, COrig, C);
                const bool ok = res <= tol;
                failures += ok ? 0 : 1;
                if (root)
                    std::printf("  %-13s %9.3f ms  %10.1f GFlop/s   || E ||_F / || Y ||_F = %.3e  %s\n", a.name, best * 1e3,
                                2.0 * m * n * k / best / 1e9, res, ok ? "ok" : "FAIL");
            }
        }
    return failures;
}

int main(int argc, char** argv) {
    Int m = 2000, n = 2000, k = 2000, nb = 96;  // the reference driver's defaults (Gemm.cpp:245-250)
    double alpha = 2, beta = 3;
    int reps = 2, gridHeight = -1;
    bool f64 = false;
    for (int i = 1; i < argc; ++i) {
        auto next = [&](const char* flag) { return !std::strcmp(argv[i], flag) && i + 1 < argc; };
        if (next("--m")) m = std::atoi(argv[++i]);
        else if (next("--n")) n = std::atoi(argv[++i]);
        else if (next("--k")) k = std::atoi(argv[++i]);
        else if (next("--nb")) nb = std::atoi(argv[++i]);
        else if (next("--alpha")) alpha = std::atof(argv[++i]);
        else if (next("--beta")) beta = std::atof(argv[++i]);
        else if (next("--reps")) reps = std::atoi(argv[++i]);
        else if (next("--gridHeight")) gridHeight = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--f64")) f64 = true;
        else { std::fprintf(stderr, "unknown argument %s\n", argv[i]); return 64; }
    }
    try {
        std::unique_ptr<Grid> g = make_grid(gridHeight);
        SetBlocksize(nb);
        int failures = f64 ? run<double>(*g, m, n, k, alpha, beta, reps) : run<float>(*g, m, n, k, float(alpha), float(beta), reps);
        // error behaviour: non-conformal dimensions throw std::logic_error like El::LogicError
        try {
            DistMatrix<float> A(8, 9, *g), B(10, 7, *g), C(8, 7, *g);
            Gemm(NORMAL, NORMAL, 1.f, A, B, 0.f, C);
            ++failures;
            if (g->Rank() == 0) std::printf("nonconformal Gemm did not throw: FAIL\n");
        } catch (const std::logic_error&) {
            if (g->Rank() == 0) std::printf("nonconformal Gemm throws std::logic_error: ok\n");
        }
        if (g->Rank() == 0) std::printf(failures ? "EL_GEMM_DRIVER_FAIL %d\n" : "EL_GEMM_DRIVER_PASS\n", failures);
        return failures ? 1 : 0;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "el_gemm_driver: %s\n", e.what());
        return 2;
    }
}
