"""TEST INFRASTRUCTURE ONLY. Regenerates tests/golden/ from the REFERENCE ITSELF.

Runs oracle/_ref/gemm_ref (the reference's own CPU El::Gemm, compiled by oracle/ref_build/Makefile
from /root/reference where it lies; single rank) with --golden, and packs what it wrote into

  tests/golden/ref_gemm_1x1.npz   per case: orientations, GemmAlgorithm, m n k, blocksize, alpha,
                                  beta, the inputs A B C0 and the reference's C (f32 and f64)
  tests/golden/ref_indexing.npz   tables of El::Shift / El::Length / El::MaxLength
                                  (include/El/core/indexing/impl.hpp) and Grid::DefaultHeight
                                  (src/core/Grid.cpp) as the reference evaluates them

Usage (in the build container, where /root/reference exists):
    make -C oracle/ref_build && python oracle/make_golden.py
The fixtures travel to the GPU box; the reference tree and this script's inputs do not need to.
"""
from __future__ import annotations

import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "gemm_ref")
OUT = os.path.join(ROOT, "tests", "golden")


def lcg_matrix(h: int, w: int, seed: int, dtype) -> np.ndarray:
    """The input generator of oracle/ref_build/gemm_ref.cpp (fill_lcg), column-major fill order."""
    mask = (1 << 64) - 1
    s = (0x9E3779B97F4A7C15 * (seed + 1)) & mask
    out = np.empty((h, w), dtype=np.float64)
    for j in range(w):
        for i in range(h):
            s = (s * 6364136223846793005 + 1442695040888963407) & mask
            out[i, j] = 2.0 * (((s >> 11) & ((1 << 53) - 1)) / float(1 << 53)) - 1.0
    return out.astype(dtype)


def main() -> int:
    if not os.path.exists(REF_BIN):
        print("oracle/_ref/gemm_ref is missing: run `make -C oracle/ref_build` first", file=sys.stderr)
        return 1
    os.makedirs(OUT, exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run([REF_BIN, "--golden", tmp], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        arrays = {}
        meta = []
        for ln in open(os.path.join(tmp, "index.txt")):
            t, cid, oa, ob, alg, m, n, k, nb, alpha, beta = ln.split()
            cid, oa, ob, alg, m, n, k, nb = map(int, (cid, oa, ob, alg, m, n, k, nb))
            dt = np.float32 if t == "f32" else np.float64
            shapes = {"A": (k, m) if oa else (m, k), "B": (n, k) if ob else (k, n), "C0": (m, n), "C": (m, n)}
            for name, (h, w) in shapes.items():
                raw = np.fromfile(os.path.join(tmp, f"{t}_{cid}_{name}.bin"), dtype=dt)
                assert raw.size == h * w, (ln, name, raw.size)
                mat = raw.reshape(w, h).T.copy()  # files are column-major
                # Inputs come from a fixed LCG (gemm_ref.cpp fill_lcg, restated as lcg_matrix below), so
                # only the reference's OUTPUT is stored, plus the inputs of case 0; every other input is
                # checked here, at generation time, to equal the restated generator bit for bit.
                if name == "C" or cid == 0:
                    arrays[f"{t}_{cid}_{name}"] = mat
                else:
                    seed = 3 * cid + {"A": 0, "B": 1, "C0": 2}[name]
                    assert np.array_equal(mat, lcg_matrix(h, w, seed, dt)), (ln, name)
            meta.append((0 if t == "f32" else 1, cid, oa, ob, alg, m, n, k, nb, float(alpha), float(beta)))
        arrays["meta"] = np.array(meta, dtype=np.float64)
        np.savez_compressed(os.path.join(OUT, "ref_gemm_1x1.npz"), **arrays)

        L, H = [], []
        for ln in open(os.path.join(tmp, "indexing.txt")):
            f = ln.split()
            (L if f[0] == "L" else H).append([int(x) for x in f[1:]])
        # L columns: n rank align stride shift length maxlength ; H columns: p height
        np.savez_compressed(os.path.join(OUT, "ref_indexing.npz"), L=np.array(L, dtype=np.int64),
                            H=np.array(H, dtype=np.int64))
    print(f"wrote {len(meta)} Gemm cases, {len(L)} index rows, {len(H)} grid heights to {OUT}")
    make_dist_golden()
    make_level3_golden()
    make_complex_golden()
    return 0


DIST_GRIDS = [(2, 1), (2, 2), (4, 2), (6, 2), (6, 3), (8, 2)]  # (processes, grid height)


def make_dist_golden() -> None:
    """The reference run MULTI-RANK (oracle/ref_build/stub/mpi_stub.cpp: one process per rank over a
    shared-memory file): per-rank local matrices after each of the 121 (U,V)->(U2,V2)
    redistributions and after 20 El::Gemm runs, on 1x2, 2x1, 2x2, 2x3, 3x2 and 2x4 grids.
    tests/golden/ref_dist.npz, per grid tag g = f"r{r}c{c}" and rank q:
      {g}_q{q}_I    [rank, height, width, row, col, vcRank, vrRank]
      {g}_q{q}_R    rows [U V U2 V2 caA raA caB raB colShiftB rowShiftB lh lw offset]   (El::Dist enum values)
      {g}_q{q}_Q    the 23 [CIRC,CIRC] cases (gather to / scatter from a root): R's columns + [rootA rootB]
      {g}_q{q}_G    rows [id oa ob alg m n k nb caC raC colShift rowShift lh lw offset]
      {g}_q{q}_bin  float64 blob the offsets index into (column-major local matrices)"""
    arrays = {}
    for (nproc, height) in DIST_GRIDS:
        tag = f"r{height}c{nproc // height}"
        with tempfile.TemporaryDirectory() as tmp:
            shm = f"/dev/shm/elb_stub_{os.getpid()}_{tag}"
            if os.path.exists(shm):
                os.remove(shm)
            procs = []
            for q in range(nproc):
                env = dict(os.environ, ELB_STUB_NP=str(nproc), ELB_STUB_RANK=str(q), ELB_STUB_SHM=shm,
                           OPENBLAS_NUM_THREADS="1")
                procs.append(subprocess.Popen([REF_BIN, "--dist-golden", tmp, str(height)], env=env,
                                              stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True))
            try:
                for q, pr in enumerate(procs):
                    _, err = pr.communicate(timeout=300)
                    assert pr.returncode == 0, f"{tag} rank {q}: {err[-2000:]}"
            finally:
                for pr in procs:
                    if pr.poll() is None:
                        pr.kill()
                if os.path.exists(shm):
                    os.remove(shm)
            for q in range(nproc):
                recs = {"I": [], "R": [], "G": [], "Q": []}
                for ln in open(os.path.join(tmp, f"rank{q}.txt")):
                    f = ln.split()
                    recs[f[0]].append([int(x) for x in f[1:]])
                assert len(recs["R"]) == 121 and len(recs["G"]) == 20 and len(recs["Q"]) == 23, (tag, q)
                arrays[f"{tag}_q{q}_I"] = np.array(recs["I"][0], dtype=np.int64)
                arrays[f"{tag}_q{q}_R"] = np.array(recs["R"], dtype=np.int64)
                arrays[f"{tag}_q{q}_G"] = np.array(recs["G"], dtype=np.int64)
                arrays[f"{tag}_q{q}_Q"] = np.array(recs["Q"], dtype=np.int64)
                arrays[f"{tag}_q{q}_bin"] = np.fromfile(os.path.join(tmp, f"rank{q}.bin"), dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "ref_dist.npz"), **arrays)
    print(f"wrote multi-rank reference runs for grids {[f'{h}x{n // h}' for n, h in DIST_GRIDS]}")


def _run_ranks(args, nproc: int, tag: str):
    """Start `nproc` processes of gemm_ref over the shared-memory MPI stub and wait for them."""
    shm = f"/dev/shm/elb_stub_{os.getpid()}_{tag}"
    if os.path.exists(shm):
        os.remove(shm)
    procs = []
    for q in range(nproc):
        env = dict(os.environ, ELB_STUB_NP=str(nproc), ELB_STUB_RANK=str(q), ELB_STUB_SHM=shm, OPENBLAS_NUM_THREADS="1")
        procs.append(subprocess.Popen([REF_BIN] + args, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True))
    try:
        for q, pr in enumerate(procs):
            _, err = pr.communicate(timeout=300)
            assert pr.returncode == 0, f"{tag} rank {q}: {err[-2000:]}"
    finally:
        for pr in procs:
            if pr.poll() is None:
                pr.kill()
        if os.path.exists(shm):
            os.remove(shm)


LEVEL3_GRIDS = [(1, 1)] + DIST_GRIDS


def make_level3_golden() -> None:
    """The reference's triangular rank-k family run multi-rank (gemm_ref --level3-golden): ScaleTrapezoid,
    MakeTrapezoidal, AxpyTrapezoid, MakeSymmetric, the four LocalTrrk overloads, Trrk, Syrk, Symm.
    tests/golden/ref_level3.npz, per grid tag g and rank q:
      {g}_q{q}_L    rows [kind uplo p1 p2 colAlign rowAlign colShift rowShift lh lw offset]
                    kind 0 ScaleTrapezoid(-0.5) 1 MakeTrapezoidal 2 AxpyTrapezoid(2.0) (p1 = offset); 3 MakeSymmetric;
                    4 LocalTrrk 5 Trrk (p1, p2 = orientation of A, B); 6 Syrk (p1 = orientation); 7 Symm (p1 = side);
                    8 Trsm (p1 = side, p2 = 2*orientation + diag; alpha 2, A = LCG + n on the diagonal, or LCG / n for UNIT)
      {g}_q{q}_bin  float64 blob (column-major local results)
    Inputs are LCG matrices (lcg_matrix) with seeds 7001 (A), 7002 (B), 7003 (C), 7004 (X); sizes and
    scalars as in oracle/ref_build/gemm_ref.cpp::level3_golden."""
    arrays = {}
    for (nproc, height) in LEVEL3_GRIDS:
        tag = f"r{height}c{nproc // height}"
        with tempfile.TemporaryDirectory() as tmp:
            _run_ranks(["--level3-golden", tmp, str(height)], nproc, "l3" + tag)
            for q in range(nproc):
                recs = [[int(x) for x in ln.split()[1:]] for ln in open(os.path.join(tmp, f"l3rank{q}.txt"))]
                assert len(recs) == 12 + 6 + 2 + 8 + 8 + 4 + 4 + 16, (tag, q, len(recs))
                arrays[f"{tag}_q{q}_L"] = np.array(recs, dtype=np.int64)
                arrays[f"{tag}_q{q}_bin"] = np.fromfile(os.path.join(tmp, f"l3rank{q}.bin"), dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "ref_level3.npz"), **arrays)
    print(f"wrote level-3 reference runs for grids {[f'{h}x{n // h}' for n, h in LEVEL3_GRIDS]}")


def lcg_complex_matrix(h: int, w: int, seed: int, dtype) -> np.ndarray:
    """fill_lcg_complex of gemm_ref.cpp: real part from stream `seed`, imaginary part from stream `seed + 1000`."""
    real = np.float32 if np.dtype(dtype) == np.complex64 else np.float64
    return (lcg_matrix(h, w, seed, real) + 1j * lcg_matrix(h, w, seed + 1000, real)).astype(dtype)


def make_complex_golden() -> None:
    """tests/golden/ref_gemm_complex_1x1.npz: the reference's own El::Gemm<Complex<float|double>> (SURVEY 8(f) N3;
    src/blas_like/level3/Gemm.cpp:504-506) on one rank — orientations NORMAL / TRANSPOSE / ADJOINT, all five
    GemmAlgorithm values, complex alpha and beta. Inputs come from the restated LCG (checked bit for bit on case 0),
    only the reference's C is stored. meta columns: type (0 c32, 1 c64), id, orientA, orientB, alg, m, n, k, nb,
    alpha re, alpha im, beta re, beta im."""
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run([REF_BIN, "--golden-complex", tmp], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        arrays, meta = {}, []
        for ln in open(os.path.join(tmp, "index_complex.txt")):
            t, cid, oa, ob, alg, m, n, k, nb, ar, ai, br, bi = ln.split()
            cid, oa, ob, alg, m, n, k, nb = map(int, (cid, oa, ob, alg, m, n, k, nb))
            dt = np.complex64 if t == "c32" else np.complex128
            shapes = {"A": (k, m) if oa else (m, k), "B": (n, k) if ob else (k, n), "C0": (m, n), "C": (m, n)}
            for name, (h, w) in shapes.items():
                if name != "C" and cid != 0:
                    continue
                raw = np.fromfile(os.path.join(tmp, f"{t}_{cid}_{name}.bin"), dtype=dt)
                assert raw.size == h * w, (ln, name, raw.size)
                mat = raw.reshape(w, h).T.copy()
                if name == "C":
                    arrays[f"{t}_{cid}_C"] = mat
                else:
                    seed = 3 * cid + {"A": 0, "B": 1, "C0": 2}[name]
                    assert np.array_equal(mat, lcg_complex_matrix(h, w, seed, dt)), (ln, name)
            meta.append((0 if t == "c32" else 1, cid, oa, ob, alg, m, n, k, nb, float(ar), float(ai), float(br), float(bi)))
        arrays["meta"] = np.array(meta, dtype=np.float64)
        np.savez_compressed(os.path.join(OUT, "ref_gemm_complex_1x1.npz"), **arrays)
    print(f"wrote {len(meta)} complex Gemm cases to {OUT}")


if __name__ == "__main__":
    if "--complex-only" in sys.argv:
        make_complex_golden()
        sys.exit(0)
    if "--level3-only" in sys.argv:
        make_level3_golden()
        sys.exit(0)
    sys.exit(main())
