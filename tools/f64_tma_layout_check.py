"""Design check for the TMA-fed fp64 GEMM — no GPU needed.

TMA lands dense boxes with the 128-byte swizzle (16-byte chunk index ^= row index & 7 inside every
8-row x 128-byte atom), so the padded pitches that make the current kernel's fragment loads
conflict-free are gone.  This script models shared-memory bank conflicts of the m8n8k4 DMMA fragment
loads (8-byte LDS, served per half-warp: 16 lanes x 8 B = 128 B = all 32 banks once) and searches the
lane -> (row) assignment that restores one wavefront per half-warp:
  * MN-contiguous tile  [k row][16 doubles of mn]  (op = 'N' for A, 'T' for B)
  * K-contiguous tile   [mn row][16 doubles of k]  (op = 'T' for A, 'N' for B)
A sum over k does not care about the order of k, and the rows of a fragment may be permuted as long as
the epilogue uses the same permutation, so both are free parameters.
"""
import itertools


def phys(row, byte_in_row):
    """byte address inside one box (rows of 128 B) after the 128-byte swizzle"""
    chunk = (byte_in_row >> 4) ^ (row & 7)
    return row * 128 + (chunk << 4) + (byte_in_row & 15)


def wavefronts(addrs):
    """number of shared-memory wavefronts for one half-warp of 8-byte loads: max multiplicity of a bank
    among DISTINCT 4-byte words (same word = broadcast)"""
    per_bank = {}
    for a in addrs:
        for w in (a // 4, a // 4 + 1):
            per_bank.setdefault(w % 32, set()).add(w)
    return max(len(v) for v in per_bank.values())


def mn_contig(krows, mm0=0):
    """lane (lr, lk) loads element (k = krows[lk], mn = mm0 + lr): worst half-warp"""
    worst = 0
    for half in (0, 1):
        addrs = [phys(krows[lk], 8 * (mm0 + lr)) for lr in range(4 * half, 4 * half + 4) for lk in range(4)]
        worst = max(worst, wavefronts(addrs))
    return worst


def k_contig(rowperm, k40=0):
    """lane (lr, lk) loads element (mn row = rowperm[lr], k = k40 + lk): worst half-warp"""
    worst = 0
    for half in (0, 1):
        addrs = [phys(rowperm[lr], 8 * (k40 + lk)) for lr in range(4 * half, 4 * half + 4) for lk in range(4)]
        worst = max(worst, wavefronts(addrs))
    return worst


def wavefronts128(addrs):
    """16-byte loads are served per quarter-warp (8 lanes x 16 B = 128 B)"""
    per_bank = {}
    for a in addrs:
        for w in range(a // 4, a // 4 + 4):
            per_bank.setdefault(w % 32, set()).add(w)
    return max(len(v) for v in per_bank.values())


def k_contig_pairs(rowperm, kpair0=0):
    """K-contiguous tile read with ONE 16-byte load per lane: lane (lr, lk) loads (k = kpair0 + 2 lk, + 1) of mn
    row rowperm[lr]; .x feeds the even k4 step, .y the odd one (same k order as the MN-contiguous {0,2,4,6} /
    {1,3,5,7} scheme, so mixed-majorness products (NN, TT) can use both). Worst quarter-warp."""
    worst = 0
    for quarter in range(4):
        addrs = [phys(rowperm[lr], 8 * (kpair0 + 2 * lk)) for lr in range(2 * quarter, 2 * quarter + 2) for lk in range(4)]
        worst = max(worst, wavefronts128(addrs))
    return worst


def main():
    print("MN-contiguous tile, natural k order {0,1,2,3}:", mn_contig([0, 1, 2, 3]), "wavefront(s) per half-warp")
    best = [(mn_contig(list(c)), c) for c in itertools.combinations(range(8), 4)]
    ok = [c for w, c in best if w == 1]
    print("  k-row sets of an 8-row atom with 1 wavefront:", ok)
    pairs = [(a, b) for a in ok for b in ok if set(a) | set(b) == set(range(8)) and a < b]
    print("  complementary pairs (two k4 steps cover the atom):", pairs)
    for mm0 in (0, 4, 8, 12):
        assert all(mn_contig(list(a), mm0) == 1 and mn_contig(list(b), mm0) == 1 for a, b in pairs), mm0
    print("K-contiguous tile, natural row order:", k_contig(list(range(8))), "wavefront(s) per half-warp")
    good = [p for p in itertools.permutations(range(8)) if k_contig(list(p)) == 1]
    print("  row permutations with 1 wavefront:", len(good), "e.g.", good[0])
    for k40 in (0, 4, 8, 12):
        assert k_contig(list(good[0]), k40) == 1
    print("K-contiguous tile, one 16-byte load per lane (k pair), natural rows:", k_contig_pairs(list(range(8))), "wavefront(s) per quarter-warp")
    good128 = [p for p in itertools.permutations(range(8)) if all(k_contig_pairs(list(p), kp) == 1 for kp in (0, 8))]
    print("  row permutations with 1 wavefront:", len(good128), "e.g.", good128[0] if good128 else None)
    print("  => one k order for every orientation: k(step s, lane lk) = 8 (s / 2) + 2 lk + (s & 1);")
    print("     MN-contiguous operands read row k directly, K-contiguous operands read the pair once per two steps")
    print("padded layout of the current kernel (pitch 132 doubles, MN-contiguous):",
          max(wavefronts([(kk * 132 + mm) * 8 for mm in range(4 * h, 4 * h + 4) for kk in range(4)]) for h in (0, 1)), "wavefront(s)")


if __name__ == "__main__":
    main()
