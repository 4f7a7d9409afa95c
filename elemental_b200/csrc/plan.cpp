// See plan.hpp. Pure host code: compiled into libelb200.so and exercised without a GPU by the
// CPU test-suite (tests/test_plan*.py) through the elb200_plan_* C entry points.
#include "plan.hpp"

#include <algorithm>

namespace elb {

int dist_stride(const GridShape& g, Dist d) {
    switch (d) {
        case MC: return g.r;
        case MR: return g.c;
        case VC:
        case VR: return g.r * g.c;
        default: return 1;
    }
}

int dist_rank(const GridShape& g, Dist d, GridCoord q) {
    switch (d) {
        case MC: return q.mc;
        case MR: return q.mr;
        case VC: return vc_of(g, q);
        case VR: return vr_of(g, q);
        default: return 0;
    }
}

// bit 0: depends on the grid row (mc); bit 1: depends on the grid column (mr)
static int constrains(Dist d) {
    switch (d) {
        case MC: return 1;
        case MR: return 2;
        case VC:
        case VR:
        case CIRC: return 3;  // ownership depends on both coordinates (one rank holds everything)
        default: return 0;
    }
}

bool valid_pair(Dist u, Dist v) {
    if (u == CIRC || v == CIRC) return u == CIRC && v == CIRC;
    if (u < MC || u > STAR || v < MC || v > STAR) return false;
    return (constrains(u) & constrains(v)) == 0;
}

int64_t local_height(const GridShape& g, const DistSpec& d, GridCoord q) {
    if (!owns_data(g, d, q)) return 0;
    const int s = dist_stride(g, d.col);
    return Length(d.height, Shift(dist_rank(g, d.col, q), d.col_align, s), s);
}
int64_t local_width(const GridShape& g, const DistSpec& d, GridCoord q) {
    if (!owns_data(g, d, q)) return 0;
    const int s = dist_stride(g, d.row);
    return Length(d.width, Shift(dist_rank(g, d.row, q), d.row_align, s), s);
}

namespace {

struct Lattice1D {
    int64_t first = 0, step = 1, count = 0;  // global indices first + t*step
};

// Global indices i in [0,n) with i = shift_a (mod stride_a) and i = shift_b (mod stride_b).
Lattice1D intersect(int64_t n, int64_t shift_a, int64_t stride_a, int64_t shift_b,
                    int64_t stride_b) {
    Lattice1D out;
    int64_t g = stride_a, h = stride_b;
    while (h) { int64_t t = g % h; g = h; h = t; }
    const int64_t lcm = stride_a / g * stride_b;
    out.step = lcm;
    if (Mod(shift_a - shift_b, g) != 0) return out;  // empty
    for (int64_t i = shift_a; i < lcm; i += stride_a) {
        if (Mod(i - shift_b, stride_b) == 0) {
            out.first = i;
            out.count = Length(n, i, lcm);
            return out;
        }
    }
    return out;
}

struct Side {
    int64_t row_shift, col_shift;
    int row_stride, col_stride;  // "row_*" = along the row INDEX i (the reference's colStride)
};

Side side_of(const GridShape& g, const DistSpec& d, GridCoord q) {
    Side s;
    s.row_stride = dist_stride(g, d.col);
    s.col_stride = dist_stride(g, d.row);
    s.row_shift = Shift(dist_rank(g, d.col, q), d.col_align, s.row_stride);
    s.col_shift = Shift(dist_rank(g, d.row, q), d.row_align, s.col_stride);
    return s;
}

// Entries owned by `src` rank in A and by `dst` rank in B, as a lattice of `of_a ? A : B`'s
// local matrix on that rank.
elb200_segment pair_lattice(const GridShape& g, const DistSpec& A, GridCoord src, const DistSpec& B,
                            GridCoord dst, bool of_a) {
    elb200_segment none{};
    if (!owns_data(g, A, src) || !owns_data(g, B, dst)) return none;  // [CIRC,CIRC]: only the root holds entries
    const Side sa = side_of(g, A, src), sb = side_of(g, B, dst);
    const Lattice1D rows = intersect(A.height, sa.row_shift, sa.row_stride, sb.row_shift, sb.row_stride);
    const Lattice1D cols = intersect(A.width, sa.col_shift, sa.col_stride, sb.col_shift, sb.col_stride);
    elb200_segment seg{};
    if (rows.count == 0 || cols.count == 0) return seg;
    const Side& me = of_a ? sa : sb;
    seg.row0 = (rows.first - me.row_shift) / me.row_stride;
    seg.row_step = rows.step / me.row_stride;
    seg.rows = rows.count;
    seg.col0 = (cols.first - me.col_shift) / me.col_stride;
    seg.col_step = cols.step / me.col_stride;
    seg.cols = cols.count;
    return seg;
}

int comm_index(const GridShape& g, Dist comm, GridCoord q) {
    return comm == MC ? q.mc : comm == MR ? q.mr : vc_of(g, q);
}
int comm_size(const GridShape& g, Dist comm) {
    return comm == MC ? g.r : comm == MR ? g.c : comm == STAR ? 1 : g.size();
}
Dist comm_for_mask(int mask) { return mask == 1 ? MC : mask == 2 ? MR : mask == 3 ? VC : STAR; }

}  // namespace

int64_t plan_send_offset(const GridShape& g, GridCoord sender, const DistSpec& A, const DistSpec& B, int dest_vc) {
    // mirrors the send loop of plan_copy: messages are laid out in increasing VC rank of the receiver
    const int free_a = 3 & ~(constrains(A.col) | constrains(A.row));
    const int sender_vc = vc_of(g, sender);
    int64_t off = 0;
    for (int v = 0; v < dest_vc; ++v) {
        if (v == sender_vc) continue;
        const GridCoord q = coord_of_vc(g, v);
        if (((free_a & 1) && sender.mc != q.mc) || ((free_a & 2) && sender.mr != q.mr)) continue;
        const elb200_segment sg = pair_lattice(g, A, sender, B, q, true);
        off += sg.rows * sg.cols;
    }
    return off;
}

CopyPlan plan_copy(const GridShape& g, GridCoord me, const DistSpec& A, const DistSpec& B) {
    CopyPlan plan;
    const int p = g.size();
    const int my_vc = vc_of(g, me);
    // A replicated entry is taken from the replica that shares the target rank's coordinates
    // in every grid dimension A does not constrain (nearest copy; keeps partial-sum layers apart).
    const int free_a = 3 & ~(constrains(A.col) | constrains(A.row));
    auto feeds = [&](GridCoord s, GridCoord q) {
        return (!(free_a & 1) || s.mc == q.mc) && (!(free_a & 2) || s.mr == q.mr);
    };

    // AllGather shapes: (U,V)->(U,*) with equal column alignment, (U,V)->(*,V) with equal row
    // alignment (reference: copy::RowAllGather / copy::ColAllGather aligned branches,
    // Copy/RowAllGather.hpp:30-76, Copy/ColAllGather.hpp:45-100).
    Dist gathered = STAR;
    if (A.col == CIRC || B.col == CIRC) {
        // Gather to / Scatter from the root (reference copy::Gather, copy::Scatter,
        // Copy/Gather.hpp, Copy/Scatter.hpp): point-to-point lattices through the general exchange
    } else if (A.col == B.col && A.col_align == B.col_align && B.row == STAR && A.row != STAR)
        gathered = A.row;
    else if (A.row == B.row && A.row_align == B.row_align && B.col == STAR && A.col != STAR)
        gathered = A.col;
    if (gathered != STAR) {
        const Dist comm = comm_for_mask(constrains(gathered));
        const int gsz = comm_size(g, comm);
        if (gsz > 1) {
            plan.kind = EX_ALLGATHER;
            plan.comm_dist = comm;
            plan.portion = std::max<int64_t>(
                1, MaxLength(A.height, dist_stride(g, A.col)) * MaxLength(A.width, dist_stride(g, A.row)));
            Message snd;
            snd.peer = my_vc;
            snd.seg.row0 = 0; snd.seg.row_step = 1; snd.seg.rows = local_height(g, A, me);
            snd.seg.col0 = 0; snd.seg.col_step = 1; snd.seg.cols = local_width(g, A, me);
            snd.seg.buf_offset = 0;
            plan.sends.push_back(snd);
            plan.send_elems = plan.portion;
            for (int k = 0; k < gsz; ++k) {
                GridCoord s = comm == MC ? GridCoord{k, me.mr}
                            : comm == MR ? GridCoord{me.mc, k} : coord_of_vc(g, k);
                Message m;
                m.peer = vc_of(g, s);
                m.seg = pair_lattice(g, A, s, B, me, false);
                m.seg.buf_offset = int64_t(k) * plan.portion;
                plan.recvs.push_back(m);
            }
            plan.recv_elems = int64_t(gsz) * plan.portion;
            return plan;
        }
    }

    for (int v = 0; v < p; ++v) {
        const GridCoord q = coord_of_vc(g, v);
        if (feeds(me, q)) {
            Message m;
            m.peer = v;
            m.seg = pair_lattice(g, A, me, B, q, true);
            if (m.count() > 0) {
                if (v == my_vc) { plan.has_self = true; plan.self_send = m; }
                else { m.seg.buf_offset = plan.send_elems; plan.send_elems += m.count(); plan.sends.push_back(m); }
            }
        }
        if (feeds(q, me)) {
            Message m;
            m.peer = v;
            m.seg = pair_lattice(g, A, q, B, me, false);
            if (m.count() > 0) {
                if (v == my_vc) plan.self_recv = m;
                else { m.seg.buf_offset = plan.recv_elems; plan.recv_elems += m.count(); plan.recvs.push_back(m); }
            }
        }
    }
    // EX_LOCAL must be agreed on by every rank (it decides whether a collective call is made):
    // it holds exactly when no pair of DISTINCT ranks exchanges anything, which each rank can
    // decide from the specs alone.
    bool any_remote = false;
    for (int vs = 0; vs < p && !any_remote; ++vs)
        for (int vq = 0; vq < p && !any_remote; ++vq) {
            if (vs == vq) continue;
            const GridCoord s = coord_of_vc(g, vs), q = coord_of_vc(g, vq);
            if (!feeds(s, q)) continue;
            const elb200_segment sg = pair_lattice(g, A, s, B, q, true);
            if (sg.rows * sg.cols > 0) any_remote = true;
        }
    plan.kind = any_remote ? EX_SENDRECV : EX_LOCAL;
    return plan;
}

ContractPlan plan_contract(const GridShape& g, GridCoord me, const DistSpec& A, const DistSpec& B) {
    ContractPlan plan;
    if (A.height != B.height || A.width != B.width) return plan;
    if (A.col == CIRC || B.col == CIRC) return plan;  // no partial-sum form of [CIRC,CIRC]
    const bool col_same = A.col == B.col, row_same = A.row == B.row;
    const bool col_collect = !col_same && A.col == STAR, row_collect = !row_same && A.row == STAR;
    if (!(col_same || col_collect) || !(row_same || row_collect)) return plan;
    if (col_same && A.col_align != B.col_align) return plan;
    if (row_same && A.row_align != B.row_align) return plan;
    int mask = 0;
    if (col_collect) mask |= constrains(B.col);
    if (row_collect) mask |= constrains(B.row);
    plan.valid = true;
    plan.comm_dist = comm_for_mask(mask);
    plan.group = comm_size(g, plan.comm_dist);
    plan.my_index = plan.comm_dist == STAR ? 0 : comm_index(g, plan.comm_dist, me);
    const int64_t hA = local_height(g, A, me), wA = local_width(g, A, me);
    const int su = dist_stride(g, B.col), sv = dist_stride(g, B.row);
    const int64_t slot_h = col_collect ? MaxLength(A.height, su) : hA;
    const int64_t slot_w = row_collect ? MaxLength(A.width, sv) : wA;
    plan.portion = std::max<int64_t>(1, slot_h * slot_w);
    for (int k = 0; k < plan.group; ++k) {
        GridCoord q = me;
        if (plan.comm_dist == MC) q.mc = k;
        else if (plan.comm_dist == MR) q.mr = k;
        else if (plan.comm_dist == VC) q = coord_of_vc(g, k);
        Message m;
        m.peer = vc_of(g, q);
        elb200_segment& s = m.seg;
        if (col_collect) {
            s.row0 = Shift(dist_rank(g, B.col, q), B.col_align, su);
            s.row_step = su;
            s.rows = Length(A.height, s.row0, su);
        } else { s.row0 = 0; s.row_step = 1; s.rows = hA; }
        if (row_collect) {
            s.col0 = Shift(dist_rank(g, B.row, q), B.row_align, sv);
            s.col_step = sv;
            s.cols = Length(A.width, s.col0, sv);
        } else { s.col0 = 0; s.col_step = 1; s.cols = wA; }
        s.buf_offset = int64_t(k) * plan.portion;
        plan.packs.push_back(m);
    }
    plan.unpack.row0 = 0; plan.unpack.row_step = 1; plan.unpack.rows = local_height(g, B, me);
    plan.unpack.col0 = 0; plan.unpack.col_step = 1; plan.unpack.cols = local_width(g, B, me);
    plan.unpack.buf_offset = 0;
    return plan;
}

DirectPanelPlan plan_direct_panel(const GridShape& g, GridCoord me, int64_t kb, int64_t a_row_align,
                                  int64_t b_col_align) {
    DirectPanelPlan p;
    const int64_t r = g.r, c = g.c;
    int64_t gcd = r, t = c;
    while (t) { const int64_t x = gcd % t; gcd = t; t = x; }
    const int64_t P = r / gcd, L = P * c;  // classes per source, lcm(r, c)
    p.row_off.assign(size_t(r) + 1, 0);
    for (int64_t q = 0; q < r; ++q) p.row_off[size_t(q) + 1] = p.row_off[size_t(q)] + Length(kb, Shift(q, b_col_align, r), r);
    p.b_row0 = p.row_off[size_t(me.mc)];
    p.b_rows = p.row_off[size_t(me.mc) + 1] - p.b_row0;
    const int64_t shA = Shift(me.mr, a_row_align, c), cntA = Length(kb, shA, c);
    for (int64_t u = 0; u < P && u < cntA; ++u) {
        const int64_t t0 = shA + u * c;                       // panel-relative k of local column u
        const int64_t q = Mod(t0 + b_col_align, r);           // its owner among the B rows
        const int64_t i0 = (t0 - Shift(q, b_col_align, r)) / r;
        PanelCols pc;
        pc.dst_col0 = p.row_off[size_t(q)] + i0;
        pc.dst_step = L / r;
        pc.src_col0 = u;
        pc.src_step = P;
        pc.count = (cntA - u + P - 1) / P;
        p.a_cols.push_back(pc);
    }
    return p;
}

}  // namespace elb
