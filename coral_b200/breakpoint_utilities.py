"""Drop-in for reference ``breakpoint_utilities.py``: same names, argument meaning and return shapes.

``cluster_bp_list``, ``bpc2bp``, ``alignment2bp`` and ``alignment2bp_l`` run in the CUDA library; this
module only flattens the reference's list layouts into arrays and rebuilds them.  The one-line interval
predicates (``interval_overlap`` & co., reference :11-67) are part of the host-side interface mirror: they
take and return Python scalars and are evaluated where they are called; their batched forms live inside
the kernels (``overlap_first`` in csrc/breakpoints.cu).
"""
import numpy as np

from . import global_names, gpu_ops
from .gpu_ops import cluster_bp_list, bpc2bp  # noqa: F401  (reference :252-286, :299-388)


def interval_overlap(int1, int2):
    """reference :11-15, evaluated literally (containment for reversed '-' intervals)."""
    return (int1[0] == int2[0] and int(int1[1]) <= int(int2[2]) and int(int2[1]) <= int(int1[2]))


def interval_include(int1, int2):
    return (int1[0] == int2[0] and int(int1[1]) >= int(int2[1]) and int(int1[2]) <= int(int2[2]))


def interval_adjacent(int1, int2):
    if int1[0] != int2[0]:
        return False
    first, second = (int1, int2) if int1[1] <= int2[1] else (int2, int1)
    return second[1] == first[2] + 1


def interval_overlap_l(int1, intl):
    for k, x in enumerate(intl):
        if interval_overlap(int1, x):
            return k
    return -1


def interval_include_l(int1, intl):
    for k, x in enumerate(intl):
        if interval_include(int1, x):
            return k
    return -1


def interval_exclusive(int1, intl):
    """reference :54-67: pieces of int1 not covered by intl, and the indices of the covering intervals."""
    hit = set()
    pieces = [list(int1)]
    for k, other in enumerate(intl):
        for j in range(len(pieces) - 1, -1, -1):
            cur = pieces[j]
            if interval_overlap(cur, other):
                hit.add(k)
                del pieces[j]
                if cur[1] < other[1]:
                    pieces.append([cur[0], cur[1], other[1] - 1, -1])
                if cur[2] > other[2]:
                    pieces.append([cur[0], other[2] + 1, cur[2], -1])
    return hit, pieces


def interval2bp(R1, R2, r=(), rgap=0):
    """reference :289-295 (KeyError for contigs outside global_names.chr_idx)."""
    c1, c2 = global_names.chr_idx[R1[0]], global_names.chr_idx[R2[0]]
    if c2 < c1 or (c2 == c1 and R2[1] < R1[2]):
        return [R1[0], R1[2], R1[3], R2[0], R2[1], global_names.neg_plus_minus[R2[3]], r, rgap, 0]
    return [R2[0], R2[1], global_names.neg_plus_minus[R2[3]], R1[0], R1[2], R1[3], (r[0], r[2], r[1]), rgap, 1]


def bp_match(bp1, bp2, rgap, bp_distance_cutoff):
    """reference :391-416 for one pair (any rgap, one cutoff per coordinate), evaluated by the device routine the peel
    loop uses (coral_bp_match); the chromosome / orientation comparison of :396 happens here, on the strings."""
    if not (bp1[0] == bp2[0] and bp1[3] == bp2[3] and bp1[2] == bp2[2] and bp1[5] == bp2[5]):
        return False
    ctx = gpu_ops.default_context()
    r = ctx.bp_match([int(bp1[1])], [int(bp1[4])], [0 if bp1[2] == '+' else 1 if bp1[2] == '-' else 2],
                     [0 if bp1[5] == '+' else 1 if bp1[5] == '-' else 2], [float(rgap)], [int(bp2[1])], [int(bp2[4])],
                     int(bp_distance_cutoff[0]), int(bp_distance_cutoff[1]))
    return bool(r[0])


def _chimeric_to_device(ctx, reads):
    """reads: list of (name, (qint, rint, qual[, nm])) -> uploads a chimeric table; returns the chr list."""
    chrs = {}
    ptr, qs, qe, tid, rs, re, st, mq = [0], [], [], [], [], [], [], []
    for _, ca in reads:
        for q, r, m in zip(ca[0], ca[1], ca[2]):
            qs.append(int(q[0])); qe.append(int(q[1])); tid.append(chrs.setdefault(r[0], len(chrs)))
            rs.append(int(r[1])); re.append(int(r[2])); st.append(0 if r[3] == '+' else 1); mq.append(int(m))
        ptr.append(len(qs))
    return chrs, (np.array(ptr), qs, qe, tid, rs, re, st, mq)


def _rows(cand, names, chr_names):
    out = []
    for k in range(cand["n"]):
        out.append([chr_names[cand["chr1"][k]], int(cand["pos1"][k]), '+-'[cand["o1"][k]], chr_names[cand["chr2"][k]],
                    int(cand["pos2"][k]), '+-'[cand["o2"][k]], (names[cand["read"][k]], int(cand["i"][k]), int(cand["j"][k])),
                    int(cand["rgap"][k]), int(cand["flag"][k]), int(cand["mq1"][k]), int(cand["mq2"][k])])
    return out


def _prepare(rn, chimeric_alignment, intervals, min_bp_match_cutoff, min_mapq, gap_, gap_mapq):
    ctx = gpu_ops.default_context()
    chrs, table = _chimeric_to_device(ctx, [(rn, chimeric_alignment)])
    for iv in intervals:
        chrs.setdefault(iv[0], len(chrs))
    names = sorted(chrs, key=chrs.get)
    ctx.set_references(names)
    ctx.set_params(min_bp_match_cutoff=min_bp_match_cutoff, min_mapq=min_mapq, gap=gap_, gap_mapq=gap_mapq)
    ctx.set_chimeric(*table)
    return ctx, chrs, names


def alignment2bp_l(rn, chimeric_alignment, min_bp_match_cutoff, min_mapq, gap_, intrvls, gap_mapq=10):
    """reference :129-186."""
    if len(chimeric_alignment[1]) < 2:
        return []
    ctx, chrs, names = _prepare(rn, chimeric_alignment, intrvls, min_bp_match_cutoff, min_mapq, gap_, gap_mapq)
    c = ctx.candidates_l([chrs[iv[0]] for iv in intrvls], [int(iv[1]) for iv in intrvls], [int(iv[2]) for iv in intrvls])
    return _rows(c, [rn], names)


def alignment2bp(rn, chimeric_alignment, min_bp_match_cutoff, min_mapq, intrvl1, intrvl2, gap_mapq=10):
    """reference :70-96."""
    if len(chimeric_alignment[1]) < 2:
        return []
    ctx, chrs, names = _prepare(rn, chimeric_alignment, [intrvl1, intrvl2], min_bp_match_cutoff, min_mapq, 100, gap_mapq)
    c = ctx.candidates_pair([0], [chrs[intrvl1[0]], chrs[intrvl2[0]]], [int(intrvl1[1]), int(intrvl2[1])],
                            [int(intrvl1[2]), int(intrvl2[2])])
    return _rows(c, [rn], names)


# ---- edit-distance filtered candidate forms (reference :99-126, :189-249) ------------------------------------------
# Reached only with bam_to_breakpoint_nanopore.nm_filter = True, which the reference never sets (ibg:40; ibg:1351 is
# commented out): host Python, kept so that the module offers every callable of the reference's.
def _query_gap(qint, a, b):
    return int(qint[b][0]) - int(qint[a][1])


def alignment2bp_nm(rn, chimeric_alignment, min_bp_match_cutoff, min_mapq, max_nm, intrvl1, intrvl2, gap_mapq=10):
    """reference :99-126: alignment2bp with both alignments' NM rate below max_nm."""
    qint, rint, mq, nm = chimeric_alignment[0], chimeric_alignment[1], chimeric_alignment[2], chimeric_alignment[3]
    n = len(rint)

    def usable(a, b):
        return mq[a] >= min_mapq and mq[b] >= min_mapq and nm[a] < max_nm and nm[b] < max_nm

    def across(a, b):        # one alignment on each interval, either way round
        return (interval_overlap(rint[a], intrvl1) and interval_overlap(rint[b], intrvl2)) or \
               (interval_overlap(rint[b], intrvl1) and interval_overlap(rint[a], intrvl2))

    out, taken = [], [0] * max(n - 1, 0)
    for i in range(n - 1):
        g = _query_gap(qint, i, i + 1)
        if g + min_bp_match_cutoff >= 0 and usable(i, i + 1) and across(i, i + 1):
            out.append(interval2bp(rint[i], rint[i + 1], (rn, i, i + 1), g) + [mq[i], mq[i + 1]])
            taken[i] = 1
    for i in range(1, n - 1):
        if not taken[i - 1] and not taken[i] and mq[i] < gap_mapq and usable(i - 1, i + 1) and across(i - 1, i + 1):
            out.append(interval2bp(rint[i - 1], rint[i + 1], (rn, i - 1, i + 1), _query_gap(qint, i - 1, i + 1)) + [mq[i - 1], mq[i + 1]])
    return out


def alignment2bp_nm_l(rn, chimeric_alignment, min_bp_match_cutoff, min_mapq, max_nm, gap_, intrvls, gap_mapq=10):
    """reference :189-249: alignment2bp_l with both alignments' NM rate below max_nm."""
    qint, rint, mq, nm = chimeric_alignment[0], chimeric_alignment[1], chimeric_alignment[2], chimeric_alignment[3]
    n = len(rint)

    def usable(a, b):
        return mq[a] >= min_mapq and mq[b] >= min_mapq and nm[a] < max_nm and nm[b] < max_nm

    def junction(a, b):      # both on the same amplicon interval, and either a strand switch or a reference jump unlike the query's
        io = interval_overlap_l(rint[a], intrvls)
        if io < 0 or io != interval_overlap_l(rint[b], intrvls):
            return False
        if rint[b][3] != rint[a][3]:
            return True
        if rint[b][3] == '+':
            ref_gap = int(rint[b][1]) - int(rint[a][2])
        elif rint[b][3] == '-':
            ref_gap = int(rint[a][2]) - int(rint[b][1])
        else:
            return False
        g = _query_gap(qint, a, b)
        return abs(g - ref_gap) > max(gap_, abs(g * 0.2))

    out, taken = [], [0] * max(n - 1, 0)
    for i in range(n - 1):
        g = _query_gap(qint, i, i + 1)
        if g + min_bp_match_cutoff >= 0 and usable(i, i + 1) and junction(i, i + 1):
            out.append(interval2bp(rint[i], rint[i + 1], (rn, i, i + 1), g) + [mq[i], mq[i + 1]])
            taken[i] = 1
    for i in range(1, n - 1):
        if not taken[i - 1] and not taken[i] and mq[i] < gap_mapq and usable(i - 1, i + 1) and junction(i - 1, i + 1):
            out.append(interval2bp(rint[i - 1], rint[i + 1], (rn, i - 1, i + 1), _query_gap(qint, i - 1, i + 1)) + [mq[i - 1], mq[i + 1]])
    return out


def sort_chrom_names(chromlist):
    def key(x):
        v = x[3:] if x.startswith("chr") else x
        return int(v) if v.isnumeric() else ord(v)
    return sorted(chromlist, key=key)
