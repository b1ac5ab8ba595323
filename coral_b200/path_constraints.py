"""Drop-in for reference ``path_constraints.py`` (``:11-375``): reads -> subpath constraints of a breakpoint graph.

Same callables, argument meaning and return shapes as the reference module:
``valid_path``, ``alignment_to_path``, ``chimeric_alignment_to_path_l / _r / _i``, ``traverse_through_sequence_edge``,
``chimeric_alignment_to_path``.  A path is a list alternating edges ``('s'|'c'|'d', index)`` and nodes
``(chr, pos, orientation)``.

These walk ONE graph for ONE read (tens of list entries): control-plane host logic.  The data-parallel part of the
step — every record of the amplicon intervals mapped onto its run of sequence edges, reference
``infer_breakpoint_graph.py:1303-1321`` — runs on the GPU (``csrc/paths.cu``, ``coral_alignment_paths``) and only comes
back here as distinct ``(first edge, last edge)`` runs to be spelled out by ``sequence_run_to_path``.
"""

import bisect

edge_type_to_index = {'s': 0, 'c': 1, 'd': 2}      # slot of each edge kind in g.nodes[node] (reference :7)


def valid_path(g, path):
    """Reference :11-45.  Odd length > 3, sequence edges at both ends, edge kinds alternate between 's' and a
    breakpoint kind, and every node is incident to both of its neighbouring edges."""
    n = len(path)
    if n <= 3 or n % 2 == 0:
        return False
    if path[0][0] != 's' or path[-1][0] != 's':
        return False
    for k, item in enumerate(path):
        if k % 2 == 0:
            if len(item) != 2:
                return False
            continue
        if len(item) != 3:
            return False
        left, right = path[k - 1], path[k + 1]
        try:
            if (left[0] == 's') == (right[0] == 's'):
                return False
            adj = g.nodes[item]
            if left[1] not in adj[edge_type_to_index[left[0]]] or right[1] not in adj[edge_type_to_index[right[0]]]:
                return False
        except Exception:       # the reference swallows everything here (:43)
            return False
    return True


# ---- shared pieces ---------------------------------------------------------------------------------------------
def _edge_index(g):
    """Per contig: (starts, ends, ids) of the sequence edges by start, when ends then ascend too (disjoint edges, what
    build_graph makes); None for a contig whose edges overlap.  Rebuilt when the edge list changes length."""
    cache = getattr(g, "_se_lookup", None)
    if cache is None or cache[0] != len(g.sequence_edges):
        by_chr = {}
        for k, e in enumerate(g.sequence_edges):
            by_chr.setdefault(e[0], []).append((e[1], k, e[2]))
        table = {}
        for c, rows in by_chr.items():
            rows.sort(key=lambda r: r[0])                     # stable: equal starts keep edge-id order
            ends = [r[2] for r in rows]
            ok = all(ends[i] < ends[i + 1] for i in range(len(ends) - 1))
            table[c] = ([r[0] for r in rows], ends, [r[1] for r in rows]) if ok else None
        cache = (len(g.sequence_edges), table)
        g._se_lookup = cache
    return cache[1]


def _touching(g, chrom, lo, hi):
    """Sequence-edge ids with interval_overlap([chrom, lo, hi], edge) (breakpoint_utilities.py:11-15), by edge start."""
    idx = _edge_index(g).get(chrom, ())
    if idx is None:       # overlapping edges on this contig: the literal scan
        ids = [k for k, e in enumerate(g.sequence_edges) if e[0] == chrom and lo <= e[2] and e[1] <= hi]
        ids.sort(key=lambda k: g.sequence_edges[k][1])
        return ids
    if not idx:
        return []
    starts, ends, ids = idx
    return ids[bisect.bisect_left(ends, lo):bisect.bisect_right(starts, hi)]


def _shared(g, k, lo, hi):
    e = g.sequence_edges[k]
    return min(e[2], hi) - max(e[1], lo)


def _concordant_index(g):
    """(chr, pos1, pos2) -> lowest concordant-edge id: what the reference's linear searches (:168-173 …) return."""
    cache = getattr(g, "_cc_lookup", None)
    if cache is None or cache[0] != len(g.concordant_edges):
        table = {}
        for ci, ce in enumerate(g.concordant_edges):
            table.setdefault((ce[0], ce[1], ce[4]), ci)
        cache = (len(g.concordant_edges), table)
        g._cc_lookup = cache
    return cache[1]


def _oriented_run(g, al):
    """Edges touched by one local alignment in the direction the read walks them: [(edge id, '+'|'-')], plus the
    alignment's (lo, hi) on the reference."""
    if al[-1] == '+':
        lo, hi, sign = al[1], al[2], '+'
    else:
        lo, hi, sign = al[2], al[1], '-'
    ids = _touching(g, al[0], lo, hi)
    if sign == '-':
        ids.reverse()
    return [(k, sign) for k in ids], lo, hi


def _exit_node(g, k, sign):      # node through which a read walking edge k in direction `sign` leaves it
    e = g.sequence_edges[k]
    return (e[0], e[2], '+') if sign == '+' else (e[0], e[1], '-')


def _entry_node(g, k, sign):
    e = g.sequence_edges[k]
    return (e[0], e[1], '-') if sign == '+' else (e[0], e[2], '+')


def _junction(g, k, k_next, sign):
    """Concordant edge between consecutive edges of a run, or None when they are not adjacent / no such edge."""
    a, b = g.sequence_edges[k], g.sequence_edges[k_next]
    if sign == '+':
        if a[2] + 1 != b[1]:
            return None
        return _concordant_index(g).get((a[0], a[2], b[1]))
    if a[1] - 1 != b[2]:
        return None
    return _concordant_index(g).get((a[0], b[2], a[1]))


def sequence_run_to_path(g, first, last):
    """The walk of alignment_to_path once its edge run [first .. last] is known (:85-90)."""
    a, b = g.sequence_edges[first], g.sequence_edges[last]
    return traverse_through_sequence_edge(g, (a[0], a[1], '-'), (b[0], b[2], '+'))[1:-1]


def trimmed_run(g, rint, min_overlap=500):
    """Edge ids of alignment_to_path after its end trimming (:60-84), or [] when fewer than three are left."""
    ids = _touching(g, rint[0], rint[1], rint[2])
    if not ids:
        return []
    size = lambda k: g.sequence_edges[k][7]
    if len(ids) > 1 and _shared(g, ids[0], rint[1], rint[2]) < min_overlap:
        del ids[0]
    while len(ids) > 1 and size(ids[0]) < min_overlap:
        del ids[0]
    if len(ids) > 1 and _shared(g, ids[-1], rint[1], rint[2]) < min_overlap:
        del ids[-1]
    while len(ids) > 1 and size(ids[-1]) < min_overlap:
        del ids[-1]
    return ids if len(ids) > 2 else []


def alignment_to_path(g, rint, min_overlap=500):
    """Reference :48-90: one alignment [chr, start, end] -> path through sequence and concordant edges."""
    ids = trimmed_run(g, rint, min_overlap)
    if not ids:
        return []
    return sequence_run_to_path(g, ids[0], ids[-1])


def chimeric_alignment_to_path_l(g, rints, ai, bp_node, min_overlap=500):
    """Reference :93-185: from alignment `ai` up to the breakpoint node it ends at.  The result starts with an
    edge and ends with a node."""
    run, lo, hi = _oriented_run(g, rints[ai])
    if not run:
        return []
    if len(run) > 1 and _shared(g, run[0][0], lo, hi) < min_overlap:
        del run[0]
    while run and g.sequence_edges[run[0][0]][7] < min_overlap:
        del run[0]
    while run and _exit_node(g, *run[-1]) != bp_node:       # the far end must be the breakpoint node
        del run[-1]
    out = []
    for n, (k, sign) in enumerate(run):
        out.append(('s', k))
        out.append(_exit_node(g, k, sign))
        if n + 1 < len(run):
            ci = _junction(g, k, run[n + 1][0], sign)
            if ci is not None:
                out.append(('c', ci))
                out.append(_entry_node(g, run[n + 1][0], sign))
    return out


def chimeric_alignment_to_path_r(g, rints, ai, bp_node, min_overlap=500):
    """Reference :188-277: from the breakpoint node into alignment `ai`.  The result starts with a node and ends
    with an edge."""
    run, lo, hi = _oriented_run(g, rints[ai])
    if not run:
        return []
    if _shared(g, run[-1][0], lo, hi) < 500:          # not parameterised in the reference either (:215, :237)
        del run[-1]
    if not run:
        return []
    while run and g.sequence_edges[run[-1][0]][7] < 500:
        del run[-1]
    while run and _entry_node(g, *run[0]) != bp_node:
        del run[0]
    out = []
    for n, (k, sign) in enumerate(run):
        out.append(_entry_node(g, k, sign))
        out.append(('s', k))
        if n + 1 < len(run):
            ci = _junction(g, k, run[n + 1][0], sign)
            if ci is not None:
                out.append(_exit_node(g, k, sign))
                out.append(('c', ci))
    return out


def chimeric_alignment_to_path_i(g, rints, ai1, ai2, di):
    """Reference :280-301: alignment ai1 -> discordant edge di -> alignment ai2 (read order = smaller index first)."""
    de = g.discordant_edges[di]
    node1, node2 = (de[0], de[1], de[2]), (de[3], de[4], de[5])
    if ai1 > ai2:
        return chimeric_alignment_to_path_l(g, rints, ai2, node2) + [('d', di)] + chimeric_alignment_to_path_r(g, rints, ai1, node1)
    return chimeric_alignment_to_path_l(g, rints, ai1, node1) + [('d', di)] + chimeric_alignment_to_path_r(g, rints, ai2, node2)


def traverse_through_sequence_edge(g, start_node, end_node):
    """Reference :304-343: walk sequence / concordant edges from start_node until end_node; the result starts and
    ends with a node.  A node without a concordant edge ends the walk early (alignments spanning two intervals)."""
    assert start_node[2] != end_node[2]
    out = [start_node]
    here = start_node
    while True:
        k = g.nodes[here][0][0]
        e = g.sequence_edges[k]
        far = (e[0], e[2], '+') if here[2] == '-' else (e[0], e[1], '-')
        out.append(('s', k))
        out.append(far)
        if far == end_node:
            return out
        try:
            ci = g.nodes[far][1][0]
        except Exception:
            return out
        ce = g.concordant_edges[ci]
        here = (ce[0], ce[1], ce[2])
        if here == far:
            here = (ce[3], ce[4], ce[5])
        out.append(('c', ci))
        out.append(here)


def chimeric_alignment_to_path(g, rints, ai_list, bp_list):
    """Reference :346-375: a block of consecutive breakpoints of one read -> one path."""
    out = []
    last = ()
    n = len(bp_list)
    for i, di in enumerate(bp_list):
        de = g.discordant_edges[di]
        node1, node2 = (de[0], de[1], de[2]), (de[3], de[4], de[5])
        a, b = ai_list[i][0], ai_list[i][1]
        if a > b:            # the read meets node2 first
            near, far, a_in, a_out = node2, node1, b, a
        else:
            near, far, a_in, a_out = node1, node2, a, b
        if i == 0:
            out = chimeric_alignment_to_path_l(g, rints, a_in, near) + [('d', di)]
        else:
            out += traverse_through_sequence_edge(g, last, near)
            out.append(('d', di))
            if i == n - 1:
                out += chimeric_alignment_to_path_r(g, rints, a_out, far)
        last = far
    return out


def longest_path_dict(path_constraints_):
    """Reference :378-417 (called by cycle_decomposition.py:1512, the consumer of compute_path_constraints): every
    path as its multiset of edges; a path whose multiset is contained in another path's is dropped — scanning from the
    last path to the first, the container being the first other path that holds it — and hands its support to the
    container when larger.  -> [edge multisets, indices of the surviving paths, supports]."""
    bags = []
    for path in path_constraints_[0]:
        bag = {}
        for edge in path[0::2]:
            bag[edge] = bag.get(edge, 0) + 1
        bags.append(bag)
    ids = list(range(len(bags)))
    support = list(path_constraints_[1][:len(bags)])
    for i in range(len(bags) - 1, -1, -1):
        mine = bags[i]
        host = -1
        for j, other in enumerate(bags):
            if j != i and all(other.get(e, 0) >= c for e, c in mine.items()):
                host = j
                break
        if host >= 0:
            support[host] = max(support[host], support[i])
            del bags[i], ids[i], support[i]
    return [bags, ids, support]
