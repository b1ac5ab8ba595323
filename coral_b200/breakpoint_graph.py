"""Breakpoint-graph container with the reference's row layouts (reference ``breakpoint_graph.py:83-363``
``BreakpointGraph`` and the AA-format writers ``:805-822``, ``:845-854``).

Only what the hot path needs in order to hand over "the GPU-built graph unchanged": edge rows are plain
lists with the reference's column order, so the reference's own ``compute_cn_lr``
(cvxopt) / cycle decomposition (Gurobi) keep working on them.  CN estimation itself is NOT re-implemented.

    sequence   [chr, l, r, sr_count, sr_flag, lr_count, lr_nc, size, cn]
    concordant [chr1, pos1, o1, chr2, pos2, o2, sr_count, sr_flag, lr_count, reads, cn]
    discordant [chr1, pos1, o1, chr2, pos2, o2, sr_count, sr_flag, sr_cn, lr_count, reads, cn]
    source     ['source', -1, '-', chr, pos, o, sr_count, sr_flag, sr_cn, lr_cn, cn]
    nodes[(chr, pos, o)] = [[seq edge ids], [concordant ids], [discordant ids], [source ids]]
"""
import warnings

from . import global_names

_NEED_NODE = "Breakpoint node must be added first."


class BreakpointGraph(object):
    def __init__(self):
        self.amplicon_intervals = []
        self.sequence_edges = []
        self.concordant_edges = []
        self.discordant_edges = []
        self.source_edges = []
        self.nodes = {}
        self.endnodes = {}
        self.max_cn = 0.0

    @staticmethod
    def _check_node(node):
        if type(node) != tuple or len(node) != 3:
            raise Exception("Breakpoint node must be of form (chr, pos, orientation).")

    def add_node(self, node):
        self._check_node(node)
        self.nodes[node] = [[], [], [], []]   # re-adding a node resets its adjacency (reference :122-124)

    def add_endnode(self, node):
        self._check_node(node)
        if node in self.endnodes:
            warnings.warn("Node corresponding to interval end already exists.")
        else:
            self.endnodes[node] = []

    def add_sequence_edge(self, chr, l, r, sr_count=-1, sr_flag='d', lr_count=-1, lr_nc=0, cn=0.0):
        a, b = (chr, l, '-'), (chr, r, '+')
        if a not in self.nodes or b not in self.nodes:
            raise Exception(_NEED_NODE)
        k = len(self.sequence_edges)
        self.nodes[a][0].append(k)
        self.nodes[b][0].append(k)
        self.sequence_edges.append([chr, l, r, sr_count, sr_flag, lr_count, lr_nc, r - l + 1, cn])

    def add_concordant_edge(self, chr1, pos1, o1, chr2, pos2, o2, sr_count=-1, sr_flag='d', lr_count=-1, reads=None, cn=0.0):
        if chr1 != chr2 or pos2 != pos1 + 1 or o1 != '+' or o2 != '-':
            raise Exception("Invalid concordant edge.")
        a, b = (chr1, pos1, o1), (chr2, pos2, o2)
        if a not in self.nodes or b not in self.nodes:
            raise Exception(_NEED_NODE)
        k = len(self.concordant_edges)
        self.nodes[a][1].append(k)
        self.nodes[b][1].append(k)
        self.concordant_edges.append([chr1, pos1, o1, chr2, pos2, o2, sr_count, sr_flag, lr_count,
                                      set() if reads is None else reads, cn])

    def add_discordant_edge(self, chr1, pos1, o1, chr2, pos2, o2, sr_count=-1, sr_flag='d', sr_cn=0.0, lr_count=-1,
                            reads=None, cn=0.0):
        a, b = (chr1, pos1, o1), (chr2, pos2, o2)
        if a not in self.nodes or b not in self.nodes:
            raise Exception(_NEED_NODE)
        k = len(self.discordant_edges)
        self.nodes[a][2].append(k)
        self.nodes[b][2].append(k)
        for nd in (a, b):
            if nd in self.endnodes:
                self.endnodes[nd].append(k)
        self.discordant_edges.append([chr1, pos1, o1, chr2, pos2, o2, sr_count, sr_flag, sr_cn, lr_count,
                                      set() if reads is None else reads, cn])

    def add_source_edge(self, chr1, pos1, o1, sr_count=0, sr_flag='d', sr_cn=0.0, lr_cn=0.0, cn=0.0):
        if (chr1, pos1, o1) not in self.nodes:
            raise Exception(_NEED_NODE)
        self.nodes[(chr1, pos1, o1)][3].append(len(self.source_edges))
        self.source_edges.append(['source', -1, '-', chr1, pos1, o1, sr_count, sr_flag, sr_cn, lr_cn, cn])

    def sort_edges(self):
        """Order sequence and concordant edges by (chr_idx, start) and rebuild their adjacency (reference :348-363)."""
        rank = global_names.chr_idx
        self.sequence_edges.sort(key=lambda e: (rank[e[0]], e[1]))
        self.concordant_edges.sort(key=lambda e: (rank[e[0]], e[1]))
        for k, e in enumerate(self.sequence_edges):
            self.nodes[(e[0], e[1], '-')][0] = [k]
            self.nodes[(e[0], e[2], '+')][0] = [k]
        for k, e in enumerate(self.concordant_edges):
            self.nodes[(e[0], e[1], e[2])][1] = [k]
            self.nodes[(e[3], e[4], e[5])][1] = [k]


def output_breakpoint_graph_lr(g, ogfile):
    """AA graph text with long-read columns (reference :805-822; `%f` CN / mean coverage, `%d` counts)."""
    rows = ["SequenceEdge: StartPosition, EndPosition, PredictedCN, AverageCoverage, Size, NumberOfLongReads\n"]
    for se in g.sequence_edges:
        rows.append("sequence\t%s:%s-\t%s:%s+\t%f\t%f\t%d\t%d\n" % (se[0], se[1], se[0], se[2], se[-1],
                                                                    se[6] * 1.0 / se[7], se[7], se[5]))
    rows.append("BreakpointEdge: StartPosition->EndPosition, PredictedCN, NumberOfLongReads\n")
    for e in g.source_edges:
        rows.append("source\t%s:%s%s->%s:%s%s\t%f\t-1\n" % (e[0], e[1], e[2], e[3], e[4], e[5], e[-1]))
    for e in g.concordant_edges:
        rows.append("concordant\t%s:%s%s->%s:%s%s\t%f\t%d\n" % (e[0], e[1], e[2], e[3], e[4], e[5], e[-1], e[8]))
    for e in g.discordant_edges:
        rows.append("discordant\t%s:%s%s->%s:%s%s\t%f\t%d\n" % (e[0], e[1], e[2], e[3], e[4], e[5], e[-1], e[9]))
    with open(ogfile, 'w') as fp:
        fp.writelines(rows)


def output_breakpoint_info_lr(g, obpfile, bp_stats):
    """Breakpoint table of --output_bp (reference :845-854); stats printed as a Python list repr."""
    with open(obpfile, 'w') as fp:
        fp.write("chr1\tpos1\tchr2\tpos2\torientation\tlr_support\tlr_info=[avg1, avg2, std1, std2, mapq1, mapq2]\n")
        for k, de in enumerate(g.discordant_edges):
            fp.write("%s\t%s\t%s\t%s\t%s%s\t%d\t%s\n" % (de[3], de[4], de[0], de[1], de[5], de[2], de[9], bp_stats[k]))
