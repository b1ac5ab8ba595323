"""Golden vectors for the coverage track of plot_amplicons.py (reference :376-412, SURVEY.md §8f rank 4, second half),
made by EXECUTING THE REFERENCE: graph_vis.open_bam / parse_graph_file / graph_amplified_intervals / plot_graph run
unmodified under the oracle's pysam and intervaltree shims, with recording stand-ins for matplotlib / pylab (the
figure is out of scope; kept are the coverage rectangles — those drawn in 'silver' — in drawing order and the y limit).

    PYTHONHASHSEED=0 python tests/golden/make_plot_cov_golden.py      # writes tests/golden/plot_cov.json.gz

Inputs: the golden pipeline datasets with the graph.txt the reference wrote for them (tests/golden/<case>.json.gz);
each once over the whole graph and once zoomed to a region (plot_bounds).
"""
import gzip
import json
import os
import sys
import tempfile
import types
from unittest import mock

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
OUT = os.path.join(HERE, "plot_cov.json.gz")
REF_SRC = os.environ.get("CORAL_REFERENCE_SRC", "/root/reference/src")
CASES = ["c1_small", "c2_small", "c3_small", "c4_small"]


class Recorder(object):
    def __init__(self):
        self.rects = []
        self.ylims = []


def install_stand_ins(rec):
    from oracle import intervaltree_shim, pysam_shim

    def rectangle(xy, width, height, **kw):
        if kw.get("color") == "silver":
            rec.rects.append([float(xy[0]), float(width), float(height)])
        return mock.MagicMock()

    mpl = mock.MagicMock()
    patches = types.ModuleType("matplotlib.patches")
    patches.Rectangle = rectangle
    patches.Arc = mock.MagicMock()
    mods = {"matplotlib": mpl, "matplotlib.gridspec": mpl.gridspec, "matplotlib.pyplot": mpl.pyplot,
            "matplotlib.patches": patches, "matplotlib.ticker": mpl.ticker, "pylab": mock.MagicMock(),
            "pysam": pysam_shim, "intervaltree": intervaltree_shim, "cycle2bed": mock.MagicMock()}
    sys.modules.update(mods)
    return mpl


def main():
    import make_golden
    from sim.generate import simulate, write_dataset
    rec = Recorder()
    mpl = install_stand_ins(rec)
    sys.path.insert(0, REF_SRC)
    import plot_amplicons as pa
    import breakpoint_utilities as bu
    vectors = []
    for case in CASES:
        with gzip.open(os.path.join(HERE, case + ".json.gz"), "rt") as fp:
            gold = json.load(fp)
        cfg = make_golden.make_cfg(case)
        data = simulate(cfg)
        d = tempfile.mkdtemp(prefix="plotcov_")
        paths = write_dataset(data, d)
        for gi, gr in enumerate(gold["stages"]["assign_cov"]["graphs"]):
            gfn = os.path.join(d, "amplicon%d_graph.txt" % (gi + 1))
            with open(gfn, "w") as fp:
                fp.write(gr["graph_txt"])
            for zoom in (False, True):
                g = pa.graph_vis()
                g.sequence_edges_by_chr, g.intervals_from_graph, g.discordant_edges = dict(), dict(), []   # class-level in the reference
                g.num_amplified_intervals = 0
                g.plot_bounds = None
                g.open_bam(paths["bam"])
                g.parse_graph_file(gfn)
                g.graph_amplified_intervals()
                ivs = {c: [list(x) for x in v] for c, v in g.intervals_from_graph.items()}
                if not ivs:
                    continue
                if zoom:        # a region inside the longest interval, cutting through it
                    c, iv = max(((c, x) for c, v in ivs.items() for x in v), key=lambda t: t[1][1] - t[1][0])
                    ln = iv[1] - iv[0]
                    g.plot_bounds = (c, iv[0] + ln // 3, iv[0] + ln // 3 + max(2000, ln // 4))
                rec.rects, rec.ylims = [], []
                ax_calls = []
                fig = mpl.pyplot.figure.return_value
                ax = fig.add_subplot.return_value
                ax.set_ylim.side_effect = lambda lo, hi: ax_calls.append([float(lo), float(hi)])
                g.plot_graph("t", os.path.join(d, "out"), hide_genes=True)
                ax.set_ylim.side_effect = None
                vectors.append({"case": case, "graph": gi, "plot_bounds": list(g.plot_bounds) if g.plot_bounds else None,
                                "intervals_from_graph": ivs, "sorted_chrs": list(bu.sort_chrom_names(ivs.keys())),
                                "rects": rec.rects, "ylims": ax_calls})
    with gzip.open(OUT, "wt") as fp:
        json.dump(vectors, fp, separators=(",", ":"))
    print(len(vectors), "vectors,", sum(len(v["rects"]) for v in vectors), "coverage windows,", os.path.getsize(OUT) // 1024, "KB")
    for v in vectors:
        print(" ", v["case"], v["graph"], v["plot_bounds"], len(v["rects"]), "windows, max height %.2f" % max([r[2] for r in v["rects"]] or [0]), v["ylims"][-2:])


if __name__ == "__main__":
    main()
