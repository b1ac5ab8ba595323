"""Golden vectors for the HSR mode's breakpoint search (the second caller of the hot
operators), made by EXECUTING THE REFERENCE: src/hsr.py `fetch` and `locate_hsrs` run unmodified under the oracle's
pysam shim, with recording stand-ins for matplotlib / pylab (the figure itself is out of scope; what is kept are the
`Breakpoint ... Support = n` lines the reference prints and the points it hands to `plt.plot`).

    python tests/golden/make_hsr_golden.py        # writes tests/golden/hsr_*.json.gz (needs /root/reference)

Cases: a cycle of which only some segments are declared ecDNA, so that its other junctions look like integration
sites; `hsr_inputs(case)` is shared with tests/test_gpu_hsr.py.
"""
import contextlib
import gzip
import io
import json
import os
import sys
import tempfile
import types

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

CASES = {
    # name: (config, kwargs, ecDNA intervals written to the cycles .bed, normal_cov, bp_match_cutoff, clustering cutoff)
    # (the flat variants hand in a CN-2 segment file: only ends in CN <= 5 segments are reported as integration loci)
    "hsr_c1": ("C1", dict(scale=0.2), [("chr7", 10_000_000, 10_500_000), ("chr7", 10_600_000, 10_900_000)], 3.0, 100, 2000, False),
    "hsr_c1_flat": ("C1", dict(scale=0.2), [("chr7", 10_000_000, 10_500_000), ("chr7", 10_600_000, 10_900_000)], 30.0, 100, 2000, True),
    "hsr_c4": ("C4", dict(scale=0.012), [("chr1", 100_000_000, 102_000_000), ("chr12", 60_000_000, 64_000_000)], 2.0, 100, 2000, False),
    "hsr_c4_flat": ("C4", dict(scale=0.012), [("chr1", 100_000_000, 102_000_000), ("chr12", 60_000_000, 64_000_000)], 2.0, 150, 1500, True),
}


def hsr_inputs(case, outdir):
    """Dataset + the argument namespace of `CoRAL.py hsr` for one case."""
    from sim import configs
    from sim.generate import simulate, write_dataset
    name, kw, ecdna, normal_cov, cut, cut_cl, flat = CASES[case]
    cfg = configs.CONFIGS[name](**kw)
    data = simulate(cfg)
    paths = write_dataset(data, outdir)
    if flat:
        paths["cn_seg"] = os.path.join(outdir, "cn_flat.bed")
        with open(paths["cn_seg"], "w") as fp:
            for (c, s, e, cn) in cfg.cn_rows:
                fp.write("%s\t%d\t%d\t2.0\n" % (c, s, e))
    bed = os.path.join(outdir, "ecdna_cycles.bed")
    with open(bed, "w") as fp:
        fp.write("#chr\tstart\tend\n")
        for c, s, e in ecdna:
            fp.write("%s\t%d\t%d\n" % (c, s, e))
    args = types.SimpleNamespace(lr_bam=paths["bam"], cycles=bed, cn_seg=paths["cn_seg"], normal_cov=normal_cov,
                                 bp_match_cutoff=cut, bp_match_cutoff_clustering=cut_cl,
                                 output_prefix=os.path.join(outdir, "out"))
    return data, args


class _Recorder(object):
    def __init__(self):
        self.points = []

    def plot(self, x, y, *a, **k):
        if a and a[0] == 'bo':
            self.points.append([x, y])

    def __getattr__(self, name):
        return lambda *a, **k: None


def install_plot_stubs():
    rec = _Recorder()
    mpl = types.ModuleType("matplotlib")
    mpl.use = lambda *a, **k: None
    mpl.rc = lambda *a, **k: None
    mpl.mlab = types.ModuleType("matplotlib.mlab")
    mpl.pyplot = rec
    pylab = types.ModuleType("pylab")
    pylab.rcParams = {}
    sys.modules.update({"matplotlib": mpl, "matplotlib.mlab": mpl.mlab, "matplotlib.pyplot": rec, "pylab": pylab})
    return rec


def main():
    from oracle import ref_runner
    only = sys.argv[1:]
    for case in CASES:
        if only and case not in only:
            continue
        d = tempfile.mkdtemp(prefix="golden_%s_" % case)
        data, args = hsr_inputs(case, d)
        ref_runner.install_shims()
        rec = install_plot_stubs()
        import hsr as ref_hsr
        import pysam
        fh = pysam.AlignmentFile(args.lr_bam, 'rb')
        read_length, chim = ref_hsr.fetch(fh)
        out = io.StringIO()
        cwd = os.getcwd()
        os.chdir(d)
        try:
            with contextlib.redirect_stdout(out):
                ref_hsr.locate_hsrs(args)
        finally:
            os.chdir(cwd)
        lines = [ln for ln in out.getvalue().splitlines()
                 if ln.startswith(("Fetched", "Found", "Breakpoint"))]
        st = {"case": case, "n_records": data.n_records, "lines": lines, "points": rec.points,
              "fetch": None if CASES[case][-1] else ref_runner.canon_chimeric(chim)}   # flat variants: same BAM
        with gzip.open(os.path.join(HERE, case + ".json.gz"), "wt") as fp:
            json.dump(st, fp, separators=(",", ":"))
        print(case, "records", data.n_records, "chimeric", len(chim), lines[:2], "integration points", len(rec.points))
        for m in ("hsr", ):
            sys.modules.pop(m, None)


if __name__ == "__main__":
    main()
