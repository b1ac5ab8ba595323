"""The five BASELINE.json configs as SimConfig factories.

``scale`` multiplies coverage (1.0 = the named shape; tests use 0.02-0.2 so the oracle finishes
in seconds).  Contigs are named from ``global_names.chr_idx`` only (other names crash
``interval2bp``, reference ``breakpoint_utilities.py:293``).  ``cn_seg.bed`` always carries >= 10 Mb
of CN-2 segments around the median (else ``read_cns`` raises IndexError, reference
``infer_breakpoint_graph.py:118-125``).
"""
from __future__ import annotations

import numpy as np

from .generate import CHR_SIZES, Cycle, SimConfig

MB = 1_000_000


def _tile_cn(chrom, start, end, amps, step=5 * MB):
    """CN rows for [start, end): CN-2 tiles of `step`, with amplified (s, e, cn) segments cut in."""
    cuts = sorted(amps)
    rows, pos = [], start
    for (s, e, cn) in cuts:
        while pos < s:
            nxt = min(s, pos + step)
            rows.append((chrom, pos, nxt, 2.0)); pos = nxt
        rows.append((chrom, s, e, cn)); pos = e
    while pos < end:
        nxt = min(end, pos + step)
        rows.append((chrom, pos, nxt, 2.0)); pos = nxt
    return rows


def c1_single_ecdna(scale=1.0, seed=20240215 + 1, cov=30.0, bg_span=30 * MB):
    """C1: one 2 Mb amplicon on chr7, 4 SVs (inversion, back-jump, deletion-like, closing), ONT 20 kb."""
    A = ('chr7', 10_000_000, 10_500_000, '+')
    B = ('chr7', 11_000_000, 11_400_000, '-')
    C = ('chr7', 10_600_000, 10_900_000, '+')
    D = ('chr7', 11_500_000, 12_000_000, '+')
    cn_rows = _tile_cn('chr7', 0, bg_span, [(10 * MB, 12 * MB, 20.0)])
    return SimConfig(
        name="C1", refs=[('chr7', CHR_SIZES['chr7'])], cn_rows=cn_rows, seeds=[('chr7', 10 * MB, 12 * MB - 1)],
        background=[('chr7', 0, bg_span)], cov=cov * scale, cycles=[Cycle([A, B, C, D], 20.0)],
        read_mean=20000, model="ont", seed=seed)


def c2_multi_segment(scale=1.0, seed=20240215 + 2, cov=50.0, bg_pad=6 * MB):
    """C2: 3 seeds on 3 chromosomes, 6 Mb total, CN 40, 12 junctions, ONT 30 kb."""
    segs = [
        ('chr1', 50_000_000, 50_700_000, '+'), ('chr8', 127_000_000, 127_600_000, '-'),
        ('chr1', 51_200_000, 51_800_000, '+'), ('chr12', 69_000_000, 69_500_000, '+'),
        ('chr8', 128_000_000, 128_500_000, '+'), ('chr12', 70_200_000, 70_700_000, '-'),
        ('chr1', 50_800_000, 51_100_000, '-'), ('chr8', 127_650_000, 127_950_000, '+'),
        ('chr12', 69_600_000, 70_100_000, '+'), ('chr1', 51_850_000, 52_000_000, '+'),
        ('chr8', 128_550_000, 129_000_000, '-'), ('chr12', 70_750_000, 71_000_000, '+'),
    ]
    spans = {'chr1': (50 * MB, 52 * MB), 'chr8': (127 * MB, 129 * MB), 'chr12': (69 * MB, 71 * MB)}
    refs = [(c, CHR_SIZES[c]) for c in ('chr1', 'chr8', 'chr12')]
    cn_rows, bg, seeds = [], [], []
    for c, (s, e) in spans.items():
        lo, hi = s - bg_pad, e + bg_pad
        cn_rows += _tile_cn(c, lo, hi, [(s, e, 40.0)], step=3 * MB)
        bg.append((c, lo, hi)); seeds.append((c, s, e - 1))
    return SimConfig(name="C2", refs=refs, cn_rows=cn_rows, seeds=seeds, background=bg, cov=cov * scale,
                     cycles=[Cycle(segs, 40.0)], read_mean=30000, model="ont", seed=seed)


def c3_smalldel_heavy(scale=1.0, seed=20240215 + 3, cov=60.0, n_del=300):
    """C3: HiFi 60x, one 3 Mb amplicon, 300 in-CIGAR deletions of 700-5000 bp + dense small D ops."""
    rng = np.random.Generator(np.random.PCG64(seed))
    s, e = 30 * MB, 33 * MB
    slots = np.linspace(s + 20_000, e - 20_000, n_del + 1)[:-1].astype(int)
    dels = []
    for a in slots:
        ln = int(rng.integers(700, 5001))
        dels.append(('chr3', int(a), int(a) + ln, float(rng.uniform(0.3, 1.0))))
    cn_rows = _tile_cn('chr3', 20 * MB, 45 * MB, [(s, e, 16.0)])
    return SimConfig(
        name="C3", refs=[('chr3', CHR_SIZES['chr3'])], cn_rows=cn_rows, seeds=[('chr3', s, e - 1)],
        background=[('chr3', 20 * MB, 45 * MB)], cov=cov * scale,
        cycles=[Cycle([('chr3', s, e, '+')], 16.0)], deletions=dels, read_mean=15000, model="hifi",
        dense_del_boost=10.0, seed=seed)


def c4_complex(scale=1.0, seed=20240215 + 4, cov=100.0, n_junc=200):
    """C4: 200 junctions over chr1/7/12/17, ONT 50 kb, CN 10-60 per cycle."""
    rng = np.random.Generator(np.random.PCG64(seed))
    chroms = ['chr1', 'chr7', 'chr12', 'chr17']
    base = {'chr1': 100 * MB, 'chr7': 50 * MB, 'chr12': 60 * MB, 'chr17': 35 * MB}
    span = 4 * MB
    per = n_junc // 4
    cycles, amps = [], {c: [] for c in chroms}
    # carve each chromosome's 4 Mb window into `per` segments, then deal them round-robin into 4 cycles
    allsegs = []
    for c in chroms:
        edges = np.sort(rng.choice(np.arange(base[c] + 40_000, base[c] + span - 40_000, 10_000), per - 1, replace=False))
        edges = np.concatenate([[base[c]], edges, [base[c] + span]])
        for i in range(per):
            allsegs.append((c, int(edges[i]), int(edges[i + 1]) - 3000, '+' if rng.random() < 0.6 else '-'))
    perm = rng.permutation(len(allsegs))
    groups = [perm[i::4] for i in range(4)]
    for gi, g in enumerate(groups):
        cycles.append(Cycle([allsegs[j] for j in g], float([10, 25, 40, 60][gi])))
    refs = [(c, CHR_SIZES[c]) for c in chroms]
    cn_rows, bg, seeds = [], [], []
    for c in chroms:
        lo, hi = base[c] - 6 * MB, base[c] + span + 6 * MB
        cn_rows += _tile_cn(c, lo, hi, [(base[c], base[c] + span, 30.0)], step=3 * MB)
        bg.append((c, lo, hi)); seeds.append((c, base[c], base[c] + span - 1))
    return SimConfig(name="C4", refs=refs, cn_rows=cn_rows, seeds=seeds, background=bg, cov=cov * scale,
                     cycles=cycles, read_mean=50000, model="ont", seed=seed)


def c5_whole_genome(scale=1.0, seed=20240215 + 5, cov=30.0, n_seeds=20, chroms=None):
    """C5: every chromosome at 30x, 20 seed intervals (2-segment cycles), 0.5 % chimeric noise reads."""
    rng = np.random.Generator(np.random.PCG64(seed))
    chroms = chroms or (['chr%d' % i for i in range(1, 23)] + ['chrX', 'chrY'])
    refs = [(c, CHR_SIZES[c]) for c in chroms]
    cn_rows, bg, seeds, cycles = [], [], [], []
    amp_by_chr = {c: [] for c in chroms}
    for k in range(n_seeds):
        c = chroms[k % len(chroms)]
        s = int(rng.integers(20, 40)) * MB + (k // len(chroms)) * 4 * MB
        amp_by_chr[c].append((s, s + 1 * MB, 12.0))
        seeds.append((c, s, s + MB - 1))
        cycles.append(Cycle([(c, s, s + 400_000, '+'), (c, s + 500_000, s + MB, '-')], 12.0))
    for c in chroms:
        cn_rows += _tile_cn(c, 0, CHR_SIZES[c] // MB * MB, amp_by_chr[c], step=10 * MB)
        bg.append((c, 0, CHR_SIZES[c] // MB * MB))
    return SimConfig(name="C5", refs=refs, cn_rows=cn_rows, seeds=seeds, background=bg, cov=cov * scale,
                     cycles=cycles, read_mean=20000, model="ont", noise_frac=0.005, seed=seed)


ALL_CHROMS = ['chr%d' % i for i in range(1, 23)] + ['chrX', 'chrY']


def c2_shard(rank, world, scale=1.0, seed=20240215 + 2, cov=50.0, bg_pad=6 * MB):
    """Weak-scaling form of C2 for the multi-GPU bench: rank r carries a full C2-shaped sample (3 seeds / 6 Mb,
    CN 40, 12 junctions, ONT 30 kb) on its own three chromosomes, ALL_CHROMS[3r : 3r + 3], against the common
    24-contig header — so the ranks' record sets, taken in rank order, are the contiguous genomic-bin shards of ONE
    coordinate-sorted file.  Returns (this rank's SimConfig, cn rows of all ranks, seeds of all ranks)."""
    assert 1 <= world <= 8
    refs = [(c, CHR_SIZES[c]) for c in ALL_CHROMS]
    lay = [(0, 0, 700_000, '+'), (1, 0, 600_000, '-'), (0, 1_200_000, 1_800_000, '+'), (2, 0, 500_000, '+'),
           (1, 1_000_000, 1_500_000, '+'), (2, 1_200_000, 1_700_000, '-'), (0, 800_000, 1_100_000, '-'),
           (1, 650_000, 950_000, '+'), (2, 600_000, 1_100_000, '+'), (0, 1_850_000, 2_000_000, '+'),
           (1, 1_550_000, 2_000_000, '-'), (2, 1_750_000, 2_000_000, '+')]
    base = 20 * MB
    all_cn, all_seeds, mine = [], [], None
    for r in range(world):
        ch = ALL_CHROMS[3 * r:3 * r + 3]
        segs = [(ch[k], base + a, base + b, o) for (k, a, b, o) in lay]
        cn_rows, bg, seeds = [], [], []
        for c in ch:
            lo, hi = base - bg_pad, base + 2 * MB + bg_pad
            cn_rows += _tile_cn(c, lo, hi, [(base, base + 2 * MB, 40.0)], step=3 * MB)
            bg.append((c, lo, hi)); seeds.append((c, base, base + 2 * MB - 1))
        all_cn += cn_rows; all_seeds += seeds
        if r == rank:
            mine = SimConfig(name="C2", refs=refs, cn_rows=cn_rows, seeds=seeds, background=bg, cov=cov * scale,
                             cycles=[Cycle(segs, 40.0)], read_mean=30000, model="ont", seed=seed + 1000 * r,
                             name_style=2 + r)       # read names are unique across the file: a letter per shard
    return mine, all_cn, all_seeds


CONFIGS = {"C1": c1_single_ecdna, "C2": c2_multi_segment, "C3": c3_smalldel_heavy, "C4": c4_complex,
           "C5": c5_whole_genome}
