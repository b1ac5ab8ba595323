"""Constants of reference ``global_names.py`` (chromosome ordering used by interval2bp / sort_edges)."""
TSTART = 0  # reference :5 (log timestamps)

neg_plus_minus = {"+": "-", "-": "+"}  # reference :9

_AUTOSOMES = ['chr%d' % i for i in range(1, 23)]
chr_idx = {c: i for i, c in enumerate(_AUTOSOMES + ['chrX', 'chrY', 'chrM'])}  # reference :13-18

chr_sizes = dict(zip(_AUTOSOMES + ['chrX', 'chrY'], [  # reference :20-25 (hg38)
    248956422, 242193529, 198295559, 190214555, 181538259, 170805979, 159345973, 145138636, 138394717, 133797422,
    135086622, 133275309, 114364328, 107043718, 101991189, 90338345, 83257441, 80373285, 58617616, 64444167,
    46709983, 50818468, 156040895, 57227415]))
