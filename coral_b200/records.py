"""In-memory record set with the small interface the drop-in class needs from its BAM handle."""
import numpy as np


def query_names(rec, rec_off, idx):
    """Read names of records `idx` in one pass over the host buffer (the per-record form is query_name)."""
    idx = np.asarray(idx, dtype=np.int64)
    if len(idx) == 0:
        return []
    base = np.asarray(rec_off, dtype=np.uint64)[idx].astype(np.int64)
    ln = rec[base + 12].astype(np.int64)                      # l_read_name, NUL included
    offs = np.cumsum(ln) - ln
    pos = np.repeat(base + 36 - offs, ln) + np.arange(int(ln.sum()), dtype=np.int64)
    return rec[pos].tobytes().decode().split("\0")[:len(idx)]


class RecordSet(object):
    def __init__(self, references, rec, rec_off, lengths=None, rec_slack=0):
        self.rec_slack = int(rec_slack)   # bytes guaranteed readable behind `rec` (64 enables the mapped placement)
        self.references = tuple(references)
        self.lengths = tuple(lengths) if lengths is not None else tuple(0 for _ in references)
        self.rec = rec
        self.rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        self.n_records = len(self.rec_off) - 1

    def query_name(self, i):
        o = int(self.rec_off[i])
        ln = int(self.rec[o + 12])
        return self.rec[o + 36:o + 36 + ln - 1].tobytes().decode()

    def query_names(self, idx):
        return query_names(self.rec, self.rec_off, idx)

    def close(self):
        pass
