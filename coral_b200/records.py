"""In-memory record set with the small interface the drop-in class needs from its BAM handle."""
import numpy as np


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

    def close(self):
        pass
