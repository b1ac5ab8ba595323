"""Drop-in for reference ``cigar_parsing.py``: same callables, same argument and return shapes; the
arithmetic runs in the CUDA library (``coral_sa_text``: SA-entry parse, the nine cigar2pos forms, rint
construction and the stable (qs, qe) order).  Host code only splits strings and rebuilds Python lists."""
import logging
import time

import numpy as np

from . import gpu_ops, global_names

_SIGS = ("SM", "MS", "SMS", "SMD", "MDS", "SMDS", "SMI", "MIS", "SMIS")


def _run(entries_per_read, read_lengths, ctx=None):
    """entries_per_read: list of lists of SA entry strings -> per-read device results."""
    ctx = ctx or gpu_ops.default_context()
    blobs, beg, end, ptr = [], [], [], [0]
    pos = 0
    for ents in entries_per_read:
        for e in ents:
            b = e.encode()
            blobs.append(b); beg.append(pos); end.append(pos + len(b)); pos += len(b)
        ptr.append(len(beg))
    return ctx.sa_text(b"".join(blobs), np.array(ptr, np.uint32), np.array(beg, np.uint32), np.array(end, np.uint32),
                       np.array(read_lengths, np.int64))


def _cigar2pos(sig, cigar, strand, read_length):
    if "".join(ch for ch in cigar if not ch.isdigit()) != sig:
        raise ValueError("CIGAR %r is not of the form %s" % (cigar, sig))
    r = _run([["x,1,%s,%s,0,0" % (strand, cigar)]], [read_length])
    if r["status"][0] != 0:
        raise ValueError("CIGAR %r is not of the form %s" % (cigar, sig))
    return int(r["qs"][0]), int(r["qe"][0]), int(r["al"][0])


def _make(sig):
    def f(cigar, strand, read_length):
        return _cigar2pos(sig, cigar, strand, read_length)
    f.__name__ = "cigar2pos" + sig
    f.__doc__ = "Convert a CIGAR of the form %s (reference cigar_parsing.py:17-215) -> (qs, qe, al)." % sig
    return f


cigar2posSM, cigar2posMS, cigar2posSMS, cigar2posSMD, cigar2posMDS, cigar2posSMDS, cigar2posSMI, cigar2posMIS, \
    cigar2posSMIS = (_make(s) for s in _SIGS)
cigar2pos_ops = {s: globals()["cigar2pos" + s] for s in _SIGS}  # reference :219-229


def alignment_from_satags(sa_list, read_length):
    """Reference cigar_parsing.py:232-269.  Returns (qint, rint, qual, nm) or ([], [], []) on the reject path."""
    if len(sa_list) == 0:
        return ([], [], [], [])
    r = _run([sa_list], [read_length])
    st = int(r["status"][0])
    if st == 2:  # cp:248-253
        logging.warning("#TIME " + '%.4f\t' % (time.time() - global_names.TSTART) + "Found chimeric alignment without match or soft clips.")
        logging.warning("#TIME " + '%.4f\t' % (time.time() - global_names.TSTART) + "\tAll CIGAR strings: %s." % (sa_list))
        return ([], [], [])
    if st == 3:
        raise KeyError("CIGAR signature not in cigar2pos_ops")
    if st != 0:
        raise ValueError("malformed SA entry in %s" % (sa_list,))
    n = len(sa_list)
    qint = [[int(r["qs"][k]), int(r["qe"][k])] for k in range(n)]
    rint = [[sa_list[int(r["src"][k])].split(',')[0], int(r["rstart"][k]), int(r["rend"][k]), '-' if r["strand"][k] else '+']
            for k in range(n)]
    qual = [int(r["mapq"][k]) for k in range(n)]
    nm = [float(int(r["nm"][k])) / (qint[k][1] - qint[k][0]) for k in range(n)]  # cp:268 (ZeroDivisionError as in the reference)
    return (qint, rint, qual, nm)
