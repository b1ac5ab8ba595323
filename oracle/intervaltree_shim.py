"""oracle/intervaltree_shim.py — TEST INFRASTRUCTURE.  Stand-in for the un-vendored ``intervaltree``
package as used by the reference (``infer_breakpoint_graph.py:88`` ``IntervalTree()``, ``:91``
``tree[a:b] = idx`` half-open, ``:178`` ``tree[pos]`` -> set of objects with ``.data``)."""
import bisect
from collections import namedtuple

Interval = namedtuple("Interval", ["begin", "end", "data"])


class IntervalTree(object):
    def __init__(self):
        self._begins = []
        self._ivs = []
        self._maxlen = 0

    def __setitem__(self, key, data):
        if not isinstance(key, slice):
            raise TypeError("IntervalTree keys must be slices")
        b, e = key.start, key.stop
        if b >= e:
            raise ValueError("IntervalTree: Null Interval objects not allowed in IntervalTree: Interval(%r, %r, %r)" % (b, e, data))
        i = bisect.bisect_right(self._begins, b)
        self._begins.insert(i, b)
        self._ivs.insert(i, Interval(b, e, data))
        self._maxlen = max(self._maxlen, e - b)

    def __getitem__(self, pos):
        if isinstance(pos, slice):
            lo, hi = pos.start, pos.stop
            return set(iv for iv in self._ivs if iv.begin < hi and iv.end > lo)
        out = set()
        i = bisect.bisect_right(self._begins, pos) - 1
        while i >= 0 and self._begins[i] >= pos - self._maxlen:
            iv = self._ivs[i]
            if iv.end > pos:
                out.add(iv)
            i -= 1
        return out

    def __len__(self):
        return len(self._ivs)
