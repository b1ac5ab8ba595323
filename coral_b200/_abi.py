"""ctypes binding of the C ABI in include/coral_b200.h (the stub INTEGRATION.md shows to maintainers).

The library is built in-tree by ``coral_b200.build``.  There is no CPU fallback: every compute entry
point needs a CUDA device and raises ``CoralError`` otherwise.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CORAL_B200_LIB") or os.path.join(_HERE, "libcoral_b200.so")   # override: kernel tuning runs

CHR_IDX = {('chr%d' % (i + 1)): i for i in range(22)}
CHR_IDX.update({'chrX': 22, 'chrY': 23, 'chrM': 24})


class CoralError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("coral_b200 error %d: %s" % (code, msg))
        self.code = code


class Params(ctypes.Structure):
    _fields_ = [("min_del_len", ctypes.c_int32), ("min_bp_match_cutoff", ctypes.c_int32),
                ("min_mapq", ctypes.c_int32), ("gap_mapq", ctypes.c_int32), ("gap", ctypes.c_int32),
                ("bp_distance_cutoff", ctypes.c_int32)]


_lib = None


def lib():
    """Load libcoral_b200.so (fails loudly when it has not been built)."""
    global _lib
    if _lib is None:
        path = os.environ.get("CORAL_B200_LIB", LIB_PATH)      # development aid: a kernel variant built beside the library
        if not os.path.exists(path):
            raise ImportError("libcoral_b200.so is missing: run `python -m coral_b200.build` (no CPU fallback exists)")
        L = ctypes.CDLL(path)
        L.coral_last_error.restype = ctypes.c_char_p
        L.coral_n_records.restype = ctypes.c_int64
        L.coral_launch_count.restype = ctypes.c_uint64
        L.coral_stage_ms.restype = ctypes.c_float
        L.coral_stage_offset_ms.restype = ctypes.c_float
        L.coral_bam_n_records.restype = ctypes.c_int64
        L.coral_bam_ref_len.restype = ctypes.c_int64
        L.coral_bam_ref_name.restype = ctypes.c_char_p
        L.coral_bam_records.restype = ctypes.c_void_p
        L.coral_bam_offsets.restype = ctypes.c_void_p
        L.coral_ingest_ref_name.restype = ctypes.c_char_p
        L.coral_ingest_ref_len.restype = ctypes.c_int64
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def ingest_host_check(file_bytes, mode=0):
    """CPU run of the device-side ingest's decoder and framing steps (coral_ingest_host_check; test hook).
    -> dict(rec, rec_off, n_candidates, used_serial).  mode 1 forces the serial chase."""
    L = lib()
    fb = np.ascontiguousarray(file_bytes, dtype=np.uint8)
    n, nb, ncand, ser = ctypes.c_int64(0), ctypes.c_uint64(0), ctypes.c_int64(0), ctypes.c_int32(0)
    err = ctypes.create_string_buffer(512)
    rc = L.coral_ingest_host_check(_p(fb), ctypes.c_uint64(len(fb)), int(mode), None, ctypes.c_uint64(0), None, ctypes.c_int64(-1),
                                   ctypes.byref(n), ctypes.byref(nb), ctypes.byref(ncand), ctypes.byref(ser), err, 512)
    if rc not in (0, -6):          # -6 = CORAL_E_NOMEM: sizing call
        raise CoralError(rc, err.value.decode())
    rec = np.zeros(nb.value + 64, np.uint8)
    off = np.zeros(n.value + 1, np.uint64)
    rc = L.coral_ingest_host_check(_p(fb), ctypes.c_uint64(len(fb)), int(mode), _p(rec), ctypes.c_uint64(nb.value), _p(off),
                                   ctypes.c_int64(n.value), ctypes.byref(n), ctypes.byref(nb), ctypes.byref(ncand),
                                   ctypes.byref(ser), err, 512)
    if rc != 0:
        raise CoralError(rc, err.value.decode())
    return {"rec": rec[:nb.value], "rec_off": off, "n_candidates": int(ncand.value), "used_serial": int(ser.value)}


class DeviceBam(object):
    """Device-side ingest (csrc/ingest.cu): the .bam file's bytes are inflated and framed on the GPU of `ctx`, which
    afterwards holds the records exactly as after load_records of a BamFile.  The handle offers what the drop-in class
    needs from its BAM handle without a host copy of the records: references, n_records, query_names."""

    def __init__(self, path, ctx):
        self.path = path
        self._ctx = ctx
        self._bytes = np.fromfile(path, dtype=np.uint8)
        self.references, self.lengths = ctx.ingest_open(self._bytes)
        self.n_records = None
        self.rec = None
        self.rec_off = None
        self.rec_slack = 0

    def run(self):
        """References must have been set on the ctx (set_references) before."""
        self.n_records, self.framed_bytes = self._ctx.ingest_run()
        self._bytes = None
        return self.n_records

    def query_name(self, i):
        return self.query_names([i])[0]

    def query_names(self, idx):
        return self._ctx.read_names(idx)

    def close(self):
        self._bytes = None


class BamFile(object):
    """Host-side ingest: inflated, framed records in one (pinned) buffer."""

    def __init__(self, path, n_threads=0):
        L = lib()
        h = ctypes.c_void_p()
        err = ctypes.create_string_buffer(512)
        rc = L.coral_bam_read(path.encode(), int(n_threads), ctypes.byref(h), err, 512)
        if rc != 0:
            raise CoralError(rc, err.value.decode())
        self._h = h
        self.path = path
        n_ref = L.coral_bam_n_ref(h)
        self.references = tuple(L.coral_bam_ref_name(h, i).decode() for i in range(n_ref))
        self.lengths = tuple(int(L.coral_bam_ref_len(h, i)) for i in range(n_ref))
        self.n_records = int(L.coral_bam_n_records(h))
        nb = ctypes.c_uint64()
        ptr = L.coral_bam_records(h, ctypes.byref(nb))
        self.nbytes = int(nb.value)
        self.rec_ptr = ptr
        buf = (ctypes.c_uint8 * max(self.nbytes, 1)).from_address(ptr)
        self.rec = np.frombuffer(buf, dtype=np.uint8, count=self.nbytes)
        self.rec_slack = 64            # coral_bam_read allocates 64 bytes past the last record (page-locked buffer)
        obuf = (ctypes.c_uint64 * (self.n_records + 1)).from_address(L.coral_bam_offsets(h))
        self.rec_off = np.frombuffer(obuf, dtype=np.uint64, count=self.n_records + 1)

    def query_name(self, i):
        o = int(self.rec_off[i])
        ln = int(self.rec[o + 12])
        return self.rec[o + 36:o + 36 + ln - 1].tobytes().decode()

    def query_names(self, idx):
        from .records import query_names
        return query_names(self.rec, self.rec_off, idx)

    def close(self):
        if self._h is not None:
            self.rec = None
            self.rec_off = None
            lib().coral_bam_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


XCHG_CHANNELS = 8
XCHG_HANDLE_BYTES = 64
# channels of the multi-GPU engine (sharded.py) and their default slot sizes (largest message per rank)
CH_SA, CH_EXTRA, CH_CHIM, CH_SDEL, CH_CONC, CH_COV, CH_MISC = 0, 1, 2, 3, 4, 5, 6
XCHG_DEFAULT_SLOTS = (32 << 20, 8 << 20, 32 << 20, 1 << 20, 32 << 20, 1 << 20, 1 << 20, 0)


class Xchg(object):
    """One rank's exchange window (csrc/xchg.cu): create on every rank, all-gather the handles over any transport,
    connect, barrier, exchange ...; barrier again before close()."""

    def __init__(self, device, rank, world, slot_bytes=XCHG_DEFAULT_SLOTS):
        L = lib()
        L.coral_xchg_last_error.restype = ctypes.c_char_p
        slots = np.zeros(XCHG_CHANNELS, np.uint64)
        slots[:len(slot_bytes)] = slot_bytes
        h = ctypes.c_void_p()
        rc = L.coral_xchg_create(int(device), int(rank), int(world), _p(slots), ctypes.byref(h))
        if rc != 0:
            raise CoralError(rc, "coral_xchg_create failed")
        self._h, self._L = h, L
        self.rank, self.world, self.device = int(rank), int(world), int(device)

    def _check(self, rc):
        if rc != 0:
            raise CoralError(rc, self._L.coral_xchg_last_error(self._h).decode())

    def handle(self):
        """-> (64-byte IPC handle as bytes, device address of the local window)."""
        hb = np.zeros(XCHG_HANDLE_BYTES, np.uint8)
        ptr = ctypes.c_uint64()
        self._check(self._L.coral_xchg_handle(self._h, _p(hb), ctypes.byref(ptr)))
        return hb.tobytes(), int(ptr.value)

    def connect(self, handles, same_process_ptrs=None):
        """handles: list of `world` 64-byte handles in rank order; same_process_ptrs: window addresses of the ranks that
        are threads of this process (0 elsewhere)."""
        hb = np.frombuffer(b"".join(handles), np.uint8).copy()
        sp = None if same_process_ptrs is None else np.ascontiguousarray(same_process_ptrs, dtype=np.uint64)
        self._check(self._L.coral_xchg_connect(self._h, _p(hb), _p(sp)))

    def set_timeout(self, seconds):
        self._check(self._L.coral_xchg_set_timeout(self._h, ctypes.c_double(seconds)))

    def timeouts(self):
        n = ctypes.c_int32()
        self._check(self._L.coral_xchg_timeouts(self._h, ctypes.byref(n)))
        return int(n.value)

    def close(self):
        if self._h is not None:
            self._L.coral_xchg_destroy(self._h)
            self._h = None


class Context(object):
    """One coral_ctx bound to one GPU."""

    def __init__(self, device=0):
        L = lib()
        h = ctypes.c_void_p()
        rc = L.coral_create(int(device), ctypes.byref(h))
        if rc != 0:
            raise CoralError(rc, "coral_create failed (no CUDA device? the hot path has no CPU fallback)")
        self._h = h
        self._L = L
        self.references = ()
        self._keep = []
        self._pass_keep = None      # inside a pipelined pass: objects the C side writes at pass_end + finalizers
        self._pass_fin = []
        self._arena = []            # page-locked chunks [address, capacity, used] of pass_begin(pinned_results=True)
        self._pinned_results = False

    # ---- pipelined pass (include/coral_b200.h: coral_pass_begin / coral_pass_end) --------------------------------
    def pass_begin(self, pinned_results=False):
        """Until pass_end() the entry points only enqueue their kernels: the arrays they return are FILLED at pass_end()
        (do not read them before), counts are those of the last complete pass over the same inputs.
        pinned_results: the arrays returned inside this pass are views of a page-locked arena owned by the ctx (their
        copies are direct DMAs, nothing is staged or memcpy'd when the pass ends) and stay valid until the next
        pass_begin(pinned_results=True) on this ctx — for callers that consume a pass's results before the next pass."""
        self._check(self._L.coral_pass_begin(self._h))
        self._pass_keep, self._pass_fin = [], []
        self._pinned_results = bool(pinned_results)
        if pinned_results:
            for ch in self._arena:
                ch[2] = 0

    def _z(self, shape, dtype=np.float64):
        """Output array: numpy zeros, or — inside a pass with pinned_results — an (uninitialised) view of the arena."""
        if self._pass_keep is None or not getattr(self, "_pinned_results", False):
            return np.zeros(shape, dtype)
        dt = np.dtype(dtype)
        count = int(np.prod(shape)) if not isinstance(shape, (int, np.integer)) else int(shape)
        nbytes = (count * dt.itemsize + 63) & ~63
        for ch in self._arena:
            if ch[1] - ch[2] >= nbytes:
                break
        else:
            cap = max(nbytes, 8 << 20)
            ptr = ctypes.c_void_p()
            self._check(self._L.coral_host_alloc(ctypes.c_uint64(cap), ctypes.byref(ptr)))
            ch = [ptr.value, cap, 0]
            self._arena.append(ch)
        buf = (ctypes.c_uint8 * max(nbytes, 1)).from_address(ch[0] + ch[2])
        ch[2] += nbytes
        return np.frombuffer(buf, dtype=dt, count=count).reshape(shape)

    def pass_end(self):
        """-> True when the pass ran without intermediate synchronisation (sizes replayed from the previous pass)."""
        rp = ctypes.c_int32()
        try:
            self._check(self._L.coral_pass_end(self._h, ctypes.byref(rp)))
            for fn in self._pass_fin:
                fn()
        finally:
            self._pass_keep, self._pass_fin = None, []
        return bool(rp.value)

    def _hold(self, *objs):
        """Keep output buffers alive until pass_end (the C side holds their addresses)."""
        if self._pass_keep is not None:
            self._pass_keep.extend(objs)

    def _check(self, rc):
        if rc != 0:
            raise CoralError(rc, self._L.coral_last_error(self._h).decode())

    def close(self):
        if self._h is not None:
            self._L.coral_destroy(self._h)
            self._h = None
            for ch in self._arena:
                self._L.coral_host_free(ctypes.c_void_p(ch[0]))
            self._arena = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_params(self, **kw):
        p = Params()
        self._L.coral_default_params(ctypes.byref(p))
        for k, v in kw.items():
            setattr(p, k, int(v))
        self._check(self._L.coral_set_params(self._h, ctypes.byref(p)))

    def set_references(self, names):
        self.references = tuple(names)
        blob = b"".join(n.encode() + b"\x00" for n in names)
        rank = np.array([CHR_IDX.get(n, -1) for n in names], dtype=np.int32)
        self._check(self._L.coral_set_references(self._h, blob, len(names), _p(rank)))

    def load_records(self, rec, rec_off, global_ordinal=None, placement="copy", slack=0):
        """rec: uint8 array (host) of framed records; rec_off: uint64[n+1].
        placement "copy": records are copied to HBM; "mapped": they stay in (page-locked) host memory and the
        kernels read them in place over PCIe — `rec` must then extend >= 64 bytes past rec_off[-1] and stay alive;
        "mapped_async": as "mapped", but the call returns after the pre-scan (what fetch needs) while the full scan keeps
        running; entry points that need it join it (wait_scan); "mapped_lazy": only the pre-scan runs, the full scan
        runs inside the first entry point that needs it."""
        rec = np.ascontiguousarray(rec, dtype=np.uint8)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        n = len(rec_off) - 1
        go = None if global_ordinal is None else np.ascontiguousarray(global_ordinal, dtype=np.uint64)
        mode = {"copy": 0, "mapped": 2, "mapped_async": 3, "mapped_lazy": 4}[placement]
        nbytes = int(rec_off[-1]) if n >= 0 and len(rec_off) else 0
        if mode >= 2:
            if rec.nbytes + int(slack) < nbytes + 64:     # `slack`: bytes the caller guarantees behind the view
                raise ValueError("mapped placement needs 64 bytes of slack after the last record")
            self._keep = [rec]
        self._check(self._L.coral_load_records(self._h, _p(rec), ctypes.c_uint64(nbytes), _p(rec_off), ctypes.c_int64(n),
                                               _p(go), mode))

    def wait_scan(self):
        """Join the scan a "mapped_async" load left running (no-op otherwise); its format errors surface here."""
        self._check(self._L.coral_wait_scan(self._h))

    def load_records_device(self, dev_ptr, nbytes, rec_off, global_ordinal=None):
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        go = None if global_ordinal is None else np.ascontiguousarray(global_ordinal, dtype=np.uint64)
        self._check(self._L.coral_load_records(self._h, ctypes.c_void_p(dev_ptr), ctypes.c_uint64(nbytes), _p(rec_off),
                                               ctypes.c_int64(len(rec_off) - 1), _p(go), 1))

    # ---- multi-GPU exchange of fetch (include/coral_b200.h) ---------------------------------------------------
    def record_keys(self):
        n = self.n_records
        out = {"name_lo": self._z(n, np.uint64), "name_hi": self._z(n, np.uint64), "sa_off": self._z(n, np.uint32),
               "sa_len": self._z(n, np.uint32), "nm": self._z(n, np.int32)}
        self._check(self._L.coral_get_record_keys(self._h, _p(out["name_lo"]), _p(out["name_hi"]), _p(out["sa_off"]),
                                                  _p(out["sa_len"]), _p(out["nm"])))
        return out

    def nm_partials(self):
        s1 = ctypes.c_double(); s2 = ctypes.c_double(); nn = ctypes.c_int64(); zd = ctypes.c_int64()
        self._check(self._L.coral_nm_partials(self._h, ctypes.byref(s1), ctypes.byref(s2), ctypes.byref(nn), ctypes.byref(zd)))
        return s1.value, s2.value, nn.value, zd.value

    def load_rows(self, text, text_off, name_lo, name_hi, flag, mapq, l_seq, nm):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        text_off = np.ascontiguousarray(text_off, dtype=np.uint64)
        a = [np.ascontiguousarray(name_lo, np.uint64), np.ascontiguousarray(name_hi, np.uint64),
             np.ascontiguousarray(flag, np.uint32), np.ascontiguousarray(mapq, np.uint32),
             np.ascontiguousarray(l_seq, np.uint32), np.ascontiguousarray(nm, np.int32)]
        n = len(text_off) - 1
        self._check(self._L.coral_load_rows(self._h, ctypes.c_int64(n), _p(text), ctypes.c_uint64(text.nbytes), _p(text_off),
                                            *[_p(x) for x in a]))

    # ---- packed device-side exchange (csrc/exchange.cu, csrc/stab.cu); addresses are device addresses ----------
    def pack_sa_rows(self):
        """-> (device address, bytes) of message 1 of the fetch exchange."""
        ptr = ctypes.c_uint64(); nb = ctypes.c_uint64()
        self._check(self._L.coral_pack_sa_rows(self._h, ctypes.byref(ptr), ctypes.byref(nb)))
        return int(ptr.value), int(nb.value)

    def pack_extra_rows(self, buf_ptr, starts):
        """Gathered messages 1 (device buffer + per-rank start offsets) -> (device address, bytes) of message 2."""
        st = np.ascontiguousarray(starts, dtype=np.uint64)
        ptr = ctypes.c_uint64(); nb = ctypes.c_uint64()
        self._check(self._L.coral_pack_extra_rows(self._h, len(st), ctypes.c_void_p(buf_ptr), _p(st), ctypes.byref(ptr),
                                                  ctypes.byref(nb)))
        return int(ptr.value), int(nb.value)

    def unpack_rows(self, buf1_ptr, starts1, buf2_ptr, starts2):
        """Both gathered message sets -> this (rows-only) ctx.  -> (n_rows, nm partial sums per rank: 4 arrays)."""
        s1 = np.ascontiguousarray(starts1, dtype=np.uint64); s2 = np.ascontiguousarray(starts2, dtype=np.uint64)
        w = len(s1)
        a = self._z(w, np.float64); b = self._z(w, np.float64); cn = self._z(w, np.int64); zd = self._z(w, np.int64)
        n = ctypes.c_int64()
        self._check(self._L.coral_unpack_rows(self._h, w, ctypes.c_void_p(buf1_ptr), _p(s1), ctypes.c_void_p(buf2_ptr), _p(s2),
                                              _p(a), _p(b), _p(cn), _p(zd), ctypes.byref(n)))
        return int(n.value), (a, b, cn, zd)

    def concordant_pack(self):
        """-> (device address, bytes) of the (edge, name hash, mask) message of the last concordant() call."""
        ptr = ctypes.c_uint64(); nb = ctypes.c_uint64()
        self._check(self._L.coral_concordant_pack(self._h, ctypes.byref(ptr), ctypes.byref(nb)))
        return int(ptr.value), int(nb.value)

    def count_name_msgs(self, n_edges, buf_ptr, starts, sizes, rb_ptr, n_rb):
        st = np.ascontiguousarray(starts, dtype=np.uint64); sz = np.ascontiguousarray(sizes, dtype=np.uint64)
        cnt = self._z(max(n_edges, 1), np.int64)
        self._check(self._L.coral_count_name_msgs(self._h, int(n_edges), len(st), ctypes.c_void_p(buf_ptr), _p(st), _p(sz),
                                                  ctypes.c_void_p(rb_ptr), ctypes.c_int64(n_rb), _p(cnt)))
        self._hold(cnt)
        return cnt[:n_edges]

    # ---- exchanges over peer memory (csrc/xchg.cu) and the distributed fetch (csrc/chimeric_merge.cu) -------------
    def xchg_allgather(self, xchg, channel, ptr, nbytes, lane=False):
        """All-gather of one device message per rank -> (gathered device buffer address, starts, sizes)."""
        w = xchg.world
        buf = ctypes.c_uint64(); st = self._z(w, np.uint64); sz = self._z(w, np.uint64)
        self._check(self._L.coral_xchg_allgather(self._h, xchg._h, int(channel), ctypes.c_uint64(ptr), ctypes.c_uint64(nbytes),
                                                 int(bool(lane)), ctypes.byref(buf), _p(st), _p(sz)))
        return int(buf.value), st, sz

    def xchg_allreduce(self, xchg, channel, values):
        v = np.ascontiguousarray(values, dtype=np.int64).copy()
        self._check(self._L.coral_xchg_allreduce_i64(self._h, xchg._h, int(channel), _p(v), len(v)))
        return v

    def region_coverage_sharded(self, xchg, channel, tid, start, stop):
        """region_coverage summed over all shards of the file (every rank calls it with the same regions)."""
        tid = np.ascontiguousarray(tid, dtype=np.int32)
        start = np.ascontiguousarray(start, dtype=np.int32)
        stop = np.ascontiguousarray(stop, dtype=np.int32)
        n = len(tid)
        bs = self._z(n, np.int64)
        rc = self._z(n, np.int64)
        self._check(self._L.coral_region_coverage_sharded(self._h, xchg._h, int(channel), _p(tid), _p(start), _p(stop), n,
                                                          _p(bs), _p(rc)))
        self._hold(bs, rc)
        return bs, rc

    def pack_chimeric(self):
        ptr = ctypes.c_uint64(); nb = ctypes.c_uint64()
        self._check(self._L.coral_pack_chimeric(self._h, ctypes.byref(ptr), ctypes.byref(nb)))
        return int(ptr.value), int(nb.value)

    def merge_chimeric(self, buf_ptr, starts):
        """Gathered partial chimeric tables -> this ctx's table in the 1-GPU order.  -> (n_reads, n_alignments)."""
        st = np.ascontiguousarray(starts, dtype=np.uint64)
        nr = ctypes.c_int64(); na = ctypes.c_int64()
        self._check(self._L.coral_merge_chimeric(self._h, len(st), ctypes.c_void_p(buf_ptr), _p(st), ctypes.byref(nr),
                                                 ctypes.byref(na)))
        return int(nr.value), int(na.value)

    def smalldel_pack(self):
        ptr = ctypes.c_uint64(); nb = ctypes.c_uint64()
        self._check(self._L.coral_smalldel_pack(self._h, ctypes.byref(ptr), ctypes.byref(nb)))
        return int(ptr.value), int(nb.value)

    def read_device(self, ptr, nbytes, lane=False):
        """Device bytes -> numpy uint8 array (filled at pass_end inside a pipelined pass)."""
        out = self._z(int(nbytes), np.uint8)
        if nbytes:
            self._check(self._L.coral_read_device(self._h, ctypes.c_uint64(ptr), ctypes.c_uint64(nbytes), _p(out), int(bool(lane))))
        self._hold(out)
        return out

    def stream_handle(self, which=0):
        self._L.coral_stream_handle.restype = ctypes.c_uint64
        return int(self._L.coral_stream_handle(self._h, int(which)))

    def adopt_stream(self, handle):
        self._check(self._L.coral_adopt_stream(self._h, ctypes.c_uint64(handle)))

    def device_view(self):
        """Device addresses (recs, offsets, name_lo, name_hi, flag_mapq, l_seq, nm, sa_off, sa_len)."""
        ptrs = (ctypes.c_uint64 * 9)()
        self._check(self._L.coral_device_view(self._h, ptrs, 9))
        return [int(x) for x in ptrs]

    def set_option(self, name, value):
        self._check(self._L.coral_set_option(self._h, name.encode(), ctypes.c_int64(int(value))))

    def rescan(self):
        self._check(self._L.coral_rescan(self._h))

    def stage_ms(self, stage):
        return float(self._L.coral_stage_ms(self._h, int(stage)))

    def stage_ms_all(self):
        """Device time of every stage's most recent run (float32[10], < 0 where a stage has not run), one call."""
        out = np.zeros(10, np.float32)
        self._check(self._L.coral_stage_ms_all(self._h, _p(out)))
        return out

    def reset_stage_times(self):
        self._L.coral_reset_stage_times(self._h)

    def stage_offset_ms(self, stage_a, which_a, stage_b, which_b):
        return float(self._L.coral_stage_offset_ms(self._h, int(stage_a), int(which_a), int(stage_b), int(which_b)))

    @property
    def n_records(self):
        return int(self._L.coral_n_records(self._h))

    @property
    def launch_count(self):
        return int(self._L.coral_launch_count(self._h))

    def record_table(self):
        n = self.n_records
        cols = {k: self._z(n, np.int32) for k in ("tid", "pos", "endpos", "flag", "mapq", "n_cigar")}
        cols.update({k: self._z(n, np.int64) for k in ("l_seq", "infer_len", "aligned_bases", "non_acgt")})
        self._check(self._L.coral_get_record_table(
            self._h, _p(cols["tid"]), _p(cols["pos"]), _p(cols["endpos"]), _p(cols["flag"]), _p(cols["mapq"]),
            _p(cols["l_seq"]), _p(cols["n_cigar"]), _p(cols["infer_len"]), _p(cols["aligned_bases"]), _p(cols["non_acgt"])))
        return cols

    def region_coverage(self, tid, start, stop):
        """Regions [start, stop) sorted by (tid, start), disjoint -> (base_sum int64[n], rec_count int64[n])."""
        tid = np.ascontiguousarray(tid, dtype=np.int32)
        start = np.ascontiguousarray(start, dtype=np.int32)
        stop = np.ascontiguousarray(stop, dtype=np.int32)
        n = len(tid)
        bs = self._z(n, np.int64)
        rc = self._z(n, np.int64)
        self._check(self._L.coral_region_coverage(self._h, _p(tid), _p(start), _p(stop), n, _p(bs), _p(rc)))
        self._hold(bs, rc)
        return bs, rc

    def region_coverage_static(self, tid, start, stop):
        """coral_region_coverage_static: the same result through the entry point that allocates nothing (fixed-size static
        device scratch; at most CORAL_STATIC_MAX_REGIONS regions, see include/coral_b200.h)."""
        tid = np.ascontiguousarray(tid, dtype=np.int32)
        start = np.ascontiguousarray(start, dtype=np.int32)
        stop = np.ascontiguousarray(stop, dtype=np.int32)
        n = len(tid)
        bs = self._z(n, np.int64)
        rc = self._z(n, np.int64)
        self._check(self._L.coral_region_coverage_static(self._h, _p(tid), _p(start), _p(stop), n, _p(bs), _p(rc)))
        return bs, rc

    def fetch(self, read_back=True):
        """coral_fetch + coral_get_chimeric -> dict of numpy arrays (see include/coral_b200.h).
        read_back=False: the table stays on the device (counts and nm sums only)."""
        nr = ctypes.c_int64(); na = ctypes.c_int64(); s1 = ctypes.c_double(); s2 = ctypes.c_double()
        nn = ctypes.c_int64(); zd = ctypes.c_int64()
        self._check(self._L.coral_fetch(self._h, ctypes.byref(nr), ctypes.byref(na), ctypes.byref(s1), ctypes.byref(s2),
                                        ctypes.byref(nn), ctypes.byref(zd)))
        nr, na = nr.value, na.value
        if self._pass_keep is not None:          # the nm sums are written at pass_end
            self._hold(s1, s2, nn, zd)
            self._pass_fin.append(lambda: out.update(nm_sum=s1.value, nm_sumsq=s2.value, nm_n=nn.value, nm_zero_div=zd.value))
        out = {"n_reads": nr, "n_aln": na, "nm_sum": s1.value, "nm_sumsq": s2.value, "nm_n": nn.value,
               "nm_zero_div": zd.value}
        if not read_back:
            return out
        return self.get_chimeric(nr, na, out)

    def get_chimeric(self, nr, na, out=None):
        """coral_get_chimeric of the table on the device (after fetch / merge_chimeric) -> dict of numpy arrays."""
        out = {} if out is None else out
        if self._pass_keep is not None and self._pinned_results:
            # one page-locked block, one packing kernel, one copy (coral_get_chimeric_packed): the arrays are views of it
            cols = (("read_first_rec", np.uint32, nr), ("read_len", np.int64, nr), ("read_status", np.int32, nr),
                    ("aln_ptr", np.uint32, nr + 1), ("qs", np.int32, na), ("qe", np.int32, na), ("tid", np.int32, na),
                    ("rstart", np.int32, na), ("rend", np.int32, na), ("strand", np.uint8, na), ("mapq", np.int32, na),
                    ("nm", np.int64, na))
            total = sum((np.dtype(dt).itemsize * n + 63) & ~63 for _, dt, n in cols)
            blk = self._z(max(total, 64), np.uint8)
            offs = np.zeros(12, np.uint64)
            self._check(self._L.coral_get_chimeric_packed(self._h, _p(blk), ctypes.c_uint64(blk.nbytes), _p(offs)))
            out.update({"n_reads": nr, "n_aln": na})
            for (name, dt, n), o in zip(cols, offs.tolist()):
                out[name] = blk[int(o):int(o) + np.dtype(dt).itemsize * n].view(dt)
            self._hold(out, blk)
            return out
        out.update({"n_reads": nr, "n_aln": na,
               "read_first_rec": self._z(nr, np.uint32), "read_len": self._z(nr, np.int64),
               "read_status": self._z(nr, np.int32), "aln_ptr": self._z(nr + 1, np.uint32)})
        for k in ("qs", "qe", "tid", "rstart", "rend", "mapq"):
            out[k] = self._z(na, np.int32)
        out["strand"] = self._z(na, np.uint8)
        out["nm"] = self._z(na, np.int64)
        self._check(self._L.coral_get_chimeric(
            self._h, _p(out["read_first_rec"]), _p(out["read_len"]), _p(out["read_status"]), _p(out["aln_ptr"]),
            _p(out["qs"]), _p(out["qe"]), _p(out["tid"]), _p(out["rstart"]), _p(out["rend"]), _p(out["strand"]),
            _p(out["mapq"]), _p(out["nm"])))
        self._hold(out)
        return out

    @staticmethod
    def _ivals(intervals_tid, start, end_incl):
        return (np.ascontiguousarray(intervals_tid, dtype=np.int32), np.ascontiguousarray(start, dtype=np.int32),
                np.ascontiguousarray(end_incl, dtype=np.int32))

    def smalldel(self, tid, start, end_incl):
        t, s, e = self._ivals(tid, start, end_incl)
        n = ctypes.c_int64()
        self._check(self._L.coral_smalldel(self._h, _p(t), _p(s), _p(e), len(t), ctypes.byref(n)))
        n = n.value
        out = {"n": n, "rec": self._z(n, np.uint32)}
        for k in ("ai", "next_start", "prev_end", "first_start", "last_end", "mapq"):
            out[k] = self._z(n, np.int32)
        if n:
            self._check(self._L.coral_get_smalldel(self._h, _p(out["rec"]), _p(out["ai"]), _p(out["next_start"]),
                                                   _p(out["prev_end"]), _p(out["first_start"]), _p(out["last_end"]),
                                                   _p(out["mapq"])))
        self._hold(out)
        return out

    def candidates_l(self, tid, start, end_incl):
        t, s, e = self._ivals(tid, start, end_incl)
        n = ctypes.c_int64()
        self._check(self._L.coral_candidates_l(self._h, _p(t), _p(s), _p(e), len(t), ctypes.byref(n)))
        n = n.value
        out = {"n": n, "read": self._z(n, np.uint32)}
        for k in ("i", "j", "chr1", "pos1", "chr2", "pos2", "rgap", "mq1", "mq2"):
            out[k] = self._z(n, np.int32)
        for k in ("o1", "o2", "flag"):
            out[k] = self._z(n, np.uint8)
        if n:
            self._check(self._L.coral_get_candidates(
                self._h, _p(out["read"]), _p(out["i"]), _p(out["j"]), _p(out["chr1"]), _p(out["pos1"]), _p(out["o1"]),
                _p(out["chr2"]), _p(out["pos2"]), _p(out["o2"]), _p(out["rgap"]), _p(out["flag"]), _p(out["mq1"]),
                _p(out["mq2"])))
        self._hold(out)
        return out

    def cluster_call(self, chr1, pos1, o1, chr2, pos2, o2, rgap, flag, mq1, mq2, min_cluster_size, bp_distance_cutoff,
                     peel_cutoff, match_cutoff, max_peels=0):
        i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
        u8 = lambda a: np.ascontiguousarray(a, dtype=np.uint8)
        chr1, pos1, chr2, pos2, rgap, mq1, mq2 = map(i32, (chr1, pos1, chr2, pos2, rgap, mq1, mq2))
        o1, o2, flag = map(u8, (o1, o2, flag))
        n = len(chr1)
        nc = ctypes.c_int64()
        self._check(self._L.coral_cluster_call(
            self._h, ctypes.c_int64(n), _p(chr1), _p(pos1), _p(o1), _p(chr2), _p(pos2), _p(o2), _p(rgap), _p(flag),
            _p(mq1), _p(mq2), ctypes.c_double(min_cluster_size), ctypes.c_int32(bp_distance_cutoff),
            ctypes.c_double(peel_cutoff), ctypes.c_int32(match_cutoff), ctypes.c_int32(max_peels), ctypes.byref(nc)))
        return self._clusters_out(n, nc.value)

    def _clusters_out(self, n, nc):
        out = {"n_clusters": nc, "order": self._z(n, np.uint32), "cluster_ptr": self._z(nc + 1, np.uint32),
               "n_sub": self._z(nc, np.int32), "sub_of": self._z(n, np.int32), "row_pos1": self._z(n, np.int32),
               "row_pos2": self._z(n, np.int32), "row_nb": self._z(n, np.int32), "row_first": self._z(n, np.uint32),
               "row_stats": self._z((n, 6), np.float64)}
        if n:
            self._check(self._L.coral_get_clusters(
                self._h, _p(out["order"]), _p(out["cluster_ptr"]), _p(out["n_sub"]), _p(out["sub_of"]), _p(out["row_pos1"]),
                _p(out["row_pos2"]), _p(out["row_nb"]), _p(out["row_first"]), _p(out["row_stats"])))
        self._hold(out)
        return out

    def cluster_call_candidates(self, n, min_cluster_size, bp_distance_cutoff, peel_cutoff, match_cutoff, max_peels=0):
        """cluster_call on the candidate table the last candidates_l / candidates_pair left on the device (`n` = its size):
        the candidates do not travel through the host."""
        nc = ctypes.c_int64()
        self._check(self._L.coral_cluster_call_candidates(
            self._h, ctypes.c_double(min_cluster_size), ctypes.c_int32(bp_distance_cutoff), ctypes.c_double(peel_cutoff),
            ctypes.c_int32(match_cutoff), ctypes.c_int32(max_peels), ctypes.byref(nc)))
        return self._clusters_out(int(n), nc.value)

    def bp_match(self, a1, a4, minus1, minus2, rgap, b1, b4, cutoff1, cutoff2):
        """coral_bp_match over n pairs -> uint8[n]."""
        i64 = lambda a: np.ascontiguousarray(a, dtype=np.int64)
        u8 = lambda a: np.ascontiguousarray(a, dtype=np.uint8)
        a1, a4, b1, b4 = map(i64, (a1, a4, b1, b4))
        minus1, minus2 = u8(minus1), u8(minus2)
        rgap = np.ascontiguousarray(rgap, dtype=np.float64)
        n = len(a1)
        out = self._z(n, np.uint8)
        self._check(self._L.coral_bp_match(self._h, ctypes.c_int64(n), _p(a1), _p(a4), _p(minus1), _p(minus2), _p(rgap), _p(b1),
                                           _p(b4), ctypes.c_int64(int(cutoff1)), ctypes.c_int64(int(cutoff2)), _p(out)))
        return out

    def concordant_rows_only(self, tid, p):
        """The (edge, record, mask) rows of this ctx's records stay on the device (concordant_pack); no count."""
        tid = np.ascontiguousarray(tid, dtype=np.int32)
        p = np.ascontiguousarray(p, dtype=np.int32)
        nr = ctypes.c_int64()
        self._check(self._L.coral_concordant(self._h, _p(tid), _p(p), len(tid), ctypes.c_int64(0), None, None, None,
                                             ctypes.byref(nr)))
        self._stab_n = int(nr.value)
        return self._stab_n

    def concordant_rows(self):
        """(edge, record, mask) rows of the last concordant / concordant_rows_only call, read back to the host."""
        nr = self._stab_n
        rows = {"edge": self._z(nr, np.uint32), "rec": self._z(nr, np.uint32), "mask": self._z(nr, np.uint32)}
        if nr:
            self._check(self._L.coral_get_concordant_rows(self._h, _p(rows["edge"]), _p(rows["rec"]), _p(rows["mask"])))
        self._hold(rows)
        return rows

    def concordant(self, tid, p, rb_edge, rb_rec, fetch_rows=True):
        """Edges sorted by (tid, p).  -> (lr_count int64[n], rows dict(edge, rec, mask))."""
        tid = np.ascontiguousarray(tid, dtype=np.int32)
        p = np.ascontiguousarray(p, dtype=np.int32)
        rb_edge = np.ascontiguousarray(rb_edge, dtype=np.uint32)
        rb_rec = np.ascontiguousarray(rb_rec, dtype=np.uint32)
        n = len(tid)
        cnt = self._z(n, np.int64)
        nr = ctypes.c_int64()
        self._check(self._L.coral_concordant(self._h, _p(tid), _p(p), n, ctypes.c_int64(len(rb_edge)), _p(rb_edge), _p(rb_rec),
                                             _p(cnt), ctypes.byref(nr)))
        self._stab_n = nr = nr.value
        self._hold(cnt)
        if fetch_rows:
            return cnt, self.concordant_rows()
        return cnt, {"edge": self._z(nr, np.uint32), "rec": self._z(nr, np.uint32), "mask": self._z(nr, np.uint32)}

    def cn_segment_index(self, seg_tid, seg_begin, seg_end, seg_idx, n_aln):
        """hash_alignment_to_seg lookups (ibg:181-210) for every alignment of the chimeric table -> (lo, hi) int32."""
        i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
        seg_tid, seg_begin, seg_end, seg_idx = map(i32, (seg_tid, seg_begin, seg_end, seg_idx))
        lo, hi = self._z(n_aln, np.int32), self._z(n_aln, np.int32)
        self._check(self._L.coral_cn_segment_index(self._h, _p(seg_tid), _p(seg_begin), _p(seg_end), _p(seg_idx),
                                                   ctypes.c_int32(len(seg_tid)), _p(lo), _p(hi)))
        return lo, hi

    def ingest_open(self, file_bytes):
        """Device-side ingest, step 1 (host only): -> (reference names, lengths).  `file_bytes` must stay alive until
        ingest_run."""
        self._ingest_bytes = np.ascontiguousarray(file_bytes, dtype=np.uint8)
        self._check(self._L.coral_ingest_open(self._h, _p(self._ingest_bytes), ctypes.c_uint64(len(self._ingest_bytes))))
        n = self._L.coral_ingest_n_ref(self._h)
        return (tuple(self._L.coral_ingest_ref_name(self._h, i).decode() for i in range(n)),
                tuple(int(self._L.coral_ingest_ref_len(self._h, i)) for i in range(n)))

    def ingest_run(self):
        """Step 2: inflate + frame on the device, then the load-time scan.  -> (n_records, framed bytes)."""
        n, nb = ctypes.c_int64(0), ctypes.c_uint64(0)
        try:
            self._check(self._L.coral_ingest_run(self._h, ctypes.byref(n), ctypes.byref(nb)))
        finally:
            self._ingest_bytes = None
        return int(n.value), int(nb.value)

    def ingest_used_serial_chase(self):
        return int(self._L.coral_ingest_used_serial_chase(self._h))

    def read_names(self, idx):
        """Read names of records `idx` (any placement; the only way to them after a device-side ingest)."""
        idx = np.ascontiguousarray(idx, dtype=np.uint32)
        if len(idx) == 0:
            return []
        out = self._z(len(idx) * 256, np.uint8)
        self._check(self._L.coral_get_read_names(self._h, _p(idx), ctypes.c_int64(len(idx)), _p(out)))
        rows = out.reshape(len(idx), 256)
        ln = (rows != 0).argmin(axis=1)          # first NUL (names are at most 254 characters + NUL)
        return [rows[k, :ln[k]].tobytes().decode() for k in range(len(idx))]

    def alignment_paths(self, iv_tid, iv_start, iv_stop, se_tid, se_l, se_r, name_rec=(), name_lo=(), name_hi=(),
                        min_overlap=500, min_mapq=20):
        """Concordant-read pass of compute_path_constraints (ibg:1303-1310).  Intervals [start, stop) in list order,
        sequence edges [l, r] sorted by (tid, l) and disjoint, the name set through records carrying the names and / or
        128-bit name hashes.  -> dict(n, rec, iv, first, last): rows in (interval, file) order."""
        i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
        iv_tid, iv_start, iv_stop, se_tid, se_l, se_r = map(i32, (iv_tid, iv_start, iv_stop, se_tid, se_l, se_r))
        name_rec = np.ascontiguousarray(name_rec, dtype=np.uint32)
        name_lo = np.ascontiguousarray(name_lo, dtype=np.uint64)
        name_hi = np.ascontiguousarray(name_hi, dtype=np.uint64)
        assert len(name_lo) == len(name_hi) and len(iv_tid) == len(iv_start) == len(iv_stop) and len(se_tid) == len(se_l) == len(se_r)
        n = ctypes.c_int64()
        self._check(self._L.coral_alignment_paths(
            self._h, _p(iv_tid), _p(iv_start), _p(iv_stop), ctypes.c_int32(len(iv_tid)), _p(se_tid), _p(se_l), _p(se_r),
            ctypes.c_int32(len(se_tid)), _p(name_rec), ctypes.c_int64(len(name_rec)), _p(name_lo), _p(name_hi),
            ctypes.c_int64(len(name_lo)), ctypes.c_int32(min_overlap), ctypes.c_int32(min_mapq), ctypes.byref(n)))
        n = int(n.value)
        out = {"n": n, "rec": self._z(n, np.uint32), "iv": self._z(n, np.int32), "first": self._z(n, np.int32),
               "last": self._z(n, np.int32)}
        if n:
            self._check(self._L.coral_get_alignment_paths(self._h, _p(out["rec"]), _p(out["iv"]), _p(out["first"]), _p(out["last"])))
        return out

    def sa_text(self, text, ent_ptr, ent_beg, ent_end, read_len):
        n_reads = len(read_len)
        ne = int(ent_ptr[-1])
        out = {"status": self._z(n_reads, np.int32)}
        for k in ("qs", "qe", "al", "rstart", "rend", "nm"):
            out[k] = self._z(ne, np.int64)
        out["strand"] = self._z(ne, np.uint8); out["mapq"] = self._z(ne, np.int32); out["src"] = self._z(ne, np.uint32)
        self._check(self._L.coral_sa_text(
            self._h, text, ctypes.c_uint64(len(text)), ctypes.c_int64(n_reads), _p(ent_ptr), _p(ent_beg), _p(ent_end),
            _p(read_len), _p(out["status"]), _p(out["qs"]), _p(out["qe"]), _p(out["al"]), _p(out["rstart"]), _p(out["rend"]),
            _p(out["strand"]), _p(out["mapq"]), _p(out["nm"]), _p(out["src"])))
        return out

    def _get_candidates(self, n):
        out = {"n": n, "read": self._z(n, np.uint32)}
        for k in ("i", "j", "chr1", "pos1", "chr2", "pos2", "rgap", "mq1", "mq2"):
            out[k] = self._z(n, np.int32)
        for k in ("o1", "o2", "flag"):
            out[k] = self._z(n, np.uint8)
        if n:
            self._check(self._L.coral_get_candidates(
                self._h, _p(out["read"]), _p(out["i"]), _p(out["j"]), _p(out["chr1"]), _p(out["pos1"]), _p(out["o1"]),
                _p(out["chr2"]), _p(out["pos2"]), _p(out["o2"]), _p(out["rgap"]), _p(out["flag"]), _p(out["mq1"]),
                _p(out["mq2"])))
        self._hold(out)
        return out

    def candidates_pair(self, reads, tid2, start2, end2):
        reads = np.ascontiguousarray(reads, dtype=np.uint32)
        t, s, e = self._ivals(tid2, start2, end2)
        assert len(t) == 2
        n = ctypes.c_int64()
        self._check(self._L.coral_candidates_pair(self._h, _p(reads), ctypes.c_int64(len(reads)), _p(t), _p(s), _p(e),
                                                  ctypes.byref(n)))
        return self._get_candidates(n.value)

    def set_chimeric(self, aln_ptr, qs, qe, tid, rstart, rend, strand, mapq):
        aln_ptr = np.ascontiguousarray(aln_ptr, dtype=np.uint32)
        i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
        qs, qe, tid, rstart, rend, mapq = map(i32, (qs, qe, tid, rstart, rend, mapq))
        strand = np.ascontiguousarray(strand, dtype=np.uint8)
        self._check(self._L.coral_set_chimeric(self._h, ctypes.c_int64(len(aln_ptr) - 1), _p(aln_ptr), _p(qs), _p(qe), _p(tid),
                                               _p(rstart), _p(rend), _p(strand), _p(mapq)))
