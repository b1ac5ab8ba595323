"""bench.py — CoRAL reconstruct hot path (breakpoint + coverage pass) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scale S]

Workload (config.workload): BASELINE.json configs[1] — synthetic multi-segment ecDNA, 3 seeds over 6 Mb,
CN 40, 50x ONT 30 kb reads (sim/configs.py::c2_multi_segment, scale 1.0).  One "step" = one pass of the hot
path over that batch of alignment records:
  record scan -> normal-coverage regions -> fetch (SA parse / chimeric table) -> small-del rows -> cluster+peel
  -> alignment2bp_l candidates -> cluster+peel -> sequence-edge coverage -> concordant-edge support,
every result table copied back to the host.
  value : records/s with the records resident in HBM when the timed region starts (device pass through the C ABI)
  e2e   : records/s through the public drop-in class (bam_to_breakpoint_nanopore) starting from a pinned HOST
          record buffer: H2D of the records, all kernels, D2H of all tables, Python materialisation, graph build
  roofline : the record-scan kernel (the only kernel that touches record bytes): algorithmic bytes / CUDA-event time
  cpu_baseline : the UNMODIFIED reference (oracle/_ref under the oracle's pysam / intervaltree stand-ins; the CPU port when
          that copy is absent) on a bounded sample, 1 core
`--impl reference` times the same reference on the GPU arm's own full-size input ("reference-on-shim": pysam is not
installable on this box).  --gpus N > 1: one process per GPU (torchrun), the file's shards exchange over peer memory
(csrc/xchg.cu; --exchange nccl = the collective baseline); --workload C4 / C5: one file cut into N shards.
"""
import argparse
import os as _os
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # streams of a ctx (scan, lane) must not share a hardware queue
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "long reads/sec & aligned Gbases/sec (breakpoint+coverage pass), % HBM roofline"
CPU_SAMPLE_SCALE = 0.06


def record_field_bytes(rec, rec_off):
    """Algorithmic bytes of the record scan: every record field but QUAL read once + 76 B of table written
    (sum(36 + l_name + 4 n_cigar + ceil(l_seq / 2) + aux))."""
    off = rec_off[:-1].astype(np.int64)

    def u32(o):
        return (rec[off + o].astype(np.int64) | (rec[off + o + 1].astype(np.int64) << 8) |
                (rec[off + o + 2].astype(np.int64) << 16) | (rec[off + o + 3].astype(np.int64) << 24))
    bs = u32(0)
    l_name = rec[off + 12].astype(np.int64)
    n_cig = rec[off + 16].astype(np.int64) | (rec[off + 17].astype(np.int64) << 8)
    l_seq = u32(20)
    aux = bs - 32 - l_name - 4 * n_cig - (l_seq + 1) // 2 - l_seq
    return int((36 + l_name + 4 * n_cig + (l_seq + 1) // 2 + aux).sum()) + 76 * len(off)


class ClockSampler(threading.Thread):
    def __init__(self, gpu=0):
        super().__init__(daemon=True)
        self.gpu = gpu
        self.rows = []
        self.stop_flag = False
        self.source = "none"

    def run(self):
        """Clocks and throttle reasons while the timed regions run.  NVML in-process (what nvidia-smi itself queries):
        spawning nvidia-smi five times a second attaches a new driver client each time, which can stall the
        cudaMalloc / cudaFree calls of the pass being timed; nvidia-smi is the fallback when pynvml is unusable."""
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self._nvml_index())
            bits = [0x8, 0x40, 0x20, 0x4]      # hw_slowdown, hw_thermal_slowdown, sw_thermal_slowdown, sw_power_cap
            sm_max = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            nv.nvmlDeviceGetCurrentClocksEventReasons(h)
        except Exception:
            return self._run_nvidia_smi()
        self.source = "nvml"
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                why = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                self.rows.append([str(sm), str(sm_max)] + ["Active" if why & b else "Not Active" for b in bits])
            except Exception:
                pass
            time.sleep(float(os.environ.get("CORAL_BENCH_SAMPLER_S", "0.02")))

    def _nvml_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        ids = [x for x in vis.split(",") if x.strip() != ""]
        return int(ids[self.gpu]) if ids and all(x.strip().isdigit() for x in ids) and self.gpu < len(ids) else self.gpu

    def _run_nvidia_smi(self):
        self.source = "nvidia-smi"
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(1.0)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 2 + k and r[2 + k] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": int(self.rows[0][1]) if self.rows[0][1].isdigit() else None, "reasons": reasons,
                "samples": len(self.rows), "source": self.source}


def hbm_peak():
    """Roofline denominator: the driver-measured copy bandwidth of this pool's B200s, else the recipe's fallback."""
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json (of measured)"
    except Exception:
        return 6650.0, "fallback 6650 (of fallback)"


def measured_traffic(workload, scale, world):
    """dram__bytes_read + write of ONE k_scan_records launch from the committed ncu capture of this workload
    (profiles/r2_scan_traffic.json: written from `ncu --set full` reports by profiles/summarize.py), or None."""
    if world != 1 or scale != 1.0:
        return None, None
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r2_scan_traffic.json")))[workload]
        return float(t["dram_bytes"]), "profiles/%s (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of one launch)" % t["file"]
    except Exception:
        return None, None


def write_inputs(cfg, outdir):
    os.makedirs(outdir, exist_ok=True)
    cns = os.path.join(outdir, "cn_seg.bed")
    with open(cns, "w") as fp:
        for (c, s, e, cn) in cfg.cn_rows:
            fp.write("%s\t%d\t%d\t%g\n" % (c, s, e, cn))
    seed = os.path.join(outdir, "seed.bed")
    with open(seed, "w") as fp:
        for (c, s, e) in cfg.seeds:
            fp.write("%s\t%d\t%d\n" % (c, s, e))
    return cns, seed


# ------------------------------------------------------------------------------------------------------------
WORKLOADS = {
    # --workload: (sim config, label)   C2 = BASELINE.json configs[1], the headline; the others are extra lines for rows
    # the C2 records never reach (a3 small deletions: C3; the non-ACGT slow path of coverage: C1n), kept in profiles/
    "C2": ("C2", {}, "C2 multi-segment ecDNA: 3 seeds/6 Mb, CN 40, 50x ONT 30 kb"),
    "C3": ("C3", {}, "C3 small-deletion-heavy amplicon: 60x HiFi 15 kb, 300 in-CIGAR deletions, dense D ops"),
    "C1": ("C1", {}, "C1 single ecDNA: 2 Mb amplicon, 4 SVs, 30x ONT 20 kb"),
    "C1n": ("full_c1n", {}, "C1 single ecDNA with 0.1 % N bases (non-ACGT slow path of coverage)"),
    # BASELINE.json configs[3] / [4] on the sub-samples the unmodified reference could be run on in the build container
    # (tests/golden/full_digest.py: sub_c4, sub_c5 — committed reference digests).  With --gpus N > 1 ONE file is cut into N
    # contiguous genomic-bin shards (strong scaling): chimeric reads cross the shard cuts at every junction
    "C4": ("sub_c4", {}, "C4 complex amplicon, 200 breakpoints / 4 chromosomes, ONT 50 kb, sub-sample at 25x of the named 100x"),
    "C5": ("sub_c5", {}, "C5 whole-genome shape on chr19-22, 6 seed intervals, 30x ONT 20 kb, 0.5 % chimeric noise reads (sub-sample)"),
}
STRONG = ("C4", "C5")          # workloads that are one file cut into N shards when N > 1


def workload_config(scale, seed_shift=0, workload="C2"):
    from sim import configs
    name, over, _ = WORKLOADS[workload]
    if name.startswith("sub_") or name.startswith("full_"):      # the very configs the committed reference digests were made on
        sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
        import full_digest as FD
        if name.startswith("sub_") and scale != 1.0:
            raise SystemExit("--workload %s is a fixed sub-sample (its reference digest is committed): use --scale 1.0" % workload)
        cname, kw, over, _ = FD.FULL_CASES[name]
        cfg = configs.CONFIGS[cname](**dict(kw, scale=kw.get("scale", 1.0) * scale))
        for k, v in over.items():
            setattr(cfg, k, v)
        return cfg
    if name == "C2":
        cfg = configs.c2_multi_segment(scale=scale, seed=20240215 + 2 + 1000 * seed_shift)
    else:
        cfg = configs.CONFIGS[name](scale=scale)
    for k, v in over.items():
        setattr(cfg, k, v)
    return cfg


def config_dict(a, world, n_rec, nbytes, aligned, n_bp, alg_bytes):
    """`config` of the JSON line — the same dict from both arms when they ran the same input."""
    return {"workload": WORKLOADS[a.workload][2], "scale": a.scale,
            "records_per_gpu": int(n_rec), "record_bytes_per_gpu": int(nbytes), "aligned_gbases_per_gpu": aligned / 1e9,
            "breakpoints_found": int(n_bp),
            "sharding": ("single GPU" if world == 1 else
                         "ONE coordinate-sorted file cut into N contiguous genomic-bin shards balanced by decoded bytes (strong "
                         "scaling): chimeric reads cross the cuts; exchange of chimeric fetch rows / parsed chimeric tables / "
                         "small-del rows / concordant rows and integer sums of coverage inside every step" if a.workload in STRONG else
                         "rank r = r-th contiguous genomic-bin shard (its own 3 chromosomes, C2-shaped) of one "
                         "24-contig file; NCCL exchange of chimeric fetch rows / small-del rows / concordant "
                         "rows and all-reduce of coverage sums inside every step"),
            "l2_policy": "inputs (%.2f GB of record fields per pass) larger than the 126 MB L2" % (alg_bytes / 1e9)}


def _reference_pass(kind, bam, cns, seed):
    """One pass of the CPU arm in THIS process (a forked child: the reference keeps its containers as class
    attributes, infer_breakpoint_graph.py:22-61, so every pass needs a fresh interpreter state).
    kind "reference": the unmodified reference (oracle/_ref or /root/reference) under the oracle's pysam / intervaltree
    stand-ins; "port": oracle/coral_oracle.py + the host steps.  -> {stage seconds}, summary."""
    import hashlib
    t = {}
    if kind == "reference":
        from oracle import ref_runner
        st = ref_runner.run(bam, cns, seed, stop_after="assign_cov")
        t = dict(st["timing_s"])
        graphs = st["stages"]["assign_cov"]["graphs"]
        n_bp = len(st["stages"]["find_breakpoints"]["new_bp_list"])
        txt = [g["graph_txt"] for g in graphs]
    else:
        import util
        t0 = time.perf_counter()
        eng = util.oracle_engine(bam, seed)
        t["open"] = time.perf_counter() - t0
        for name, fn, args in (("read_cns", eng.read_cns, (cns,)), ("fetch", eng.fetch, ()),
                               ("hash_alignment_to_seg", eng.hash_alignment_to_seg, ()),
                               ("find_amplicon_intervals", eng.find_amplicon_intervals, ()),
                               ("find_smalldel_breakpoints", eng.find_smalldel_breakpoints, ()),
                               ("find_breakpoints", eng.find_breakpoints, ()), ("build_graph", eng.build_graph, ()),
                               ("assign_cov", eng.assign_cov, ())):
            t0 = time.perf_counter()
            fn(*args)
            t[name] = time.perf_counter() - t0
        n_bp = len(eng.new_bp_list)
        from coral_b200.breakpoint_graph import output_breakpoint_graph_lr
        import tempfile
        txt = []
        for g in eng.lr_graph:
            fd, pth = tempfile.mkstemp(suffix=".txt")
            os.close(fd)
            output_breakpoint_graph_lr(g, pth)
            txt.append(open(pth).read())
            os.unlink(pth)
    return t, {"breakpoints": n_bp, "graph_txt_sha1": [hashlib.sha1(x.encode()).hexdigest() for x in txt]}


def _forked_pass(kind, bam, cns, seed):
    """_reference_pass in a forked child (copy-on-write view of the decoded records) -> (stage seconds, summary)."""
    r, w = os.pipe()
    pid = os.fork()
    if pid == 0:
        code = 1
        try:
            os.close(r)
            res = _reference_pass(kind, bam, cns, seed)
            with os.fdopen(w, "w") as fp:
                json.dump(res, fp)
            code = 0
        except BaseException:       # noqa
            import traceback
            traceback.print_exc()
        finally:
            os._exit(code)
    os.close(w)
    with os.fdopen(r) as fp:
        txt = fp.read()
    _, status = os.waitpid(pid, 0)
    if status != 0 or not txt:
        raise RuntimeError("reference pass failed (status %d)" % status)
    return json.loads(txt)


REF_STAGES = ("open", "read_cns", "fetch", "hash_alignment_to_seg", "find_amplicon_intervals", "find_smalldel_breakpoints",
              "find_breakpoints", "build_graph", "assign_cov")


def cpu_arm(a, world, scale, max_steps, budget_s):
    """The reference's own CPU implementation of the path on this box's host cores, from decoded records in host memory
    (where the GPU arm's e2e starts) through graph + coverage: the SAME stage set as the GPU arm's e2e.  One step = one
    full pass over the workload at `scale`; as many steps (<= max_steps) as fit in budget_s after the first."""
    from sim.generate import simulate
    from oracle import pysam_shim, make_ref
    import tempfile
    kind = "reference" if make_ref.ref_src() else "port"
    if world > 1:
        from sim import configs
        cfg = configs.c2_shard(0, world, scale=scale)[0]          # rank 0's shard: C2-shaped, its own cn rows / seeds
    else:
        cfg = workload_config(scale, workload=a.workload)
    data = simulate(cfg)
    n_rec, nbytes = int(data.n_records), int(data.rec_off[-1])
    alg_bytes = record_field_bytes(data.rec, data.rec_off)
    cns, seed = write_inputs(cfg, tempfile.mkdtemp(prefix="coral_ref_"))
    bam = pysam_shim.register_memory_bam("bench", cfg.refs, data.rec[:nbytes].tobytes(), data.rec_off)
    del data
    fh = pysam_shim.AlignmentFile(bam)
    aligned = int(fh._mb.sum())
    del fh
    steps, tot, stages, summ = 0, 0.0, {}, None
    t_begin = time.perf_counter()
    while steps < max_steps:
        st, summ = _forked_pass(kind, bam, cns, seed)
        dt = sum(st[k] for k in REF_STAGES)
        tot += dt
        steps += 1
        stages = st
        if time.perf_counter() - t_begin + dt > budget_s:
            break
    return {"kind": kind, "n_rec": n_rec, "nbytes": nbytes, "aligned": aligned, "alg_bytes": alg_bytes, "steps": steps,
            "seconds": tot, "stage_s": stages, "summary": summ, "scale": scale}


def reference_arm(a, rank, world):
    """`--impl reference`: the unmodified reference (kind "reference"; the CPU port when oracle/_ref is absent) on the GPU
    arm's own input — C2 at scale 1.0, the same simulated records.  One pass takes minutes of single-threaded Python,
    so the arm runs as many of the requested steps as fit in ~4 minutes (at least one) without warm-up, and says so."""
    if rank != 0:
        return
    r = cpu_arm(a, world, a.scale, max_steps=max(1, a.steps), budget_s=float(os.environ.get("CORAL_REF_BUDGET_S", "240")))
    v = r["n_rec"] * r["steps"] / r["seconds"]
    sample = ("the GPU arm's own input (%s at scale %g: %d records, %.3f aligned Gbases%s), decoded records in host memory "
              "-> graph + coverage, all stages of the GPU arm's e2e; %d of the %d requested steps (one pass is minutes of "
              "single-threaded Python), no warm-up" % (a.workload, r["scale"], r["n_rec"], r["aligned"] / 1e9,
                                                      "; rank 0's shard of the weak-scaled file" if world > 1 else "",
                                                      r["steps"], a.steps))
    out = {"impl": "reference", "metric": METRIC, "value": v, "unit": "records/s", "n_gpus": world, "steps": r["steps"],
           "warmup": 0, "steps_requested": a.steps, "warmup_requested": a.warmup,
           "ms_per_step": r["seconds"] / r["steps"] * 1e3, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "u8/int32/int64 (f64 stats)", "data": "synthetic",
           "config": config_dict(a, world, r["n_rec"], r["nbytes"], r["aligned"], r["summary"]["breakpoints"], r["alg_bytes"]),
           "aligned_gbases_per_s": r["aligned"] * r["steps"] / r["seconds"] / 1e9,
           "result": r["summary"],
           "cpu_baseline": {"value": v, "unit": "records/s", "cores": 1, "kind": r["kind"], "sample": sample,
                            "stage_s": r["stage_s"],
                            "note": ("the unmodified reference (oracle/_ref, byte-identical copy of /root/reference/src) under "
                                     "the oracle's pysam / intervaltree stand-ins: pysam is not installable on this box, its inner "
                                     "loops are C in the stand-in too (oracle/covlib.c)" if r["kind"] == "reference" else
                                     "CPU port of the reference path (oracle/coral_oracle.py): oracle/_ref is absent")},
           "e2e": {"value": v, "unit": "records/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


# ------------------------------------------------------------------------------------------------------------
def build_plan(eng, cn_seg):
    """Run the public pipeline once and capture the inputs of every device call (fixed for the timed passes)."""
    from coral_b200 import infer_breakpoint_graph  # noqa: F401
    eng.read_cns(cn_seg)
    segs, total = eng._median_cn_segments()
    segs = sorted(segs, key=lambda s: (eng._tid[s[0]], s[1]))
    plan = {"normal": ([eng._tid[s[0]] for s in segs], [s[1] for s in segs], [s[2] + 1 for s in segs])}
    eng.fetch()
    eng.hash_alignment_to_seg()
    eng.find_amplicon_intervals()
    plan["intervals"] = eng._interval_arrays()
    eng.find_smalldel_breakpoints()
    plan["smalldel_cand"] = eng._smalldel_candidates()[0]
    eng.find_breakpoints()
    eng.build_graph()
    batches, uniq, owners, rbps = eng._edge_queries()
    rb_edge, rb_rec = eng._rbps_rows(rbps)
    b = batches[0]
    plan["seq_edges"] = ([r[0] for r in b], [r[1] for r in b], [r[2] for r in b])
    plan["conc"] = (np.array([u[0] for u in uniq], np.int32), np.array([u[1] for u in uniq], np.int32),
                    np.array(rb_edge, np.uint32), np.array(rb_rec, np.uint32))
    plan["uniq"], plan["rbps"] = uniq, rbps
    plan["cutoff"] = eng.min_cluster_cutoff
    eng.assign_cov()
    return plan


def device_pass(ctx, plan, gpu_ops, pipelined=True):
    """One pass of the hot path on device-resident records; every result table is read back.
    pipelined: as ONE stream of kernels (coral_pass_begin / coral_pass_end): the calls below only enqueue, the result
    arrays are filled by one deferred copy each when the pass ends with its single synchronisation; launch sizes that
    depend on device results are those of the previous pass over the same records, checked on the device."""
    if pipelined:
        ctx.pass_begin(pinned_results=True)     # result arrays in page-locked memory: direct DMA, valid until the next pass
    ctx.rescan()
    out = {}
    # first what only needs the columns of the pre-scan (core fields, name hash, SA / NM): in a pipelined pass these
    # kernels run next to the full scan, which is joined by the first call below that reads its CIGAR / SEQ columns
    out["fetch"] = ctx.fetch()
    cand = ctx.candidates_l(*plan["intervals"])
    out["cand"] = cand
    if cand["n"]:
        out["cl2"] = ctx.cluster_call_candidates(cand["n"], plan["cutoff"], 2000, plan["cutoff"], 100)
    out["normal"] = ctx.region_coverage(*plan["normal"])
    out["smalldel"] = ctx.smalldel(*plan["intervals"])
    c = plan["smalldel_cand"]
    if len(c["chr1"]):
        out["cl1"] = gpu_ops.cluster_arrays(ctx, c, plan["cutoff"], 2000, plan["cutoff"], 100)
    out["seq"] = ctx.region_coverage(*plan["seq_edges"])
    out["conc"] = ctx.concordant(*plan["conc"])
    if pipelined:
        out["replayed"] = ctx.pass_end()
    return out


SHARD_MS = {}


def sharded_device_pass(eng, plan, gpu_ops, pipelined=False):
    """The same pass on one shard of the file, with the exchanges of SURVEY.md §8e inside (all ranks in lockstep).
    pipelined (needs the peer-memory transport): the whole pass — packing kernels, exchanges over NVLink, the rank's
    share of fetch, merges, coverage, concordant counts — is enqueued as one stream of kernels per lane on both
    contexts (records ctx + rows-only ctx) and synchronised once."""
    ctx, cctx = eng._ctx, eng._cctx
    out = {}
    t = [time.perf_counter()]

    def lap(name):
        t.append(time.perf_counter())
        SHARD_MS[name] = SHARD_MS.get(name, 0.0) + (t[-1] - t[-2]) * 1e3
    if pipelined:
        ctx.pass_begin(pinned_results=True); cctx.pass_begin(pinned_results=True)
    ctx.rescan(); lap("scan")
    # the chimeric chain first: inside a pipelined pass it runs on the second lane next to the stages below
    out["fetch"], _ = eng._run_fetch(); lap("fetch_exchange")
    cand = cctx.candidates_l(*plan["intervals"])
    out["cand"] = cand
    if cand["n"]:
        out["cl2"] = cctx.cluster_call_candidates(cand["n"], plan["cutoff"], 2000, plan["cutoff"], 100)
    lap("candidates+cluster")
    out["normal"] = eng._region_counts(*plan["normal"]); lap("normal_cov+sum")
    out["smalldel"] = eng._smalldel_gather(*plan["intervals"]); lap("smalldel+gather")
    c = plan["smalldel_cand"]
    if len(c["chr1"]):
        out["cl1"] = gpu_ops.cluster_arrays(ctx, c, plan["cutoff"], 2000, plan["cutoff"], 100)
    out["seq"] = eng._region_counts(*plan["seq_edges"]); lap("seq_cov+sum")
    cnt = eng._concordant_counts(plan["uniq"], plan["rbps"])[0]        # local stab rows, combined by name hash
    out["conc"] = cnt; lap("concordant+gather")
    if pipelined:
        r1 = cctx.pass_end()
        r2 = ctx.pass_end()
        out["replayed"] = bool(r1 and r2)
        lap("pass_end")
    return out


def result_bytes(out):
    n = 0
    for v in out.values():
        items = v.values() if isinstance(v, dict) else (v if isinstance(v, (tuple, list)) else [v])
        for x in items:
            if isinstance(x, np.ndarray):
                n += x.nbytes
            elif isinstance(x, dict):
                n += sum(y.nbytes for y in x.values() if isinstance(y, np.ndarray))
    return n


FILE_SAMPLE_SCALE = 0.1


def e2e_from_file(a, device, scale):
    """`.bam` file bytes -> graph + coverage through the public class, both ingest forms, on a sample of the workload
    written as a real BGZF BAM: (1) device_ingest — the compressed bytes cross PCIe, BGZF inflate and record framing run
    on the GPU (csrc/ingest.cu), the records never exist in host memory; (2) host reader — zlib on all host threads,
    records framed into a page-locked buffer, kernels read them in place (what `e2e` above starts from).  The timed
    region of each pass starts with the path of the file and ends after closebam().  The inflate share of the reference
    arm (pysam / htslib, one thread) is `cpu_baseline.stage_s.open` resp. the reference line's `stage_s`."""
    import tempfile
    import numpy as np
    from sim.generate import simulate, write_dataset
    from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
    from coral_b200.host_steps import gc_paused
    cfg = workload_config(scale, workload=a.workload)
    data = simulate(cfg)
    paths = write_dataset(data, tempfile.mkdtemp(prefix="coral_file_"))
    bam_bytes = os.path.getsize(paths["bam"])
    res = {"sample": "%s at scale %g (%d records, %.2f GB decoded) as a BGZF BAM of %.2f GB" %
                     (a.workload, scale, data.n_records, int(data.rec_off[-1]) / 1e9, bam_bytes / 1e9),
           "records": int(data.n_records), "bam_bytes": int(bam_bytes), "decoded_bytes": int(data.rec_off[-1])}
    digests = {}
    for kind, kw in (("device_ingest", dict(placement="device_ingest")), ("host_reader", dict(placement="auto"))):
        ms, stage = [], {}
        for it in range(4):
            with gc_paused():
                t0 = time.perf_counter()
                eng = bam_to_breakpoint_nanopore(paths["bam"], paths["seed"], device=device, **kw)
                t1 = time.perf_counter()
                eng.read_cns(paths["cn_seg"]); eng.fetch(); eng.hash_alignment_to_seg(); eng.find_amplicon_intervals()
                eng.find_smalldel_breakpoints(); eng.find_breakpoints(); eng.build_graph(); eng.assign_cov()
                txt = [g.graph_text() if hasattr(g, "graph_text") else None for g in eng.lr_graph]
                n_bp = len(eng.new_bp_list)
                cov = eng.normal_cov
                eng.closebam()
                del eng
                t2 = time.perf_counter()
            if it:                     # the first pass warms the allocator / page-locks
                ms.append((t2 - t0) * 1e3)
                stage = {"open+ingest+scan": (t1 - t0) * 1e3, "pipeline": (t2 - t1) * 1e3}
        digests[kind] = (n_bp, cov, txt)
        res[kind] = {"ms_per_pass": [round(x, 2) for x in ms], "records_per_s": data.n_records / (float(np.mean(ms)) * 1e-3),
                     "bam_GB_per_s": bam_bytes / (float(np.mean(ms)) * 1e6), "last_pass_ms": stage}
    res["identical_results"] = digests["device_ingest"] == digests["host_reader"]
    res["h2d_bytes_per_pass"] = {"device_ingest": int(bam_bytes), "host_reader": "the record fields the scan pulls over PCIe (QUAL never crosses)"}
    return res


DIGEST_OF = {"C2": "full_c2", "C3": "full_c3", "C1": "full_c1", "C1n": "full_c1n", "C4": "sub_c4", "C5": "sub_c5"}


def parity_check(a, rank, world, local, make_engine, cns, comm):
    """The parity bit of the JSON line (outside every timed region; a mismatch makes the run exit non-zero).
    N = 1: the whole pipeline through the public class on the bench workload itself, digested
    (tests/golden/full_digest.py) and compared with the committed digest of what the UNMODIFIED reference produced on
    the same input (tests/golden/full_c2.json.gz, made by tests/golden/make_full_digests.py): breakpoints, support sets,
    chimeric table, coverage, byte-identical graph.txt.
    N > 1, over the real process group (NCCL): (1) two reference-generated fixtures with cross-shard chimeric reads on
    every junction (c4_small, c5_mini) cut into N genomic-bin shards -> every stage equal to the fixture on rank 0 and
    identical on all ranks; (2) the weak-scaled bench workload: every rank's own amplicon graph (its shard through the
    1-GPU class) must appear byte-identical among the graphs of the sharded run."""
    import gzip
    import hashlib
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import full_digest as FD
    import util
    from coral_b200.host_steps import gc_paused
    if world == 1:
        case = DIGEST_OF.get(a.workload)
        path = os.path.join(ROOT, "tests", "golden", "%s.json.gz" % case)
        if a.scale != 1.0 or case is None or not os.path.exists(path):
            return {"parity_ok": None, "parity": "no committed reference digest for this workload / scale"}
        with gzip.open(path, "rt") as fp:
            gold = json.load(fp)
        with gc_paused():
            eng = make_engine(a.e2e_placement)
            st = util.run_pipeline(eng, cns, defer_cov=True)
            eng.closebam()
        bad = FD.compare_digests(FD.digest_state(st), gold["stages"])
        return {"parity_ok": not bad,
                "parity": {"against": "tests/golden/%s.json.gz (digest of the unmodified reference's state on this input)" % case,
                           "stages_compared": sorted(gold["stages"]), "mismatches": [list(b) for b in bad],
                           "graph_txt_sha1": [hashlib.sha1(g["graph_txt"].encode()).hexdigest()
                                              for g in st["assign_cov"]["graphs"]]}}
    import torch
    import torch.distributed as dist
    if a.workload in STRONG:
        # the file of this workload is the one the committed digest of the UNMODIFIED reference was made on: the sharded
        # run over the real process group must reproduce it (rank 0 compares; every rank must agree with rank 0)
        case = DIGEST_OF[a.workload]
        with gzip.open(os.path.join(ROOT, "tests", "golden", "%s.json.gz" % case), "rt") as fp:
            gold = json.load(fp)
        with gc_paused():
            eng = make_engine("copy")
            st = util.run_pipeline(eng, cns)
            eng.closebam()
        dig = FD.digest_state(st)
        bad = FD.compare_digests(dig, gold["stages"])
        dig_cmp = json.loads(json.dumps(dig))
        dig_cmp.get("fetch", {}).pop("nm_stats", None)         # f64 sums in rank order: within 1e-9, compared above on rank 0
        h = hashlib.sha1(json.dumps(dig_cmp, sort_keys=True).encode()).digest()[:8]
        t = torch.tensor([int.from_bytes(h, "little") >> 1, len(bad)], dtype=torch.int64, device="cuda")
        lo, hi = t.clone(), t.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool(lo[0] == hi[0])
        return {"parity_ok": int(hi[1]) == 0 and same,
                "parity": {"against": "tests/golden/%s.json.gz (digest of the unmodified reference's state on this file), sharded run "
                                      "over the real process group" % case, "stages_compared": sorted(gold["stages"]),
                           "mismatches": [list(b) for b in bad], "identical_on_all_ranks": same,
                           "graph_txt_sha1": [hashlib.sha1(g["graph_txt"].encode()).hexdigest() for g in st["assign_cov"]["graphs"]]}}
    import make_golden
    import test_golden_cpu as G
    from sim.generate import simulate
    from coral_b200 import sharded
    from coral_b200.records import RecordSet
    import tempfile
    res, ok = {}, True
    for case in ("c4_small", "c5_mini"):
        cfg = make_golden.make_cfg(case)
        data = simulate(cfg)
        recs = RecordSet([c for c, _ in cfg.refs], data.rec, data.rec_off)
        c_cns, c_seed = write_inputs(cfg, tempfile.mkdtemp(prefix="coral_parity_%d_" % rank))
        eng = sharded.make_sharded_engine(recs, c_seed, comm, device=local, xchg=make_engine.xchg)
        st = util.jsonify(util.run_pipeline(eng, c_cns, stop_after="assign_cov"))
        eng.closebam()
        want = {k: v for k, v in G.load_golden(case)["stages"].items() if k != "compute_path_constraints"}
        bad = G.compare_states(st, want)
        h = hashlib.sha1(json.dumps(st["assign_cov"], sort_keys=True).encode()).digest()[:8]
        t = torch.tensor([int.from_bytes(h, "little") >> 1, len(bad)], dtype=torch.int64, device="cuda")
        lo, hi = t.clone(), t.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool(lo[0] == hi[0])
        res[case] = {"mismatches_max_over_ranks": int(hi[1]), "identical_on_all_ranks": same, "records": int(data.n_records)}
        ok = ok and int(hi[1]) == 0 and same
    # the bench workload: this rank's shard alone through the 1-GPU class against the sharded run
    from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
    from sim import configs
    own = configs.c2_shard(rank, world, scale=a.scale)[0]
    o_cns, o_seed = write_inputs(own, tempfile.mkdtemp(prefix="coral_parity_own_%d_" % rank))
    with gc_paused():
        eng = make_engine("copy")
        st = util.run_pipeline(eng, cns, stop_after="assign_cov")
        eng.closebam()
        solo = bam_to_breakpoint_nanopore(None, o_seed, device=local, records=eng_records(make_engine), placement="copy")
        st1 = util.run_pipeline(solo, o_cns, stop_after="assign_cov")
        solo.closebam()
    all_txt = set(g["graph_txt"] for g in st["assign_cov"]["graphs"])
    mine = [g["graph_txt"] for g in st1["assign_cov"]["graphs"]]
    t = torch.tensor([int(len(mine) > 0 and all(x in all_txt for x in mine)), len(st["assign_cov"]["graphs"])],
                     dtype=torch.int64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    res["bench_workload"] = {"every_ranks_own_graph_found_in_the_sharded_run": bool(t[0]), "graphs": int(t[1])}
    ok = ok and bool(t[0])
    return {"parity_ok": ok, "parity": res}


def eng_records(make_engine):
    return make_engine.records


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--scale", type=float, default=float(os.environ.get("CORAL_BENCH_SCALE", "1.0")))
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS),
                    help="C2 = BASELINE.json configs[1] (the headline, default); C3 / C1 / C1n: extra lines for profiles/ "
                         "(small-deletion stage, non-ACGT coverage path); N=1 only")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (development aid)")
    ap.add_argument("--no-pipeline", action="store_true",
                    help="device pass call by call (a synchronisation inside every entry point) instead of as one pipelined "
                         "pass (development aid: what the pass mode buys)")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1: the exchanges of the sharded pass as kernels over peer memory (NVLink, csrc/xchg.cu; the pass "
                         "is then one stream of kernels per rank) or as torch.distributed / NCCL collectives issued by the "
                         "host between kernels (the baseline this is measured against)")
    ap.add_argument("--overlap-scan", action="store_true", help="N > 1: start the exchange chain next to the record scan (measured slower; development aid)")
    ap.add_argument("--e2e-placement", default="mapped_async", choices=["mapped", "mapped_async", "copy"],
                    help="e2e arm: kernels read the pinned host records in place over PCIe (mapped; mapped_async = fetch and "
                         "the host steps after it run next to the scan, the default at N=1: 172 vs 200 ms per C2 pass) or after a full "
                         "H2D copy")
    ap.add_argument("--no-file", action="store_true", help="skip the e2e_from_file leg (development aid)")
    ap.add_argument("--only", default="", help="'scan': time the record-scan kernel alone; 'e2e': per-pass stage times of "
                                               "the e2e arm alone (development aids)")
    ap.add_argument("--cprofile", action="store_true", help="cProfile of 20 device passes on rank 0 (development aid)")
    ap.add_argument("--profile", action="store_true",
                    help="bracket ONE device pass with cudaProfilerStart/Stop and exit (for ncu --profile-from-start off)")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        reference_arm(a, rank, world)
        return
    if world > 1 and a.e2e_placement == "mapped_async":
        a.e2e_placement = "mapped"
    warm = max(a.warmup, 3)

    import torch
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)

    from sim.generate import simulate
    from coral_b200 import _abi, gpu_ops
    from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
    from coral_b200.records import RecordSet
    from coral_b200.host_steps import gc_paused
    import tempfile

    comm = xchg = None
    if world > 1:
        # weak scaling: rank r holds a C2-shaped sample on its own three chromosomes of one 24-contig file, i.e. the
        # r-th contiguous genomic-bin shard of ONE coordinate-sorted BAM (sim/configs.py::c2_shard)
        import types
        from sim import configs
        from coral_b200 import sharded
        if a.workload in STRONG:
            cfg = cfg_files = workload_config(a.scale, workload=a.workload)     # every rank simulates the file, keeps its shard
        elif a.workload == "C2":
            cfg, all_cn, all_seeds = configs.c2_shard(rank, world, scale=a.scale)
            cfg_files = types.SimpleNamespace(cn_rows=all_cn, seeds=all_seeds)
        else:
            raise SystemExit("--workload %s runs on one GPU only" % a.workload)
        comm = sharded.TorchComm(device=torch.device("cuda", local))
        xchg = sharded.make_xchg(comm, local) if a.exchange == "peer" else None
    else:
        cfg = workload_config(a.scale, workload=a.workload)
        cfg_files = cfg
    pinned = {}

    def alloc(nbytes):
        pinned["t"] = torch.empty(nbytes + 64, dtype=torch.uint8).pin_memory()   # 64 B of slack: mapped placement
        return pinned["t"].numpy()[:nbytes]

    strong = world > 1 and a.workload in STRONG
    if strong:
        # one file, N contiguous shards balanced by decoded bytes (sharded.shard_ranges): this rank's records go into its
        # own page-locked buffer, re-based offsets; the whole file only exists while it is being cut
        whole = simulate(cfg)
        off = np.asarray(whole.rec_off, dtype=np.uint64)
        r0, r1 = sharded.shard_ranges(off, world)[rank]
        b0, b1 = int(off[r0]), int(off[r1])
        buf = alloc(b1 - b0)
        buf[:] = whole.rec[b0:b1]
        data = types.SimpleNamespace(rec=buf, rec_off=(off[r0:r1 + 1] - np.uint64(b0)).astype(np.uint64), n_records=r1 - r0)
        del whole
    else:
        data = simulate(cfg, alloc=alloc)
    n_rec = data.n_records
    nbytes = int(data.rec_off[-1])
    recs = RecordSet([c for c, _ in cfg.refs], data.rec, data.rec_off, rec_slack=64)
    cns, seed = write_inputs(cfg_files, tempfile.mkdtemp(prefix="coral_bench_%d_" % rank))

    def make_engine(placement):
        if world > 1:
            return sharded.make_sharded_engine(recs, seed, comm, device=local, placement=placement, presharded=True, xchg=xchg)
        return bam_to_breakpoint_nanopore(None, seed, device=local, records=recs, placement=placement)
    make_engine.records = recs
    make_engine.xchg = xchg
    alg_bytes = record_field_bytes(data.rec, data.rec_off)

    if a.only == "scan":
        ctx = _abi.Context(local)
        ctx.set_references([c for c, _ in cfg.refs])
        ctx.load_records(data.rec, data.rec_off)
        if os.environ.get("CORAL_SCAN_ORDER"):          # development aid: the record queue's order (0 file, 1 big first, 2 by size)
            ctx.set_option("scan_order", int(os.environ["CORAL_SCAN_ORDER"]))
        for _ in range(warm):
            ctx.rescan()
        ms = []
        for _ in range(a.steps):
            ctx.rescan()
            ms.append(ctx.stage_ms(0))
        k_ms = float(np.mean(ms))
        peak, peak_src = hbm_peak()
        print(json.dumps({"only": "scan", "workload": a.workload, "records": n_rec, "alg_bytes": alg_bytes, "record_bytes": nbytes,
                          "kernel_ms": k_ms, "min_ms": float(np.min(ms)), "GBps": alg_bytes / k_ms / 1e6,
                          "frac": alg_bytes / k_ms / 1e6 / peak, "peak": peak, "peak_source": peak_src}), flush=True)
        return

    e2e_stage = {}

    def e2e_pass():
        with gc_paused():           # as reconstruct_graph() does
            eng = _e2e_pass()
            t = time.perf_counter()
            del eng                 # released by reference count inside the timed region
            e2e_stage["release"] = (time.perf_counter() - t) * 1e3

    def _e2e_pass():
        t = time.perf_counter()
        eng = make_engine(a.e2e_placement)
        for name, fn, args in (("load+scan", None, ()), ("read_cns", eng.read_cns, (cns,)), ("fetch", eng.fetch, ()),
                               ("hash", eng.hash_alignment_to_seg, ()), ("scan_join", eng._ctx.wait_scan, ()),
                               ("intervals", eng.find_amplicon_intervals, ()),
                               ("smalldel", eng.find_smalldel_breakpoints, ()), ("breakpoints", eng.find_breakpoints, ()),
                               ("graph", eng.build_graph, ()), ("assign_cov", eng.assign_cov, ())):
            if fn is not None:
                fn(*args)
            t2 = time.perf_counter()
            e2e_stage[name] = (t2 - t) * 1e3
            t = t2
        eng.closebam()
        e2e_stage["closebam"] = (time.perf_counter() - t) * 1e3
        return eng

    if a.only == "e2e":             # development aid: per-pass stage times of the e2e arm alone
        rows = []
        sampler = ClockSampler(local)
        sampler.start()
        for _ in range(warm + a.steps):
            t = time.perf_counter()
            e2e_pass()
            rows.append(dict(e2e_stage, total=(time.perf_counter() - t) * 1e3))
        sampler.stop_flag = True
        sampler.join(timeout=2)
        print(json.dumps({"only": "e2e", "placement": a.e2e_placement, "clocks": sampler.summary(), "passes": rows}), flush=True)
        return

    # plan + resident context
    eng0 = make_engine("copy")
    plan = build_plan(eng0, cns)
    ctx = eng0._ctx
    if world > 1:
        eng0.materialise = rank == 0       # the replicated result tables are delivered to rank 0's host only
        if xchg is not None and a.overlap_scan:
            # the exchange chain of a sharded pass (three all-gathers, the rank's share of fetch, the merge: ~100 small
            # kernels) next to the record scan: coral_rescan = pre-scan + full scan on its own stream with 74 block slots
            # left free.  Measured at N = 2: 2.42 ms per step against 2.15 ms with the chain behind the scan (the pre-scan
            # costs 0.2 ms and the small kernels run at half speed next to the scan) -> off by default
            ctx.set_option("pass_overlap_scan", 1)
            ctx.set_option("pass_scan_reserve", 74)
    tab = ctx.record_table()
    aligned = int(tab["aligned_bases"].sum())
    n_bp = len(eng0.new_bp_list)
    def one_pass():
        if world > 1:
            return sharded_device_pass(eng0, plan, gpu_ops, pipelined=(xchg is not None and not a.no_pipeline))
        return device_pass(ctx, plan, gpu_ops, pipelined=not a.no_pipeline)

    def n_launches():
        return ctx.launch_count + (eng0._cctx.launch_count if world > 1 else 0)

    import gc
    gc.collect()
    gc.freeze()          # the plan's engine holds ~10^6 Python objects: a generation-2 collection inside the timed loop costs
                         # milliseconds (measured: 20 timed steps averaged 2.19 ms, 200 steps 1.80 ms, kernel times equal)
    for _ in range(warm):
        out = one_pass()
    SHARD_MS.clear()
    if world > 1:
        eng0.fetch_laps.clear()
    d2h = result_bytes(out)
    if a.cprofile:
        import cProfile
        import pstats
        pr = cProfile.Profile()
        e2e_pass()
        pr.enable()
        for _ in range(20 if a.only != "e2e" else 3):
            one_pass() if a.only != "e2e" else e2e_pass()
        pr.disable()
        if rank == 0:
            pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
            pstats.Stats(pr).sort_stats("tottime").print_stats(40)
        if world > 1:
            dist.destroy_process_group()
        return
    if a.profile:
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStart()
        one_pass()
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStop()
        print(json.dumps({"profile": "one device pass", "records": n_rec}), flush=True)
        return
    sampler = ClockSampler(local)
    sampler.start()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    l0 = n_launches()
    barrier()
    gc.disable()         # as reconstruct_graph() runs its pass (host_steps.gc_paused)
    t0 = time.perf_counter()
    scan_ms, prescan_ms, stage_ms = [], [], np.zeros(8)
    replayed = 0
    for _ in range(a.steps):
        ctx.reset_stage_times()            # stage_ms below: only the stages this step ran
        replayed += int(bool(one_pass().get("replayed")))
        # the roofline kernel: in a pipelined pass the aux walk of the scan runs as a separate pre-scan kernel, so the
        # time that goes with the scan's algorithmic bytes is pre-scan + scan
        sm = np.maximum(ctx.stage_ms_all().astype(np.float64), 0.0)      # one call: CUDA-event times of this step's stages
        scan_ms.append(sm[0] + sm[9])
        prescan_ms.append(sm[9])
        stage_ms += sm[:8]
    barrier()
    dt = time.perf_counter() - t0
    gc.enable()
    launches = n_launches() - l0
    # e2e through the public class, from pinned host records
    e2e_pass()
    barrier()
    t1 = time.perf_counter()
    e2e_steps = max(1, min(a.steps, 5))
    e2e_ms = []
    for _ in range(e2e_steps):
        tp = time.perf_counter()
        e2e_pass()
        e2e_ms.append((time.perf_counter() - tp) * 1e3)
    barrier()
    dt_e2e = time.perf_counter() - t1
    sampler.stop_flag = True
    sampler.join(timeout=2)
    parity = parity_check(a, rank, world, local, make_engine, cns, comm)
    # SURVEY 8f-2 (the step after the path): compute_path_constraints through the public class, on its own clock
    pc_info = None
    if world == 1:
        eng = make_engine("copy")
        with gc_paused():
            for fn, args in ((eng.read_cns, (cns,)), (eng.fetch, ()), (eng.hash_alignment_to_seg, ()),
                             (eng.find_amplicon_intervals, ()), (eng.find_smalldel_breakpoints, ()),
                             (eng.find_breakpoints, ()), (eng.build_graph, ()), (eng.assign_cov, ())):
                fn(*args)
            eng.compute_path_constraints()
            l1 = eng._ctx.launch_count
            t2 = time.perf_counter()
            eng.compute_path_constraints()
            pc_ms = (time.perf_counter() - t2) * 1e3
        pc_info = {"ms_through_the_class": pc_ms, "kernel_ms": eng._ctx.stage_ms(8), "launches": int(eng._ctx.launch_count - l1),
                   "constraints": int(sum(len(v[0]) for v in eng.path_constraints.values())),
                   "sequence_edges": int(sum(len(g.sequence_edges) for g in eng.lr_graph)),
                   "note": "not part of value / e2e; the CPU port's time for the same step is cpu_baseline.stage_s.path_constraints"}
        eng.closebam()

    # SURVEY 8f-3 (the step before the path): .bam bytes -> graph, on a bounded sample of the workload written as a real
    # BGZF file (the full-size file takes minutes to deflate on the host; the sample is stated in the line)
    file_info = None
    if world == 1 and not a.no_file and a.scale >= 0.5:
        file_info = e2e_from_file(a, local, FILE_SAMPLE_SCALE * a.scale)

    tt = torch.tensor([dt, dt_e2e], dtype=torch.float64, device="cuda")
    tot = torch.tensor([n_rec, aligned, launches], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    dt, dt_e2e = float(tt[0]), float(tt[1])
    n_rec_all, aligned_all, launches_all = int(tot[0]), int(tot[1]), int(tot[2])
    peak, peak_src = hbm_peak()
    traffic, traffic_src = measured_traffic(a.workload, a.scale, world)
    k_ms = float(np.mean(scan_ms))
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    names = ["scan", "coverage", "fetch", "smalldel", "candidates", "cluster", "bpc2bp", "concordant"]
    out = {
        "metric": METRIC, "value": n_rec_all * a.steps / dt, "unit": "records/s",
        "n_gpus": world, "steps": a.steps, "warmup": warm, "ms_per_step": dt / a.steps * 1e3,
        "higher_is_better": True, "scaling": ("strong" if strong else "weak"), "vs_baseline": None,
        "dtype": "u8/int32/int64 (f64 stats)", "data": "synthetic",
        "config": config_dict(a, world, n_rec, nbytes, aligned, n_bp, alg_bytes),
        "aligned_gbases_per_s": aligned_all * a.steps / dt / 1e9,
        "e2e": {"value": n_rec_all * e2e_steps / dt_e2e, "unit": "records/s", "ms_per_step": dt_e2e / e2e_steps * 1e3,
                "h2d_bytes_per_step": (alg_bytes - 76 * n_rec if a.e2e_placement.startswith("mapped") else nbytes) + 8 * (n_rec + 1),
                "d2h_bytes_per_step": d2h, "placement": a.e2e_placement, "steps": e2e_steps,
                "ms_per_pass": [round(x, 2) for x in e2e_ms], "stage_ms": e2e_stage},
        "gpu_launches": int(launches_all),
        "device_pass": {"pipelined": ((world == 1 or xchg is not None) and not a.no_pipeline), "steps_without_intermediate_sync": replayed,
                        "exchange": (None if world == 1 else a.exchange),
                        "note": "one stream of kernels per step (coral_pass_begin / coral_pass_end): one synchronisation and one "
                                "deferred copy per result table; launch sizes that depend on device results are replayed from the "
                                "previous pass over the same records and checked on the device"},
        "roofline": {"bound": "hbm", "kernel": "k_scan_records", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": k_ms,
                     "kernel_ms_parts": {"k_prescan_records": float(np.mean(prescan_ms)), "k_scan_records": k_ms - float(np.mean(prescan_ms)),
                                         "note": "inside a pipelined pass the scan is two kernels (pre-scan: record core, name hash, "
                                                 "aux walk; scan: CIGAR / SEQ); kernel_ms is their sum"},
                     "whole_pass": {"achieved": alg_bytes / (dt / a.steps) / 1e9, "frac": alg_bytes / (dt / a.steps) / 1e9 / peak,
                                    "note": "the same algorithmic bytes over the whole device pass (every stage, results read back)"}},
        "stage_ms_per_step": {n: float(stage_ms[i] / a.steps) for i, n in enumerate(names)},
        "clocks": sampler.summary(),
    }
    out.update(parity)
    if pc_info is not None:
        out["path_constraints_step"] = pc_info
    if file_info is not None:
        out["e2e_from_file"] = file_info
    if world > 1:
        out["sharded_stage_ms_per_step"] = {k: v / a.steps for k, v in SHARD_MS.items()}
        out["fetch_exchange_ms_per_step"] = {k: v / a.steps for k, v in eng0.fetch_laps.items()}
    if rank == 0:
        if world == 1 and not a.no_cpu:
            try:       # bounded sample (about 10-30 s of CPU work): the same workload at scale CPU_SAMPLE_SCALE, two passes
                r = cpu_arm(a, 1, CPU_SAMPLE_SCALE * a.scale, max_steps=2, budget_s=60.0)
                out["cpu_baseline"] = {
                    "value": r["n_rec"] * r["steps"] / r["seconds"], "unit": "records/s", "cores": 1, "kind": r["kind"],
                    "sample": "%s at scale %g (%d records, %.3f aligned Gbases), decoded records in host memory -> graph + "
                              "coverage (the stage set of e2e), %d pass(es); the full-size figure is the `--impl reference` line"
                              % (a.workload, r["scale"], r["n_rec"], r["aligned"] / 1e9, r["steps"]),
                    "stage_s": r["stage_s"]}
            except Exception as ex:  # noqa
                out["cpu_baseline"] = {"error": repr(ex)}
        print(json.dumps(out), flush=True)
    if world > 1:
        if xchg is not None:
            sharded.close_xchg(comm, xchg)
        dist.destroy_process_group()
    if out.get("parity_ok") is False:
        sys.exit(3)


if __name__ == "__main__":
    if os.environ.get("PYTHONHASHSEED") != "0":
        # the reference iterates a set of read names (infer_breakpoint_graph.py:432): the order of new_bp_list, and with
        # it what the parity digests compare, is only defined under a fixed string-hash seed (the fixtures use 0)
        os.environ["PYTHONHASHSEED"] = "0"
        os.execv(sys.executable, [sys.executable] + sys.argv)
    # stdout carries the JSON line only: libraries that write to fd 1 (NCCL prints its version there) go to stderr
    sys.stdout.flush()
    _json_fd = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(_json_fd, "w")
    main()
    sys.stdout.flush()
