"""Parity pins at BASELINE.json's NAMED sizes: run the UNMODIFIED reference (oracle/ref_runner.py, fresh subprocess,
PYTHONHASHSEED=0) on C1 / C2 / C3 at scale 1.0 and on stated sub-samples of C4 / C5, and commit compact digests
(tests/golden/full_digest.py) as tests/golden/<case>.json.gz.  Only runnable where /root/reference exists.

    python tests/golden/make_full_digests.py [case ...] [--stop_after assign_cov]
"""
import gzip
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import full_digest as FD  # noqa: E402


def main():
    from sim.generate import simulate, write_dataset
    args = sys.argv[1:]
    stop_after = None
    if "--stop_after" in args:
        i = args.index("--stop_after")
        stop_after = args[i + 1]
        del args[i:i + 2]
    for case in FD.FULL_CASES:
        if args and case not in args:
            continue
        t0 = time.time()
        cfg = FD.make_cfg(case)
        data = simulate(cfg)
        d = tempfile.mkdtemp(prefix="full_%s_" % case, dir=os.environ.get("CORAL_TMP", "/tmp"))
        paths = write_dataset(data, d)
        n_records, decoded = int(data.n_records), int(data.rec_off[-1])
        del data
        t_sim = time.time() - t0
        out = os.path.join(d, "state.json")
        env = dict(os.environ, PYTHONHASHSEED="0")
        cmd = [sys.executable, "-m", "oracle.ref_runner", "--bam", paths["bam"], "--cn_seg", paths["cn_seg"],
               "--seed", paths["seed"], "--out", out]
        if stop_after:
            cmd += ["--stop_after", stop_after]
        t1 = time.time()
        subprocess.check_call(cmd, cwd=ROOT, env=env)
        t_ref = time.time() - t1
        st = json.load(open(out))
        dg = {"case": case, "what": FD.FULL_CASES[case][3], "n_records": n_records, "decoded_bytes": decoded,
              "reference_timing_s": st["timing_s"], "reference_wall_s": t_ref, "stages": FD.digest_state(st["stages"])}
        with gzip.open(os.path.join(HERE, case + ".json.gz"), "wt") as fp:
            json.dump(dg, fp, separators=(",", ":"))
        s = dg["stages"]
        print(case, "records", n_records, "decoded GB %.2f" % (decoded / 1e9), "chimeric", s["fetch"]["chimeric"]["n"],
              "bps", len(s["find_breakpoints"]["new_bp_list"]) if "find_breakpoints" in s else None,
              "graphs", len(s["assign_cov"]["graphs"]) if "assign_cov" in s else None,
              "sim s %.0f ref s %.0f" % (t_sim, t_ref), json.dumps(st["timing_s"]), flush=True)
        shutil.rmtree(d, ignore_errors=True)


if __name__ == "__main__":
    main()
