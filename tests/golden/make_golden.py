"""Generate the golden fixtures by EXECUTING THE REFERENCE (only possible where /root/reference exists).

    python tests/golden/make_golden.py            # writes tests/golden/<case>.json.gz and l2_kat.json

Each case = one deterministic synthetic dataset (sim/) run through the unmodified reference under the
oracle's pysam/intervaltree shims (oracle/ref_runner.py) in a fresh subprocess, PYTHONHASHSEED=0.
The fixtures pin oracle/coral_oracle.py, coral_b200/host_steps.py and the CUDA path.
"""
import gzip
import hashlib
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

CASES = {
    # name: (config factory name, kwargs, overrides applied to the SimConfig)
    "c1_small": ("C1", dict(scale=0.05), {}),
    "c1_n": ("C1", dict(scale=0.04, seed=99), {"p_n": 0.001}),
    "c2_small": ("C2", dict(scale=0.03), {}),
    "c2_oneseed": ("C2", dict(scale=0.03, seed=7), {"seeds": "first"}),
    "c3_small": ("C3", dict(scale=0.03, n_del=40), {}),
    "c4_small": ("C4", dict(scale=0.012), {}),
    "c5_mini": ("C5", dict(scale=0.08, n_seeds=3, chroms=['chr20', 'chr21', 'chr22']), {}),
}


def make_cfg(case):
    from sim import configs
    name, kw, over = CASES[case]
    cfg = configs.CONFIGS[name](**kw)
    for k, v in over.items():
        if k == "seeds" and v == "first":
            cfg.seeds = cfg.seeds[:1]
        else:
            setattr(cfg, k, v)
    return cfg


def digest_names(names):
    return {"n": len(names), "sha1": hashlib.sha1("\n".join(sorted(names)).encode()).hexdigest()}


def compact(state):
    """Replace the large name sets of concordant edges by (count, sha1)."""
    for g in state["stages"].get("assign_cov", {}).get("graphs", []):
        for ce in g["concordant_edges"]:
            ce[9] = digest_names(ce[9])
    return state


def main():
    from sim.generate import simulate, write_dataset
    only = sys.argv[1:]
    for case in CASES:
        if only and case not in only:
            continue
        cfg = make_cfg(case)
        data = simulate(cfg)
        d = tempfile.mkdtemp(prefix="golden_%s_" % case)
        paths = write_dataset(data, d)
        out = os.path.join(d, "state.json")
        env = dict(os.environ, PYTHONHASHSEED="0")
        subprocess.check_call([sys.executable, "-m", "oracle.ref_runner", "--bam", paths["bam"], "--cn_seg", paths["cn_seg"],
                               "--seed", paths["seed"], "--out", out], cwd=ROOT, env=env)
        st = compact(json.load(open(out)))
        st["n_records"] = data.n_records
        st["case"] = case
        with gzip.open(os.path.join(HERE, case + ".json.gz"), "wt") as fp:
            json.dump(st, fp, separators=(",", ":"))
        s = st["stages"]
        print(case, "records", data.n_records, "chimeric", len(s["fetch"]["chimeric"]),
              "indel reads", len(s["find_smalldel_breakpoints"]["large_indel_alignments"]),
              "bps", len(s["find_breakpoints"]["new_bp_list"]),
              "intervals", len(s["find_amplicon_intervals"]["amplicon_intervals"]),
              "size KB", os.path.getsize(os.path.join(HERE, case + ".json.gz")) // 1024, st["timing_s"])


if __name__ == "__main__":
    main()
