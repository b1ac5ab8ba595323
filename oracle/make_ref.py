"""oracle/make_ref.py — TEST INFRASTRUCTURE.  Copy recipe for ``oracle/_ref/``: the UNMODIFIED reference as a travelling
baseline.

The reference (suhas-r/CoRAL, BSD-3-Clause, pure Python) cannot be pip-installed into a runnable package here — its
dependencies pysam, cvxopt, gurobipy and intervaltree are absent and there is no network — but its hot path runs
unchanged under the oracle's stand-ins for those imports (oracle/ref_runner.py).  This recipe copies the reference's
``src/*.py`` and LICENSE, byte for byte, from where they lie (``/root/reference``) into ``oracle/_ref/`` — git-ignored,
so no reference source enters the history, but NOT gpurun-ignored, so the copy travels to the GPU box like the built
``.so`` files.  There ``bench.py --impl reference`` and the ``cpu_baseline`` leg time THE REFERENCE ITSELF
(``cpu_baseline.kind = "reference"``); without the copy they fall back to the CPU port (``kind = "port"``).

Nothing in the product (``coral_b200/``) imports ``oracle/_ref``; only ``bench.py``'s reference legs and ``tests/`` do.

    python -m oracle.make_ref          # no-op when /root/reference is absent
"""
import filecmp
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("CORAL_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")


def ref_src():
    """Directory holding the reference's modules: the travelling copy if present, else the original tree, else None."""
    for d in (os.path.join(DST, "src"), os.path.join(REF_ROOT, "src")):
        if os.path.exists(os.path.join(d, "infer_breakpoint_graph.py")):
            return d
    return None


def make_ref(verbose=False):
    src = os.path.join(REF_ROOT, "src")
    if not os.path.isdir(src):
        return None
    os.makedirs(os.path.join(DST, "src"), exist_ok=True)
    n = 0
    for name in sorted(os.listdir(src)):
        if not name.endswith(".py"):
            continue
        a, b = os.path.join(src, name), os.path.join(DST, "src", name)
        if not (os.path.exists(b) and filecmp.cmp(a, b, shallow=False)):
            shutil.copyfile(a, b)
            n += 1
    lic = os.path.join(REF_ROOT, "LICENSE")
    if os.path.exists(lic):
        shutil.copyfile(lic, os.path.join(DST, "LICENSE"))
    with open(os.path.join(DST, "README"), "w") as fp:
        fp.write("Unmodified copy of %s/src/*.py (suhas-r/CoRAL, BSD-3-Clause, see LICENSE) made by oracle/make_ref.py.\n"
                 "Git-ignored; used only by bench.py's reference legs and tests/ as the CPU baseline / checker.\n" % REF_ROOT)
    if verbose:
        print("oracle/_ref: %d file(s) copied" % n)
    return os.path.join(DST, "src")


if __name__ == "__main__":
    print(make_ref(verbose=True))
