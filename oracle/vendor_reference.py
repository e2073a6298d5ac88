"""TEST / BASELINE INFRASTRUCTURE ONLY -- places an UNMODIFIED copy of the reference's Python sources
under the git-ignored baseline/_ref/ so that the real reference can be TIMED on the GPU box's host cores
(bench.py: cpu_baseline leg and --impl reference, the cases it can run: cfg1).  /root/reference does not
exist on that box; baseline/_ref/ travels with the snapshot (git-ignored, not gpurun-ignored) and never
enters the history.  The reference has no setup.py / pyproject (Dockerfile + Makefile only), so there is
nothing to pip-install: the recipe is a plain copy of src/*.py and tests/__context__.py.

    python oracle/vendor_reference.py          # run by __graft_entry__.build() when /root/reference exists

Nothing in the product path reads baseline/_ref/.
"""
import glob
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEST = os.path.join(ROOT, "baseline", "_ref")


def vendor(ref_root="/root/reference"):
    if not os.path.isdir(os.path.join(ref_root, "src")):
        return None
    for sub, pattern in (("src", "*.py"), ("tests", "__context__.py")):
        os.makedirs(os.path.join(DEST, sub), exist_ok=True)
        for f in glob.glob(os.path.join(ref_root, sub, pattern)):
            shutil.copyfile(f, os.path.join(DEST, sub, os.path.basename(f)))
    lic = os.path.join(ref_root, "LICENSE")
    if os.path.exists(lic):
        shutil.copyfile(lic, os.path.join(DEST, "LICENSE"))
    return DEST


if __name__ == "__main__":
    d = vendor(*sys.argv[1:2])
    print(d if d else "reference tree not present: nothing vendored")
