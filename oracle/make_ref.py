"""Recipe for oracle/_ref/: a verbatim copy of the reference's ``ldm`` python package (and the Shapes3D yaml) so that the
UNMODIFIED reference modules can be timed and cross-checked on the GPU box, where /root/reference does not exist.

  python -m oracle.make_ref          (run in the build container; __graft_entry__.build() runs it when the checkout is there)

oracle/_ref/ is git-ignored (no reference source enters the repository's history) and NOT gpurun-ignored, so it travels
with the snapshot exactly like the built libencdiff_b200.so.  Nothing under encdiff_b200/ reads it; users are
oracle/ref_shim.py -> tests/ (cross-checks through the reference's own instantiate_from_config / LatentDiffusion /
DDIMSampler) and bench.py's reference + comparator legs.  Test / baseline infrastructure only.
"""
from __future__ import annotations

import os
import shutil
import sys

SRC = os.environ.get("ENCDIFF_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
YAMLS = ["configs/latent-diffusion/shapes3d-vq-4-16-encdiff.yaml"]


def make(verbose=True) -> bool:
    if not os.path.isdir(os.path.join(SRC, "ldm")):
        if verbose:
            print(f"oracle.make_ref: no reference checkout at {SRC}; keeping {DST} as it is")
        return os.path.isdir(os.path.join(DST, "ldm"))
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    n = 0
    for root, dirs, files in os.walk(os.path.join(SRC, "ldm")):
        dirs[:] = [d for d in dirs if d != "__pycache__"]
        for f in files:
            if f.endswith(".py"):
                rel = os.path.relpath(os.path.join(root, f), SRC)
                os.makedirs(os.path.dirname(os.path.join(DST, rel)), exist_ok=True)
                shutil.copyfile(os.path.join(SRC, rel), os.path.join(DST, rel))
                n += 1
    for y in YAMLS:
        if os.path.exists(os.path.join(SRC, y)):
            os.makedirs(os.path.dirname(os.path.join(DST, y)), exist_ok=True)
            shutil.copyfile(os.path.join(SRC, y), os.path.join(DST, y))
    with open(os.path.join(DST, "README"), "w") as fh:
        fh.write("Verbatim copy of SelenaGeRuiqi/EncDiff `ldm/**/*.py` made by oracle/make_ref.py. Not part of the repository.\n")
    if verbose:
        print(f"oracle.make_ref: copied {n} files into {DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if make() else 1)
