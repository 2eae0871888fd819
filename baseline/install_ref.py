#!/usr/bin/env python
"""Install the UNMODIFIED reference package for the hot path into the git-ignored baseline/_ref/.

    python baseline/install_ref.py            # needs /root/reference (build container only)

The reference has no setup.py / pyproject, so `pip install --target baseline/_ref /root/reference`
has nothing to build (outcome recorded in DESIGN.md section 4); its hot path is three pure-Python
files, which this recipe copies byte for byte:

    /root/reference/build_hypergraph/{__init__,similarity_kernel,preprocess_hypergraph}.py
        -> baseline/_ref/build_hypergraph/

baseline/_ref/ is listed in .gitignore (reference sources never enter this repository's history)
but not in .gpurunignore, so the copy travels to the GPU box with the snapshot, like the built
.so.  Users: bench.py --impl reference (the timed CPU arm) and tests/ (parity against the real
reference on the box's CPU) through baseline/ref_loader.py.  Nothing under multimodal-fusion_b200/
may import it.  A MANIFEST with the sha256 of every copied file is written beside the copy.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("HG_REFERENCE_ROOT", "/root/reference")
FILES = ("__init__.py", "similarity_kernel.py", "preprocess_hypergraph.py")


def install(verbose: bool = True) -> bool:
    src = os.path.join(REF, "build_hypergraph")
    if not os.path.isdir(src):
        if verbose:
            print(f"install_ref: {src} not present (GPU box?) - keeping whatever baseline/_ref holds")
        return os.path.isfile(os.path.join(HERE, "_ref", "build_hypergraph", "preprocess_hypergraph.py"))
    dst = os.path.join(HERE, "_ref", "build_hypergraph")
    os.makedirs(dst, exist_ok=True)
    lines = []
    for name in FILES:
        shutil.copyfile(os.path.join(src, name), os.path.join(dst, name))
        with open(os.path.join(dst, name), "rb") as f:
            lines.append(f"{hashlib.sha256(f.read()).hexdigest()}  build_hypergraph/{name}")
    with open(os.path.join(HERE, "_ref", "MANIFEST.sha256"), "w") as f:
        f.write("\n".join(lines) + "\n")
    if verbose:
        print("install_ref: copied", ", ".join(FILES), "->", dst)
    return True


if __name__ == "__main__":
    sys.exit(0 if install() else 1)
