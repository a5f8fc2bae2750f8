"""Recipe: make the UNMODIFIED reference importable on the GPU box.  TEST / MEASUREMENT INFRASTRUCTURE ONLY.

    python oracle/install_reference.py        # needs /root/reference (the read-only mount of the build container)

The reference (AvalonEnjoyer/pytorch-ViT-paper-replication) is 973 lines of Python with no setup.py / pyproject, so
`pip install --target baseline/_ref /root/reference` has nothing to build. Its hot path is importable as it lies:
`models/vit.py`, `models/vit_no_classifier.py` and `going_modular/going_modular/engine.py` (plus the small modules
next to them). `/root/reference` does not exist on the GPU box, and `gpurun` ships the repository's working tree, so
this script copies those files byte for byte into `baseline/_ref/` -- a directory that is git-ignored (never committed,
never edited) but not gpurun-ignored. `bench.py --impl reference`, `bench.py`'s `cpu_baseline` / `gpu_eager_baseline`
legs and `oracle/reference_runner.py` import the reference from there (or straight from /root/reference when that is
mounted); the product package never does. __graft_entry__.build() runs this when the mount is present.
"""
from __future__ import annotations

import hashlib
import shutil
import sys
from pathlib import Path

REF = Path("/root/reference")
DEST = Path(__file__).resolve().parent.parent / "baseline" / "_ref"
FILES = [
    "models/vit.py",
    "models/vit_no_classifier.py",
    "going_modular/going_modular/engine.py",
    "going_modular/going_modular/utils.py",
    "going_modular/going_modular/data_setup.py",
    "going_modular/going_modular/predictions.py",
    "going_modular/going_modular/model_builder.py",
]


def install(quiet: bool = False) -> bool:
    if not REF.exists():
        if not quiet:
            print(f"{REF} is not mounted: nothing to install (an existing {DEST} is left alone)")
        return DEST.exists()
    manifest = []
    for rel in FILES:
        src = REF / rel
        dst = DEST / rel
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(src, dst)
        manifest.append(f"{hashlib.sha256(dst.read_bytes()).hexdigest()}  {rel}")
    (DEST / "MANIFEST.sha256").write_text("\n".join(manifest) + "\n")
    if not quiet:
        print(f"installed {len(FILES)} reference files into {DEST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if install() else 1)
