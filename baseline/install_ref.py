"""Recipe for `baseline/_ref/`: the UNMODIFIED reference modules that `bench.py --impl reference` and the
`cpu_baseline` leg time on the GPU box's host cores.

The reference (MattiDeBeer/image-segmentation) is a plain source tree — no setup.py / pyproject.toml — so
`pip install /root/reference` has nothing to install; its Python packages `models/` and `customDatasets/` are copied
byte for byte instead.  `baseline/_ref/` is git-ignored (never part of this repo's history) and NOT gpurun-ignored, so it
travels to the GPU box with the snapshot, where /root/reference does not exist.  Run in the build container:

    python baseline/install_ref.py          (also done by __graft_entry__.build() when /root/reference is present)

Importing them needs stand-ins for three packages that are absent from the image and only used for plotting / training
augmentation / the Dice loss object (oracle/refload.py registers them); the arithmetic that is timed — the Dataset-wrapper
perturbations, UNet.forward, IoU, PixelAccuracy — is the reference's own code.
"""
import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
PACKAGES = ("models", "customDatasets")


def install(src_root: str = "/root/reference", quiet: bool = False) -> bool:
    if not os.path.isdir(os.path.join(src_root, "models")):
        return False
    os.makedirs(DEST, exist_ok=True)
    manifest = []
    for pkg in PACKAGES:
        dst = os.path.join(DEST, pkg)
        if os.path.isdir(dst):
            shutil.rmtree(dst)
        shutil.copytree(os.path.join(src_root, pkg), dst, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
        for dirpath, _, files in os.walk(dst):
            for f in sorted(files):
                p = os.path.join(dirpath, f)
                with open(p, "rb") as fh:
                    manifest.append(f"{hashlib.sha256(fh.read()).hexdigest()}  {os.path.relpath(p, DEST)}")
    with open(os.path.join(DEST, "MANIFEST.sha256"), "w") as fh:
        fh.write("\n".join(sorted(manifest)) + "\n")
    if not quiet:
        print(f"[baseline] copied {len(manifest)} reference files from {src_root} to {DEST}")
    return True


def available() -> bool:
    return os.path.isfile(os.path.join(DEST, "models", "UNet.py"))


if __name__ == "__main__":
    sys.exit(0 if install(*(sys.argv[1:2])) else 1)
