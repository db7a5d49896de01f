"""Installs the UNMODIFIED reference (pure Python) under baseline/_ref/ so that `bench.py --impl reference` and the
`cpu_baseline` leg can time the reference's own code on the GPU box, where /root/reference does not exist.

    python baseline/install_ref.py        (run by __graft_entry__.build() when /root/reference is present)

`pip install --target baseline/_ref /root/reference` does not apply: the reference root is not one package (five
extension packages under source/, each with Isaac Sim / Kit dependencies that are not installable here).  What the timed
path needs is the packages' Python sources, imported through the stub loader tests/golden/ref_harness.py exactly as the
golden fixtures were generated (only Omniverse / PhysX / Warp / trimesh / gymnasium / tensordict are stubbed).  The files
are copied byte for byte; nothing is patched.  baseline/_ref/ is git-ignored (not product source) and NOT
gpurun-ignored, so it travels to the box with the snapshot."""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
SRC = os.environ.get("B200LAB_REFERENCE_SRC", "/root/reference")
PACKAGES = ("isaaclab", "isaaclab_tasks", "isaaclab_assets", "isaaclab_rl", "rsl_rl")
KEEP = (".py", ".toml", ".yaml", ".yml", ".json", ".txt", ".cfg")


def install(force: bool = False) -> str | None:
    if not os.path.isdir(os.path.join(SRC, "source", "isaaclab")):
        return DST if os.path.isdir(os.path.join(DST, "source", "isaaclab")) else None
    stamp = os.path.join(DST, "INSTALLED")
    if os.path.exists(stamp) and not force:
        return DST
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    n = 0
    for pkg in PACKAGES:
        root = os.path.join(SRC, "source", pkg)
        for dirpath, dirnames, files in os.walk(root):
            dirnames[:] = [d for d in dirnames if d not in ("docs", "test", "tests", "__pycache__", "data")]
            for f in files:
                if not f.endswith(KEEP):
                    continue
                rel = os.path.relpath(os.path.join(dirpath, f), SRC)
                out = os.path.join(DST, rel)
                os.makedirs(os.path.dirname(out), exist_ok=True)
                shutil.copyfile(os.path.join(dirpath, f), out)
                n += 1
    with open(stamp, "w") as fh:
        fh.write(f"{n} files copied unmodified from {SRC}/source/{{{','.join(PACKAGES)}}}\n")
    return DST


if __name__ == "__main__":
    print(install(force="-f" in sys.argv))
