"""Recipe that puts the UNMODIFIED reference package where the GPU box can import it: baseline/_ref/.

`python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref
/root/reference` cannot work in this image: the reference's build backend (hatchling, pyproject.toml:43-45)
is neither installed nor in /opt/wheelhouse, and there is no network.  The reference is pure Python and its
wheel is declared as `packages = ["src/bloqade"]` (pyproject.toml:50-51), so this recipe produces what that
install would: the package tree copied byte for byte plus a dist-info METADATA (the package asks
importlib.metadata for its own version, src/bloqade/analog/__init__.py:28).

baseline/_ref/ is git-ignored (reference sources never enter this repository's history) and NOT
gpurun-ignored, so it travels to the GPU box with the snapshot.  The absent third-party imports (plotext,
bokeh, braket, plum, simplejson ...) are stubbed at import time by baseline/refshim.py, exactly as in the
build container.  Run by `__graft_entry__.build()` whenever /root/reference is present.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
TARGET = os.path.join(HERE, "_ref")
SOURCE = "/root/reference"


def _version(pyproject: str) -> str:
    for line in open(pyproject):
        if line.startswith("version"):
            return line.split("=", 1)[1].strip().strip('"')
    return "0"


def _tree_stamp(root: str) -> str:
    n, size, newest = 0, 0, 0.0
    for d, _, files in os.walk(root):
        for f in files:
            if f.endswith(".py"):
                st = os.stat(os.path.join(d, f))
                n, size, newest = n + 1, size + st.st_size, max(newest, st.st_mtime)
    return f"{n}:{size}:{newest:.0f}"


def install(force: bool = False) -> bool:
    """Returns True when baseline/_ref holds the reference afterwards."""
    src = os.path.join(SOURCE, "src", "bloqade")
    if not os.path.isdir(src):
        return os.path.isdir(os.path.join(TARGET, "bloqade", "analog"))
    stamp_file = os.path.join(TARGET, ".stamp")
    stamp = _tree_stamp(src)
    if not force and os.path.exists(stamp_file) and open(stamp_file).read() == stamp:
        return True
    if os.path.isdir(TARGET):
        shutil.rmtree(TARGET)
    os.makedirs(TARGET)
    shutil.copytree(src, os.path.join(TARGET, "bloqade"), ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    version = _version(os.path.join(SOURCE, "pyproject.toml"))
    info = os.path.join(TARGET, f"bloqade_analog-{version.replace('-', '.')}.dist-info")
    os.makedirs(info)
    with open(os.path.join(info, "METADATA"), "w") as f:
        f.write(f"Metadata-Version: 2.1\nName: bloqade-analog\nVersion: {version}\n")
    with open(os.path.join(info, "INSTALLER"), "w") as f:
        f.write("baseline/install_ref.py (copy of src/bloqade; hatchling unavailable offline)\n")
    with open(stamp_file, "w") as f:
        f.write(stamp)
    return True


if __name__ == "__main__":
    ok = install(force="--force" in sys.argv)
    print("baseline/_ref:", "ready" if ok else "reference not available")
