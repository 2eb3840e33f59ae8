"""Place the UNMODIFIED reference under ``baseline/_ref`` (git-ignored, travels with gpurun).

Step 1 is the sanctioned offline install:

    python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
        --target baseline/_ref /root/reference

The reference is a directory of scripts with neither ``setup.py`` nor ``pyproject.toml``, so pip
answers "Directory '/root/reference' is not installable" (recorded in DESIGN.md).  Step 2 therefore
copies the source tree byte for byte (``master/``, ``slave/``, ``README.md``; the stale
``__pycache__`` directories are skipped) and writes a SHA-256 manifest so that the harness can
prove at run time that it executes unmodified files.
"""
from __future__ import annotations

import hashlib
import json
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
DEST = HERE / "_ref"
SRC = Path("/root/reference")


def sha256(p: Path) -> str:
    return hashlib.sha256(p.read_bytes()).hexdigest()


def main() -> int:
    if not SRC.is_dir():
        print(f"{SRC} not present; nothing to install", file=sys.stderr)
        return 1
    r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
                        "/opt/wheelhouse", "--target", str(DEST), str(SRC)], capture_output=True, text=True)
    pip_msg = (r.stdout + r.stderr).strip().splitlines()[-1] if (r.stdout + r.stderr).strip() else ""
    print("pip install:", "ok" if r.returncode == 0 else f"failed ({pip_msg})")
    DEST.mkdir(exist_ok=True)
    manifest = {}
    for f in sorted(SRC.rglob("*")):
        if f.is_dir() or "__pycache__" in f.parts:
            continue
        rel = f.relative_to(SRC)
        out = DEST / rel
        out.parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(f, out)
        manifest[str(rel)] = sha256(out)
    (DEST / "MANIFEST.json").write_text(json.dumps({"source": str(SRC), "pip": pip_msg, "sha256": manifest}, indent=1))
    print(f"copied {len(manifest)} files into {DEST}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
