"""Stages the UNMODIFIED reference sources of the hot path into ``oracle/_ref/`` so that they travel to the GPU box.

TEST INFRASTRUCTURE ONLY.  The reference is pure Python: "building" it means placing the two files its xattn path consists
of — ``src/models/fusion.py`` and ``src/models/temporal.py`` (they import only ``torch`` and each other) — under
``oracle/_ref/src/models/`` next to empty package markers.  ``oracle/_ref/`` is git-ignored (no reference source enters the
history) but NOT gpurun-ignored, so ``bench.py --impl reference`` / ``cpu_baseline`` can time the reference's own module
(``kind: "reference"``) on the GPU box's host cores, where ``/root/reference`` does not exist.  ``__graft_entry__.build()``
runs this whenever ``/root/reference`` is present; MANIFEST.json records the SHA-256 of what was staged.

    python oracle/stage_ref.py [/root/reference]
"""
from __future__ import annotations

import hashlib
import json
import shutil
import sys
from pathlib import Path

FILES = ("src/models/fusion.py", "src/models/temporal.py")
DEST = Path(__file__).resolve().parent / "_ref"


def stage(reference_root: str = "/root/reference") -> bool:
    root = Path(reference_root)
    if not all((root / f).is_file() for f in FILES):
        return False
    manifest = {}
    for f in FILES:
        dst = DEST / f
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(root / f, dst)
        manifest[f] = hashlib.sha256(dst.read_bytes()).hexdigest()
    for pkg in (DEST / "src", DEST / "src" / "models"):
        (pkg / "__init__.py").write_text("")          # package markers only: the reference's own __init__ imports librosa-bound code
    (DEST / "MANIFEST.json").write_text(json.dumps({"source": str(root), "sha256": manifest}, indent=1))
    return True


if __name__ == "__main__":
    ok = stage(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
    print("staged" if ok else "reference sources not found; nothing staged", DEST)
