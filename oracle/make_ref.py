"""TEST INFRASTRUCTURE ONLY — recipe that stages the UNMODIFIED reference package for the GPU box.

    python -m oracle.make_ref            # /root/reference/ggpm/*.py -> oracle/_ref/ggpm/*.py (byte-identical copies)

The reference is pure Python (no build step), so "building" it is copying its package directory.  The copy lands
in ``oracle/_ref/`` only, which is git-ignored (it never enters the history — the reference's sources are not part
of this repo) but NOT gpurun-ignored, so it travels to the GPU box with the snapshot exactly like the built
``libggpm_b200.so``.  There, ``oracle/ref_loader.py`` imports it (rdkit stubbed) so that

  * ``bench.py --impl reference`` and the ``cpu_baseline`` leg time the reference's OWN ``HierMPNEncoder`` /
    ``MotifEncoder`` (``cpu_baseline.kind = "reference"``), and
  * the wrapper-level drop-in tests (tests/test_wrappers.py) can build ``HierPropertyVAE`` / ``HierVAE`` /
    ``PropertyVAE`` around the drop-in encoder on the B200.

A manifest with the sha256 of every staged file is written next to the copies; ``verify()`` re-checks it, and —
when /root/reference is present — that the copies still equal the originals.  ``__graft_entry__.build()`` runs this
recipe whenever /root/reference exists (this container); the GPU box only uses the staged files.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("GGPM_REFERENCE_ROOT", "/root/reference")
REF_DST = os.path.join(HERE, "_ref")
PACKAGE = "ggpm"


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def source_available() -> bool:
    return os.path.isfile(os.path.join(REF_SRC, PACKAGE, "encoder.py"))


def staged() -> bool:
    return os.path.isfile(os.path.join(REF_DST, PACKAGE, "encoder.py")) and \
        os.path.isfile(os.path.join(REF_DST, "MANIFEST.json"))


def stage() -> str:
    """copy the reference package (only its .py files) into oracle/_ref/; returns the destination root"""
    if not source_available():
        raise RuntimeError("reference checkout not found at %s" % REF_SRC)
    src_pkg, dst_pkg = os.path.join(REF_SRC, PACKAGE), os.path.join(REF_DST, PACKAGE)
    os.makedirs(dst_pkg, exist_ok=True)
    manifest = {}
    for name in sorted(os.listdir(src_pkg)):
        if not name.endswith(".py"):
            continue
        s, d = os.path.join(src_pkg, name), os.path.join(dst_pkg, name)
        if not (os.path.exists(d) and _sha(d) == _sha(s)):
            shutil.copyfile(s, d)
        manifest[PACKAGE + "/" + name] = _sha(d)
    commit = None
    sub = os.path.join(REF_SRC, ".SUBMODULES.json")
    if os.path.exists(sub):
        try:
            commit = json.load(open(sub)).get("commit")
        except Exception:
            commit = None
    with open(os.path.join(REF_DST, "MANIFEST.json"), "w") as f:
        json.dump({"source": REF_SRC, "commit": commit, "files": manifest}, f, indent=1, sort_keys=True)
    return REF_DST


def verify() -> bool:
    """staged files match their manifest (and the originals when those are present)"""
    if not staged():
        return False
    man = json.load(open(os.path.join(REF_DST, "MANIFEST.json")))
    for rel, sha in man["files"].items():
        p = os.path.join(REF_DST, rel)
        if not os.path.exists(p) or _sha(p) != sha:
            return False
        o = os.path.join(REF_SRC, rel)
        if os.path.exists(o) and _sha(o) != sha:
            return False
    return True


if __name__ == "__main__":
    print(stage())
    sys.exit(0 if verify() else 1)
