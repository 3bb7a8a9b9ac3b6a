"""TEST INFRASTRUCTURE ONLY -- stage the UNMODIFIED reference modules for the timed CPU arm.

bench.py --impl reference has to time the reference's own CPU implementation of the path on the
GPU box, which has no /root/reference.  This recipe (run by __graft_entry__.build() in the build
container, where the reference tree is mounted) copies the three modules of the path and the
bundled spectrum, byte for byte, into oracle/_ref/randomfield/ -- a git-ignored directory that
travels to the GPU box with the snapshot, like baseline/_ref of the base contract -- and writes
a sha256 manifest next to them.  Nothing under oracle/_ref/ is ever committed, and nothing in the
product imports it: oracle/ref_replay.py loads the modules from there when /root/reference is
absent.

    python oracle/make_ref.py            # copy + manifest
    python oracle/make_ref.py --check    # verify the staged files against the manifest
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("RF_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")
FILES = ["randomfield/transform.py", "randomfield/powertools.py", "randomfield/random.py",
         "randomfield/generate.py",          # only for the lensing loop (ref_replay.lensing_replay)
         "randomfield/data/default_power.dat"]


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


def stage():
    """Copy the files and write the manifest.  Returns the manifest dict."""
    man = {"source": SRC, "files": {}}
    sub = os.path.join(SRC, ".SUBMODULES.json")
    for rel in FILES:
        src, dst = os.path.join(SRC, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        man["files"][rel] = _sha(dst)
        if _sha(src) != man["files"][rel]:
            raise RuntimeError("copy of %s does not match its source" % rel)
    if os.path.exists(sub):
        man["submodules_sha256"] = _sha(sub)
    with open(os.path.join(DST, "MANIFEST.json"), "w") as f:
        json.dump(man, f, indent=1, sort_keys=True)
    return man


def check():
    """True if oracle/_ref holds every file with the hash the manifest recorded."""
    try:
        with open(os.path.join(DST, "MANIFEST.json")) as f:
            man = json.load(f)
        return all(_sha(os.path.join(DST, rel)) == man["files"][rel] for rel in FILES)
    except (OSError, KeyError, ValueError):
        return False


if __name__ == "__main__":
    if "--check" in sys.argv:
        ok = check()
        print("oracle/_ref: %s" % ("ok" if ok else "missing or modified"))
        sys.exit(0 if ok else 1)
    if not os.path.exists(os.path.join(SRC, FILES[0])):
        sys.exit("reference tree not found at %s" % SRC)
    print(json.dumps(stage(), indent=1))
