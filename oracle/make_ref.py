"""Recipe for oracle/_ref/: the reference's OWN hot-path modules, staged so that they travel to the GPU box.

TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/__init__.py).  The reference is plain Python (no build system,
no native code), so "building" it is: read the six files of the caption path where they lie under
/root/reference, apply the in-memory patches recorded in oracle/ref_shim.py:_PATCHES (three py3/CPU
incompatibilities + one timing print), and write the result under oracle/_ref/.  That directory is git-ignored
(no reference source enters the history) but NOT gpurun-ignored, so `bench.py --impl reference` and the
cpu_baseline leg time the reference itself (`cpu_baseline.kind = "reference"`) on the GPU box's host cores.

    python oracle/make_ref.py            # also run by __graft_entry__.build() when /root/reference exists
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
if os.path.dirname(HERE) not in sys.path:
    sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_shim  # noqa: E402

OUT = os.path.join(HERE, "_ref")
FILES = ("CaptionGenerator.py", "models/__init__.py", "models/MA_attention_8_NEW.py", "models/OldModel_NEW.py",
         "models/sst_model.py", "misc/__init__.py", "misc/utils.py")


def main(src_root="/root/reference", quiet=False):
    if not os.path.isfile(os.path.join(src_root, "CaptionGenerator.py")):
        if not quiet:
            print("make_ref: no reference checkout at %s; oracle/_ref left as it is" % src_root)
        return False
    for rel in FILES:
        src = open(os.path.join(src_root, rel)).read()
        for old, new in ref_shim._PATCHES.get(rel, []):
            if old not in src:
                raise RuntimeError("reference changed: %r not found in %s" % (old, rel))
            src = src.replace(old, new)
        dst = os.path.join(OUT, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        with open(dst, "w") as f:
            f.write(src)
    with open(os.path.join(OUT, "PATCHED"), "w") as f:
        f.write("staged from %s by oracle/make_ref.py with the patches of oracle/ref_shim.py:_PATCHES\n" % src_root)
    if not quiet:
        print("make_ref: %d files -> %s" % (len(FILES), OUT))
    return True


if __name__ == "__main__":
    main()
