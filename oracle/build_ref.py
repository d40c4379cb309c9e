"""Recipe for oracle/_ref: the reference's own hot-path files, unmodified, for the CPU arm of bench.py.

TEST / BASELINE INFRASTRUCTURE - nothing under sparse_coding__b200/ imports it.

The reference (johnathan217/sparse_coding_) is pure Python: "building" it means putting the four modules of the
path - autoencoders/{ensemble, sae_ensemble, learned_dict, fista}.py - where the GPU box can import them
(/root/reference does not exist there).  This script copies them byte for byte from /root/reference into
oracle/_ref/autoencoders/ and writes their SHA-256 to oracle/_ref/MANIFEST.json.  oracle/_ref/ is git-ignored (the
sources stay out of the repository's history) but travels to the GPU box with the snapshot, like a built .so.
The three packages those modules import that cannot be installed offline (optree, torchopt, torchtyping) come from
oracle/shims/ (see oracle/ref_loader.py).

    python oracle/build_ref.py            # run by __graft_entry__.build() when /root/reference is present
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_ROOT = os.environ.get("SC_REFERENCE_SRC", "/root/reference")
OUT = os.path.join(HERE, "_ref")
FILES = ["autoencoders/ensemble.py", "autoencoders/sae_ensemble.py", "autoencoders/learned_dict.py",
         "autoencoders/fista.py"]


def available() -> bool:
    """oracle/_ref holds the four modules (built here or shipped with the snapshot)."""
    return all(os.path.isfile(os.path.join(OUT, f)) for f in FILES)


def build(verbose: bool = False) -> bool:
    if not os.path.isfile(os.path.join(REFERENCE_ROOT, FILES[0])):
        return available()
    manifest = {}
    for f in FILES:
        src, dst = os.path.join(REFERENCE_ROOT, f), os.path.join(OUT, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        with open(dst, "rb") as fh:
            manifest[f] = hashlib.sha256(fh.read()).hexdigest()
    init = os.path.join(REFERENCE_ROOT, "autoencoders", "__init__.py")
    if os.path.isfile(init):
        shutil.copyfile(init, os.path.join(OUT, "autoencoders", "__init__.py"))
    with open(os.path.join(OUT, "MANIFEST.json"), "w") as fh:
        json.dump({"source": REFERENCE_ROOT, "sha256": manifest}, fh, indent=1)
    if verbose:
        print(f"oracle/_ref: {len(FILES)} reference modules copied from {REFERENCE_ROOT}")
    return True


if __name__ == "__main__":
    ok = build(verbose=True)
    sys.exit(0 if ok else 1)
