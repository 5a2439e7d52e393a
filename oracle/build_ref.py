"""Recipe for ``oracle/_ref/``: the reference's OWN loss module, staged so that it can travel to the GPU box.

TEST INFRASTRUCTURE ONLY (see oracle/loss_oracle.py): used by ``bench.py --impl reference`` / ``cpu_baseline`` as
the timed CPU arm and by tests as a second checker.  The product package never imports it.

The reference (bioscan-ml/bioscan-clip @ 382d5ff) is pure Python; the path BASELINE.json names lives in ONE file,
``bioscanclip/model/loss_func.py`` (ContrastiveLoss :23-54, gather_features :58-91, ClipLoss :94-165), which
imports nothing but torch (horovod is optional and guarded, :13-16).  ``/root/reference`` exists only in the build
container, so this script copies that file -- byte for byte, unmodified -- to

    oracle/_ref/bioscanclip_ref/loss_func.py        (+ an empty __init__.py and a SOURCE note)

``oracle/_ref/`` is listed in .gitignore (reference sources never enter the history) but not in .gpurunignore, so the
copy ships to the GPU box with the snapshot like the built ``.so``.  ``__graft_entry__.build()`` runs this recipe
whenever ``/root/reference`` is present; on the GPU box the prebuilt copy is used as it is.

    python oracle/build_ref.py
"""
from __future__ import annotations

import hashlib
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = "/root/reference"
REF_FILE = os.path.join(REF_ROOT, "bioscanclip", "model", "loss_func.py")
OUT_DIR = os.path.join(HERE, "_ref", "bioscanclip_ref")
OUT_FILE = os.path.join(OUT_DIR, "loss_func.py")


def available() -> bool:
    """True when the staged reference module exists (here after build(), on the GPU box from the snapshot)."""
    return os.path.exists(OUT_FILE)


def build_ref(verbose: bool = False) -> bool:
    """Stage the reference loss module.  Returns True if oracle/_ref is usable afterwards."""
    if not os.path.exists(REF_FILE):
        return available()
    os.makedirs(OUT_DIR, exist_ok=True)
    shutil.copyfile(REF_FILE, OUT_FILE)
    for d in (os.path.join(HERE, "_ref"), OUT_DIR):
        init = os.path.join(d, "__init__.py")
        if not os.path.exists(init):
            open(init, "w").close()
    with open(REF_FILE, "rb") as f:
        digest = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(OUT_DIR, "SOURCE"), "w") as f:
        f.write(f"unmodified copy of {REF_FILE}\nsha256 {digest}\nstaged by oracle/build_ref.py\n")
    if verbose:
        print(f"staged {REF_FILE} -> {OUT_FILE} (sha256 {digest[:16]}...)")
    return True


def load_reference_loss():
    """-> the reference's (ContrastiveLoss, ClipLoss) classes from oracle/_ref, or None when it was never staged."""
    if not available():
        return None
    import importlib.util

    spec = importlib.util.spec_from_file_location("bioscanclip_ref_loss_func", OUT_FILE)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.ContrastiveLoss, mod.ClipLoss


if __name__ == "__main__":
    ok = build_ref(verbose=True)
    print("oracle/_ref", "ready" if ok else "unavailable (no /root/reference and no staged copy)")
