"""Stage the UNMODIFIED reference python package under oracle/_ref/ (TEST INFRASTRUCTURE; git-ignored, travels to the
GPU box with the gpurun snapshot like the built .so files).

    python -m oracle.build_ref          # needs /root/reference; writes oracle/_ref/multimeditron/{model,dataset,utils}
                                        # and oracle/_ref/mock_dataset/*.jpg

The reference is pure Python: nothing is compiled, the files are copied byte for byte so that the
`-m gpu` tests and `bench.py --impl torch_gpu` can run the reference's own `MultiModalModelForCausalLM`,
`DataCollatorForMultimodal` and `ImageModality` on the B200 (where /root/reference does not exist).  The package root
`__init__.py` and `cli/`, `train/`, `ui/` are NOT staged (they import ray / deepspeed / hydra, absent from the image);
oracle/ref_shim.py seeds the package module instead.  Nothing under multimeditron_b200/ imports oracle/_ref.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

SRC = "/root/reference/src/multimeditron"
MOCK = "/root/reference/mock_dataset"
HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
PARTS = ("model", "dataset", "utils")


def build(verbose: bool = True) -> bool:
    """-> True when oracle/_ref is (now) populated.  A no-op without /root/reference (the GPU box uses the staged copy)."""
    if not os.path.isdir(SRC):
        return os.path.isdir(os.path.join(DST, "multimeditron", "model"))
    manifest = {}
    pkg = os.path.join(DST, "multimeditron")
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    os.makedirs(pkg)
    for part in PARTS:
        for root, _dirs, files in os.walk(os.path.join(SRC, part)):
            for f in files:
                if not f.endswith(".py"):
                    continue
                s = os.path.join(root, f)
                rel = os.path.relpath(s, SRC)
                d = os.path.join(pkg, rel)
                os.makedirs(os.path.dirname(d), exist_ok=True)
                shutil.copyfile(s, d)
                manifest["multimeditron/" + rel] = hashlib.sha256(open(s, "rb").read()).hexdigest()
    if os.path.isdir(MOCK):
        os.makedirs(os.path.join(DST, "mock_dataset"), exist_ok=True)
        for f in sorted(os.listdir(MOCK)):
            if f.lower().endswith((".jpg", ".jpeg", ".png")):
                shutil.copyfile(os.path.join(MOCK, f), os.path.join(DST, "mock_dataset", f))
                manifest["mock_dataset/" + f] = hashlib.sha256(open(os.path.join(MOCK, f), "rb").read()).hexdigest()
    json.dump(manifest, open(os.path.join(DST, "MANIFEST.json"), "w"), indent=1, sort_keys=True)
    if verbose:
        print(f"staged {len(manifest)} reference files under {DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if build() else 1)
