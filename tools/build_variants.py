#!/usr/bin/env python
"""Builds libdash_b200.so with several sets of kernel tuning macros into
build_variants/<name>.so (A/B timing on the GPU box with tools/ab_variants.sh).

    python tools/build_variants.py            # all
    python tools/build_variants.py aes3 ...   # some"""
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dash_b200 import build  # noqa: E402

VARIANTS = {
    "base": "",
    "nofrac": "-DDASHB_FRAC=0",
    "aes3": "-DDASHB_AES_UNROLL=3",
    "sign9": "-DDASHB_SIGN_MINB=9",
    "sign10": "-DDASHB_SIGN_MINB=10",
    "rs10": "-DDASHB_RS_MINB=10",
    "rs12": "-DDASHB_RS_MINB=12",
    "sign64": "-DDASHB_SIGN_THREADS=64 -DDASHB_SIGN_MINB=16",
    "sign256": "-DDASHB_SIGN_THREADS=256 -DDASHB_SIGN_MINB=4",
    "hash_inline": "-DDASHB_HASH_INLINE=1",
}

if __name__ == "__main__":
    out = os.path.join(ROOT, "build_variants")
    os.makedirs(out, exist_ok=True)
    keep = os.path.join(out, "_default.so")
    if os.path.exists(build.LIB):
        shutil.copy(build.LIB, keep)
    for name in sys.argv[1:] or VARIANTS:
        os.environ["DASHB_NVCC_EXTRA"] = VARIANTS[name]
        build.build(force=True)
        shutil.copy(build.LIB, os.path.join(out, name + ".so"))
        print(name, VARIANTS[name], flush=True)
    os.environ["DASHB_NVCC_EXTRA"] = ""
    if os.path.exists(keep):
        shutil.move(keep, build.LIB)
