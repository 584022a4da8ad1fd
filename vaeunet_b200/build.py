"""Build libvaeunet_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m vaeunet_b200.build [--force] [--verbose]

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
OUT = PKG / "libvaeunet_b200.so"
OBJ = PKG / "build"
SOURCES = ["vu_api.cu", "vu_conv_igemm.cu", "vu_conv_halo.cu", "vu_conv_flat.cu", "vu_conv_wgrad.cu", "vu_layout.cu", "vu_small_cin_mma.cu", "vu_gate_latent.cu", "vu_decoder_concat.cu",
           "vu_uncertainty.cu", "vu_sort_metrics.cu", "vu_tiles.cu", "vu_loss.cu", "vu_train.cu", "vu_train_bwd.cu", "vu_train_small.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "--use_fast_math" if False else "-DVU_NO_FAST_MATH",
         "-I", str(PKG.parent / "include"), "-I", str(CSRC)]


def _sources():
    return [s for s in SOURCES if (CSRC / s).exists()]


def _stamp() -> str:
    h = hashlib.sha256()
    for f in sorted(list(CSRC.glob("*")) + [PKG.parent / "include" / "vaeunet_b200.h", Path(__file__)]):
        h.update(f.name.encode())
        h.update(f.read_bytes())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> Path:
    stamp_file = OBJ / "stamp.txt"
    stamp = _stamp()
    if not force and OUT.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return OUT
    OBJ.mkdir(exist_ok=True)
    srcs = _sources()

    def compile_one(src: str):
        obj = OBJ / (src.replace(".cu", ".o"))
        cmd = [NVCC, *FLAGS, "-c", str(CSRC / src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(compile_one, srcs))
    cmd = [NVCC, "-shared", "-o", str(OUT), *map(str, objs), "-cudart", "static",
           "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp_file.write_text(stamp)
    return OUT


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(p)
