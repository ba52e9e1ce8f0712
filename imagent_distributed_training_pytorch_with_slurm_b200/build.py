"""Build the C-ABI shared library (in-tree) with nvcc for sm_100a.

    python -m imagent_distributed_training_pytorch_with_slurm_b200.build

nvcc cross-compiles without a GPU.  The product is ``libb200_imagenet.so`` next to this file; it is
git-ignored but travels to the GPU box with the working-tree snapshot.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libb200_imagenet.so")
SOURCES = ["api.cu", "sgd.cu", "ce_topk.cu", "batchnorm.cu", "input.cu", "maxpool.cu", "pack.cu", "peer_exchange.cu"]
# the JPEG decode glue links libnvjpeg and therefore lives in a library of its own (loaded only for decoder="nvjpeg")
JPEG_LIB = os.path.join(HERE, "libb200_jpeg.so")
JPEG_SOURCES = ["jpeg.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "--shared",
              "-Xcompiler", "-fPIC", "-cudart", "shared"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; the CUDA library cannot be built")


def needs_build():
    if not os.path.exists(LIB) or not os.path.exists(JPEG_LIB):
        return True
    t = min(os.path.getmtime(LIB), os.path.getmtime(JPEG_LIB))
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", h)
                                                                 for h in ("b200_imagenet.h", "b200_jpeg.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          [os.path.join(CSRC, s) for s in SOURCES] + ["-o", LIB + ".tmp"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    os.replace(LIB + ".tmp", LIB)
    if verbose:
        sys.stderr.write(r.stderr)
    cmd = [_nvcc()] + NVCC_FLAGS + [os.path.join(CSRC, s) for s in JPEG_SOURCES] + ["-lnvjpeg", "-o", JPEG_LIB + ".tmp"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    os.replace(JPEG_LIB + ".tmp", JPEG_LIB)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose="-v" in sys.argv))
