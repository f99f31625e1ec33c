"""Build libccm_orb_cuda.so (hand-written CUDA for sm_100a) in-tree with nvcc. No torch, no JIT cache."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libccm_orb_cuda.so")
SRC = os.path.join(HERE, "csrc", "ccm_orb_cuda.cu")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def sources():
    d = os.path.join(HERE, "csrc")
    out = [os.path.join(d, f) for f in sorted(os.listdir(d))]
    out.append(os.path.join(HERE, "..", "include", "ccm_orb.h"))
    out.append(os.path.join(HERE, "..", "include", "ccm_match.h"))
    return out


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(s) > t for s in sources() if os.path.exists(s))


def build(force=False, verbose=False):
    """Compile the CUDA library if it is missing or older than its sources. Returns the library path."""
    if not force and not stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("CCM_NVCC_EXTRA", "").split()          # experiments only, e.g. -DFAST_RESIDENT_THREADS=1536
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB, SRC]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
