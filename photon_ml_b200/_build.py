"""Builds photon_ml_b200/lib/libphoton_ml_b200.so with nvcc for sm_100a (in-tree, so the .so
travels to the GPU box with the repository snapshot)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIB_DIR, "libphoton_ml_b200.so")
STAMP = os.path.join(LIB_DIR, "build.sha256")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-O2", "--expt-relaxed-constexpr"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _digest(paths, flags):
    """Content hash of what a build depends on: the sources, the headers and the compiler flags (modification times
    say nothing after a checkout or a copy: a binary that merely LOOKS newer must not be reused)."""
    import hashlib
    h = hashlib.sha256(" ".join(flags).encode())
    for path in sorted(paths):
        h.update(os.path.basename(path).encode())
        with open(path, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def build_library(force=False, verbose=False):
    os.makedirs(LIB_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(os.path.dirname(HERE), "include", "photon_ml_b200.h"))
    sources = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cu")]
    want = _digest(sources + headers, NVCC_FLAGS)
    have = open(STAMP).read().strip() if os.path.exists(STAMP) else ""
    if not force and not verbose and os.path.exists(LIB) and have == want:
        return LIB
    # every object is rebuilt whenever anything changed: five files, compiled in parallel in well under a minute
    objs, procs = [], []
    for path in sources:
        obj = os.path.join(LIB_DIR, os.path.basename(path)[:-3] + ".o")
        objs.append(obj)
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
        procs.append((os.path.basename(path), subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write("---- %s ----\n%s\n" % (src, out))
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    subprocess.check_call([_nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
    with open(STAMP, "w") as f:
        f.write(want + "\n")
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
