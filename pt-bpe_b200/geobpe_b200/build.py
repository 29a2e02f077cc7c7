"""In-tree build of the sm_100a CUDA library (explicit nvcc; no JIT cache, no torch extension).

    python -m geobpe_b200.build            # builds pt-bpe_b200/geobpe_b200/libgeobpe_b200.so
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(PKG))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libgeobpe_b200.so")
SOURCES = ["k_select.cu", "k3_binning.cu", "k1_paircount.cu", "k1_slots.cu", "k_hloop.cu", "k_hloop2.cu", "k_induce.cu", "k_export.cu", "k_recon.cu", "k2_merge.cu", "k45_geometry.cu", "k6_pam.cu"]


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found; the geobpe_b200 kernels cannot be built")
    return p


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(ROOT, "include", "geobpe_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """one nvcc -c per source (in parallel), then one link; objects live in a git-ignored build/ directory"""
    if not force and not needs_build():
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    objdir = os.path.join(PKG, "build")
    os.makedirs(objdir, exist_ok=True)
    base = [nvcc_path(), "-Xcompiler", "-fPIC", "-O3", "-std=c++17", "-lineinfo",
            "-gencode", "arch=compute_100a,code=sm_100a", "-fmad=false",
            "-I", os.path.join(ROOT, "include"), "-I", CSRC]
    if verbose:
        base += ["-Xptxas", "-v"]
    hdr_t = max(os.path.getmtime(os.path.join(CSRC, f)) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h")))
    hdr_t = max(hdr_t, os.path.getmtime(os.path.join(ROOT, "include", "geobpe_b200.h")))

    def one(src):
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        path = os.path.join(CSRC, src)
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(path), hdr_t):
            return obj, ""
        r = subprocess.run(base + ["-c", path, "-o", obj], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n" + r.stdout + r.stderr)
        return obj, r.stdout + r.stderr
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        res = list(ex.map(one, SOURCES))
    r = subprocess.run([nvcc_path(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a"] + [o for o, _ in res] + ["-o", LIB],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    if verbose:
        print("".join(t for _, t in res))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
