"""Build libsage_b200.so (sm_100a only) in-tree with nvcc.

    python -m sageattention_b200.build [--force]

The library has no torch / pybind dependency: it is a plain C-ABI shared object
(include/sage_b200.h) that the Python layer binds with ctypes.
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
ROOT = os.path.dirname(PKG)
LIB = os.path.join(PKG, "libsage_b200.so")
OBJ = os.path.join(PKG, "csrc", "_obj")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-w", f"-I{os.path.join(ROOT, 'include')}"]
# (source, extra flags).  prepass.cu is built with --use_fast_math ON PURPOSE: the reference's
# quantizers are (setup.py:45-54) and bit-exact codes need the same div.approx/.ftz sequences.
UNITS = [
    ("capi.cu", []),
    ("prepass.cu", ["--use_fast_math"]),
    ("attn_f16_sm100.cu", []),
    ("attn_f8_sm100.cu", []),
    ("fwd.cu", []),
]


def source_hash():
    """sha256 over the library's sources (csrc/*.cu, *.cuh, *.h and include/sage_b200.h), in name order."""
    import hashlib
    h = hashlib.sha256()
    files = sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h")) and not f.startswith("_variant_"))
    for f in files:
        h.update(f.encode())
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(fh.read())
    with open(os.path.join(ROOT, "include", "sage_b200.h"), "rb") as fh:
        h.update(fh.read())
    return h.hexdigest()


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def _deps(src):
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(ROOT, "include", "sage_b200.h"))
    return [src] + hdrs


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    objs, rebuilt = [], False
    procs = []
    digest = source_hash()
    stamp = os.path.join(OBJ, "capi.hash")
    for name, extra in UNITS:
        src = os.path.join(CSRC, name)
        obj = os.path.join(OBJ, name + ".o")
        objs.append(obj)
        stale = force or not os.path.exists(obj) or any(
            os.path.getmtime(d) > os.path.getmtime(obj) for d in _deps(src))
        if name == "capi.cu":
            # carries the hash of ALL sources: rebuilt whenever that changes
            extra = extra + [f'-DSAGE_SRC_HASH="{digest}"']
            stale = stale or not os.path.exists(stamp) or open(stamp).read() != digest
        if stale:
            cmd = [_nvcc(), "-c", src, "-o", obj] + ARCH + COMMON + extra
            if verbose:
                cmd += ["-Xptxas", "-v"]
            procs.append((name, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
            rebuilt = True
    for name, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError(f"nvcc failed on {name}")
        if verbose:
            print(out)
    with open(stamp, "w") as fh:
        fh.write(digest)
    if rebuilt or not os.path.exists(LIB):
        cmd = [_nvcc(), "-shared", "-o", LIB] + objs + ARCH
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if p.returncode != 0:
            sys.stderr.write(p.stdout)
            raise RuntimeError("link failed")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
