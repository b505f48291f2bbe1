"""Build a compile-time-flag variant of the two attention units (development tool):
    python tools/build_flags.py NAME [-DSAGE_FUSE=2 -DSAGE_MAXTREE=1 ...]   ->  tools/_bin/libsage_NAME.so
Select it at run time with SAGE_B200_LIB=tools/_bin/libsage_NAME.so (tools/sweep_variants.py does)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sageattention_b200 import build as B  # noqa: E402


def build_variant(name, flags, only=None):
    """`only`: apply the flags to this attention unit alone (e.g. -DSAGE_TRACE, whose dump symbol both units define)."""
    B.build()
    out = os.path.join(ROOT, "tools", "_bin")
    os.makedirs(out, exist_ok=True)
    objs, procs = [], []
    for unit, extra in B.UNITS:
        if (only is None and unit.startswith("attn_")) or unit == only:
            obj = os.path.join(out, f"{name}_{unit}.o")
            cmd = [B._nvcc(), "-c", os.path.join(B.CSRC, unit), "-o", obj] + B.ARCH + B.COMMON + extra + flags
            procs.append(subprocess.Popen(cmd))
        else:
            obj = os.path.join(B.OBJ, unit + ".o")
        objs.append(obj)
    for p in procs:
        if p.wait() != 0:
            raise SystemExit("nvcc failed")
    lib = os.path.join(out, f"libsage_{name}.so")
    subprocess.check_call([B._nvcc(), "-shared", "-o", lib] + objs + B.ARCH)
    return lib


if __name__ == "__main__":
    args = sys.argv[2:]
    only = None
    if "--only" in args:
        i = args.index("--only")
        only = args[i + 1]
        del args[i:i + 2]
    print(build_variant(sys.argv[1], args, only))
