"""Build a variant of libsage_b200.so from an alternative attention source (development tool):
    python tools/build_variant.py NAME path/to/attn_variant.cu [extra nvcc flags]  ->  tools/_bin/libsage_NAME.so
Select it at run time with SAGE_B200_LIB=tools/_bin/libsage_NAME.so."""
import os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sageattention_b200 import build as B
name, src = sys.argv[1], os.path.abspath(sys.argv[2])
B.build()
out = os.path.join(ROOT, "tools", "_bin"); os.makedirs(out, exist_ok=True)
tmp = os.path.join(B.CSRC, f"_variant_{name}.cu"); shutil.copy(src, tmp)
obj = os.path.join(out, f"attn_{name}.o")
try:
    subprocess.check_call([B._nvcc(), "-c", tmp, "-o", obj] + B.ARCH + B.COMMON + sys.argv[3:])
finally:
    os.remove(tmp)
# the variant replaces the unit(s) whose entry points it defines (older experiments define both)
text = open(src).read()
drop = {"attn_f8_sm100.cu"} if 'int sage_attn_i8_f8(' in text else set()
if 'int sage_attn_i8_f16(' in text:
    drop.add("attn_f16_sm100.cu")
objs = [os.path.join(B.OBJ, n + ".o") for n, _ in B.UNITS if n not in drop] + [obj]
lib = os.path.join(out, f"libsage_{name}.so")
subprocess.check_call([B._nvcc(), "-shared", "-o", lib] + objs + B.ARCH)
print(lib)
