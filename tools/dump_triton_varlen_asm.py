"""Compile the reference's Triton varlen quantiser on the GPU box and print how its arithmetic was lowered
(development tool: which instructions the two divisions and the rounding became decides bit-exactness of
sage_quant_per_block_varlen_int8)."""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from oracle import build_ref  # noqa: E402

m = build_ref.load_triton("quant_per_block_varlen")
q = torch.randn(300, 2, 128, device="cuda", dtype=torch.float16)
k = torch.randn(300, 2, 128, device="cuda", dtype=torch.float16)
cu = torch.tensor([0, 100, 300], device="cuda", dtype=torch.int32)
out = m.per_block_int8(q, k, cu, cu, 200, 200, sm_scale=0.0883)
torch.cuda.synchronize()
fn = m.quant_per_block_int8_kernel
caches = getattr(fn, "device_caches", None) or {}
n = 0
for dev, tup in caches.items():
    for key, ck in tup[0].items():
        n += 1
        ptx = ck.asm.get("ptx", "")
        print(f"==== compiled kernel {n}: PTX arithmetic lines")
        for line in ptx.splitlines():
            if re.search(r"\b(div|rcp|mul\.f32|mul\.rn|fma|cvt\.r|max\.|abs\.|selp|copysign|and\.b32|or\.b32|add\.f32|mov\.f32|mov\.b32.*0f)", line):
                print("   ", line.strip())
        sass = ck.asm.get("sass", None)
        if sass is None:
            try:
                import subprocess, tempfile
                with tempfile.NamedTemporaryFile(suffix=".cubin", delete=False) as f:
                    f.write(ck.asm["cubin"])
                sass = subprocess.run(["cuobjdump", "-sass", f.name], capture_output=True, text=True).stdout
            except Exception as e:
                sass = f"(no sass: {e})"
        print(f"==== compiled kernel {n}: SASS arithmetic lines")
        for line in sass.splitlines():
            if re.search(r"(MUFU|FMUL|FFMA|FADD|F2I|I2F|FMNMX|FSEL|FSETP|LOP3|FCHK|BRA|CALL)", line):
                print("   ", re.sub(r"/\*[0-9a-f]{4,}\*/", "", line).split(";")[0].strip())
print("kernels:", n)
