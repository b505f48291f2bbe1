"""TEST INFRASTRUCTURE (oracle/): build the reference's own CUDA sources as GPU oracles.

Compiles, *from where they lie under /root/reference* (nothing is copied into this repo):

  _ref_fused : csrc/fused/fused.cu                      (+ oracle/ref_bind_fused.cpp)
  _ref_sm80  : csrc/qattn/qk_int_sv_f16_cuda_sm80.cu    (+ csrc/qattn/pybind_sm80.cpp)
  _ref_sm89  : csrc/qattn/sm89_*.cu (7 files)           (+ csrc/qattn/pybind_sm89.cpp)

with the reference's NVCC flags (/root/reference/setup.py:45-54) plus
``-Dhip_bfloat16=nv_bfloat16`` (the fork maps BFloat16 to a HIP type,
/root/reference/csrc/dispatch_utils.h:76-88) and ``-gencode arch=compute_100,code=sm_100``
(the reference has no sm_100 target of its own, setup.py:41).

Outputs go ONLY to oracle/_ref/ (git-ignored, travels to the GPU box with gpurun).  The
product (sageattention_b200/) never imports these; only tests/ and bench.py's comparison
legs do, as checkers / baselines.

Usage:  python oracle/build_ref.py [fused] [sm80] [sm89]      (default: all three)
"""
import os
import shutil
import subprocess
import sys
import sysconfig
import time

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("SAGE_REFERENCE_ROOT", "/root/reference")
CSRC = os.path.join(REF, "csrc")

NVCC_FLAGS = [
    "-O3", "-std=c++17",
    "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__",
    "--use_fast_math", "--threads=4", "-diag-suppress=174", "-w",
    "-Dhip_bfloat16=nv_bfloat16",
    "-gencode", "arch=compute_100,code=sm_100",
    "--expt-relaxed-constexpr", "--expt-extended-lambda",
    "-Xcompiler", "-fPIC",
]

SM89_FILES = [
    "sm89_qk_int8_sv_f8_accum_f32_attn.cu",
    "sm89_qk_int8_sv_f8_accum_f32_fuse_v_scale_attn.cu",
    "sm89_qk_int8_sv_f8_accum_f32_fuse_v_scale_fuse_v_mean_attn.cu",
    "sm89_qk_int8_sv_f8_accum_f32_attn_inst_buf.cu",
    "sm89_qk_int8_sv_f8_accum_f16_attn_inst_buf.cu",
    "sm89_qk_int8_sv_f8_accum_f32_fuse_v_scale_attn_inst_buf.cu",
    "sm89_qk_int8_sv_f8_accum_f16_fuse_v_scale_attn_inst_buf.cu",
]

TARGETS = {
    "fused": dict(name="_ref_fused",
                  cu=[os.path.join(CSRC, "fused", "fused.cu")],
                  cpp=[os.path.join(HERE, "ref_bind_fused.cpp")]),
    "sm80": dict(name="_ref_sm80",
                 cu=[os.path.join(CSRC, "qattn", "qk_int_sv_f16_cuda_sm80.cu")],
                 cpp=[os.path.join(CSRC, "qattn", "pybind_sm80.cpp")]),
    "sm89": dict(name="_ref_sm89",
                 cu=[os.path.join(CSRC, "qattn", f) for f in SM89_FILES],
                 cpp=[os.path.join(CSRC, "qattn", "pybind_sm89.cpp")]),
}


def _torch_paths():
    import torch
    from torch.utils import cpp_extension as ce
    inc = ce.include_paths("cuda")
    lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = 1 if torch._C._GLIBCXX_USE_CXX11_ABI else 0
    return inc, lib, abi


def _run(cmd):
    t = time.time()
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if p.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + p.stdout[-6000:] + "\n")
        raise SystemExit(f"build failed: {cmd[0]} rc={p.returncode}")
    return time.time() - t


def build(which):
    if not os.path.isdir(REF):
        print(f"[build_ref] {REF} not present (GPU box?) - using prebuilt oracle/_ref if any")
        return
    tgt = TARGETS[which]
    name = tgt["name"]
    os.makedirs(os.path.join(OUT, "obj"), exist_ok=True)
    so = os.path.join(OUT, name + ".so")
    srcs = tgt["cu"] + tgt["cpp"]
    if os.path.exists(so) and all(os.path.getmtime(so) > os.path.getmtime(s) for s in srcs):
        print(f"[build_ref] {name}: up to date")
        return
    inc, lib, abi = _torch_paths()
    incs = [f"-I{p}" for p in inc] + [f"-I{sysconfig.get_paths()['include']}",
                                      f"-I{os.path.join(CSRC, 'qattn')}", f"-I{os.path.join(CSRC, 'fused')}"]
    defs = [f"-DTORCH_EXTENSION_NAME={name}", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
            "-DTORCH_API_INCLUDE_EXTENSION_H"]
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    objs = []
    # compile the .cu files a few at a time (each nvcc already uses --threads=4)
    procs = []
    for s in tgt["cu"]:
        o = os.path.join(OUT, "obj", os.path.basename(s) + ".o")
        objs.append(o)
        cmd = [nvcc, "-c", s, "-o", o] + NVCC_FLAGS + incs + defs
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        if len(procs) >= 3:
            _drain(procs)
    _drain(procs)
    for s in tgt["cpp"]:
        o = os.path.join(OUT, "obj", name + "_" + os.path.basename(s) + ".o")
        objs.append(o)
        dt = _run(["g++", "-c", s, "-o", o, "-O2", "-std=c++17", "-fPIC", "-w"] + incs + defs)
    dt = _run(["g++", "-shared", "-o", so] + objs +
              [f"-L{lib}", "-L/usr/local/cuda/lib64", "-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_python",
               "-lc10_cuda", "-ltorch_cuda", "-lcudart", f"-Wl,-rpath,{lib}"])
    print(f"[build_ref] {name}: built {so}")


def _drain(procs):
    while procs:
        s, p = procs.pop(0)
        t = time.time()
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out[-6000:])
            raise SystemExit(f"nvcc failed on {s}")
        print(f"[build_ref] compiled {os.path.basename(s)}")


# The reference's Triton quantisers / varlen kernels are Python: "building" them for the GPU box means staging
# the files, unmodified, next to the compiled .so files (oracle/_ref/stage/, git-ignored like the rest of
# oracle/_ref, travels with gpurun).  Triton JIT-compiles them on the box.  Checker only, never product.
TRITON_FILES = ["quant_per_thread.py", "quant_per_block.py", "quant_per_block_varlen.py",
                "attn_qk_int8_block_varlen.py", "attn_qk_int8_per_block_causal_varlen.py",
                "attn_qk_int8_per_block.py", "attn_qk_int8_per_block_causal.py",
                "attn_qk_int8_per_thread.py", "attn_qk_int8_per_thread_causal.py"]


def stage_triton():
    src_dir = os.path.join(REF, "sageattention", "triton")
    if not os.path.isdir(src_dir):
        print(f"[build_ref] {src_dir} not present (GPU box?) - using staged files if any")
        return
    dst_dir = os.path.join(OUT, "stage")
    os.makedirs(dst_dir, exist_ok=True)
    for f in TRITON_FILES:
        shutil.copyfile(os.path.join(src_dir, f), os.path.join(dst_dir, f))
    print(f"[build_ref] staged {len(TRITON_FILES)} Triton reference files -> {dst_dir}")


def load_triton(stem):
    """Import a staged reference Triton module (e.g. "quant_per_thread"); None if not staged."""
    import importlib.util
    path = os.path.join(OUT, "stage", stem + ".py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("ref_triton_" + stem, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load(which):
    """Import a prebuilt oracle module from oracle/_ref (returns None if absent)."""
    import importlib.util
    name = TARGETS[which]["name"]
    so = os.path.join(OUT, name + ".so")
    if not os.path.exists(so):
        return None
    import torch  # noqa: F401  (libtorch must be loaded first)
    spec = importlib.util.spec_from_file_location(name, so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    todo = sys.argv[1:] or ["fused", "sm80", "sm89", "triton"]
    for w in todo:
        if w == "triton":
            stage_triton()
        else:
            build(w)
