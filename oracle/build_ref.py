"""TEST INFRASTRUCTURE — build recipe for oracle/_ref: the reference's OWN CPU implementation of the
hot path, compiled from the sources where they lie under /root/reference (never copied here).

    python oracle/build_ref.py        ->  oracle/_ref/libsgl_cpu_ref.so

Sources: sgl-kernel/csrc/cpu/{gemm.cpp, gemm_int8.cpp, gemm_fp8.cpp, decode.cpp, extend.cpp} (+ their headers
common.h / gemm.h / vec.h) and the registration shim oracle/ref_register.cpp.  Plain g++ commands — the
reference's own CMake build is not run (it needs libnuma and builds every op).  The kernels need
AVX512F + AVX512-BF16 + AMX-BF16 (sgl-kernel/csrc/cpu/vec.h:3-5); both the build container and the GPU
box's host have them (gpurun_out/cpuflags.txt).  oracle/_ref/ is git-ignored but travels with gpurun.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_ref")
LIB = os.path.join(OUT_DIR, "libsgl_cpu_ref.so")
REF_CPU = "/root/reference/sgl-kernel/csrc/cpu"
SOURCES = ["gemm.cpp", "gemm_int8.cpp", "gemm_fp8.cpp", "decode.cpp", "extend.cpp"]
ISA = ["-mavx512f", "-mavx512bw", "-mavx512vl", "-mavx512dq", "-mavx512vnni", "-mavx512bf16", "-mavx512fp16",
       "-mamx-tile", "-mamx-bf16", "-mamx-int8", "-mfma", "-mf16c", "-mavx2"]


def reference_sources_available() -> bool:
    return all(os.path.exists(os.path.join(REF_CPU, s)) for s in SOURCES)


def build(verbose: bool = False) -> str | None:
    """Returns the library path, or None when /root/reference is absent (GPU box: prebuilt file is used)."""
    if not reference_sources_available():
        return LIB if os.path.exists(LIB) else None
    import torch
    from torch.utils import cpp_extension

    os.makedirs(OUT_DIR, exist_ok=True)
    srcs = [os.path.join(REF_CPU, s) for s in SOURCES] + [os.path.join(HERE, "ref_register.cpp")]
    newest = max(os.path.getmtime(s) for s in srcs)
    if os.path.exists(LIB) and os.path.getmtime(LIB) >= newest:
        return LIB
    inc = []
    for p in cpp_extension.include_paths():
        inc += ["-isystem", p]
    import sysconfig

    inc += ["-isystem", sysconfig.get_paths()["include"]]
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    flags = ["-O3", "-std=c++17", "-fPIC", "-fopenmp", "-Wno-unknown-pragmas", "-DSGLANG_CPU_FP8_CVT_FTZ",
             f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "-DTORCH_API_INCLUDE_EXTENSION_H"] + ISA + inc + ["-I", REF_CPU]

    def cc(src):
        obj = os.path.join(OUT_DIR, os.path.basename(src).replace(".cpp", ".o"))
        cmd = ["g++"] + flags + ["-c", src, "-o", obj]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"g++ failed for {src}:\n{r.stderr[-4000:]}")
        return obj

    with ThreadPoolExecutor(max_workers=5) as ex:
        objs = list(ex.map(cc, srcs))
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-shared", "-o", LIB] + objs + ["-fopenmp", f"-L{libdir}", "-ltorch", "-ltorch_cpu", "-lc10",
                                                   f"-Wl,-rpath,{libdir}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stderr[-4000:]}")
    for o in objs:
        os.remove(o)
    return LIB


_LOADED = False


def load():
    """torch.ops.b200_ref.* or None if the library is absent / the host CPU lacks the ISA."""
    global _LOADED
    import torch

    if _LOADED:
        return torch.ops.b200_ref
    if not os.path.exists(LIB):
        return None
    flags = open("/proc/cpuinfo").read()
    if not all(f in flags for f in ("avx512f", "avx512_bf16", "amx_bf16", "amx_int8", "avx512_fp16")):
        return None
    torch.ops.load_library(LIB)
    _LOADED = True
    return torch.ops.b200_ref


if __name__ == "__main__":
    print(build(verbose=True))
