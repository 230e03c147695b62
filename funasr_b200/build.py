"""Build libfunasr_b200.so in-tree with nvcc for sm_100a (no torch extension machinery: the
boundary is a plain C-ABI, see include/funasr_b200.h).  `python -m funasr_b200.build`."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libfunasr_b200.so")
COMPAT_LIB = os.path.join(HERE, "libfunasr_b200_compat.so")
TORCH_LIB = os.path.join(HERE, "libfunasr_b200_torch.so")
SOURCES = ["api.cu", "frontend.cu", "rowwise.cu", "gemm_sm100.cu", "attention_sm100.cu", "cif_decode.cu", "conformer.cu", "relpos_attention_sm100.cu", "bicif.cu", "runtime.cu", "stack_sm100.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]


def _nvcc() -> str:
    cand = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "bin", "nvcc")
    return cand if os.path.exists(cand) else "nvcc"


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    headers = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "gemm_host.inc"), os.path.join(CSRC, "attention_host.inc"), os.path.join(CSRC, "attention_softmax.inc"), os.path.join(CSRC, "attention_output.inc"), os.path.join(CSRC, "sm100_ptx.cuh"), os.path.join(CSRC, "stack.cuh"), os.path.join(os.path.dirname(HERE), "include", "funasr_b200.h")]
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    objs, jobs = [], []
    for s in SOURCES:
        src, obj = os.path.join(CSRC, s), os.path.join(objdir, s[:-3] + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            jobs.append([_nvcc()] + NVCC_FLAGS + ["-c", src, "-o", obj])

    def run(cmd):
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n%s\n%s" % (" ".join(cmd), r.stdout + r.stderr))

    with ThreadPoolExecutor(max_workers=6) as ex:
        list(ex.map(run, jobs))
    if jobs or force or _stale(LIB, objs):
        run([_nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
    # the reference's C++ offline-stream API (funasrruntime.h names and signatures) over the library: plain C++, no CUDA
    compat_src = os.path.join(CSRC, "funasr_compat.cpp")
    if force or _stale(COMPAT_LIB, [compat_src, LIB, headers[-1]]):
        run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", compat_src, "-o", COMPAT_LIB, "-L" + HERE,
             "-lfunasr_b200", "-Wl,-rpath,$ORIGIN"])
    # TORCH_LIBRARY(funasr_b200, ...): the torch custom-op layer over the C-ABI (csrc/torch_shim.cpp; g++ with torch's headers)
    shim_src = os.path.join(CSRC, "torch_shim.cpp")
    if force or _stale(TORCH_LIB, [shim_src, LIB, headers[-1]]):
        import torch

        tdir = os.path.dirname(torch.__file__)
        cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
        run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden",
             "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI), "-DUSE_CUDA",
             "-I" + os.path.join(tdir, "include"), "-I" + os.path.join(tdir, "include", "torch", "csrc", "api", "include"),
             "-I" + cuda_inc, shim_src, "-o", TORCH_LIB, "-L" + HERE, "-lfunasr_b200", "-L" + os.path.join(tdir, "lib"),
             "-ltorch", "-ltorch_cpu", "-lc10", "-lc10_cuda", "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + os.path.join(tdir, "lib")])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
