"""In-tree build of the native code (no JIT cache, so the .so files travel with the repo snapshot).

  lib/libsdpm.so     CUDA kernels + C ABI (include/sdpm.h), sm_100a only
  lib/libsdsynth.so  synthetic key-frame generator (input generation for tests / bench)
  lib/test_pm_class  C++ test driver of host/ProbabilityMapping (links libsdpm.so)

nvcc cross-compiles sm_100a without a GPU.  Numerics flags are part of the contract
(-fmad=false, IEEE div/sqrt, no flush-to-zero): see csrc/sdpm_kernels.cuh.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "lib")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math,-Wall",
    "-Xptxas", "-v",
]
# the image exports CXX=/opt/gcc/bin/g++ (a wrapper without libgomp); use the system g++
GXX = shutil.which("g++") or "g++"
HOST_FLAGS = ["-O3", "-march=x86-64-v3", "-std=c++17", "-fPIC", "-ffp-contract=off", "-fno-fast-math", "-Wall"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd, log_name=None):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log_name:
        with open(os.path.join(LIB, log_name), "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd))
    return r.stdout


def nvcc_path():
    for p in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if p and os.path.exists(p):
            return p
    raise RuntimeError("nvcc not found")


def build_sdpm(force=False):
    csrc = os.path.join(HERE, "csrc")
    srcs = [os.path.join(csrc, f) for f in ("sdpm_api.cu", "sdpm_kernels.cuh", "pair_geometry.h")]
    srcs.append(os.path.join(ROOT, "include", "sdpm.h"))
    out = os.path.join(LIB, "libsdpm.so")
    if force or _newer(out, srcs + [__file__]):
        os.makedirs(LIB, exist_ok=True)
        _run([nvcc_path()] + NVCC_FLAGS + os.environ.get("SDPM_NVCC_EXTRA", "").split() + ["-shared", "-o", out, os.path.join(csrc, "sdpm_api.cu")], "libsdpm.build.log")
    return out


def build_synth(force=False):
    src = os.path.join(HERE, "synth", "scene.cc")
    out = os.path.join(LIB, "libsdsynth.so")
    if force or _newer(out, [src]):
        os.makedirs(LIB, exist_ok=True)
        _run([GXX] + HOST_FLAGS + ["-fopenmp", "-shared", "-o", out, src])
    return out


def build_host_test(force=False):
    host = os.path.join(HERE, "host")
    srcs = [os.path.join(host, f) for f in ("ProbabilityMapping.cc", "ProbabilityMapping.h", "test_pm_class.cc")]
    if not all(os.path.exists(s) for s in srcs):
        return None
    out = os.path.join(LIB, "test_pm_class")
    lib = build_sdpm(force)
    if force or _newer(out, srcs + [lib]):
        _run([GXX] + HOST_FLAGS + ["-I", os.path.join(ROOT, "include"), "-o", out, srcs[0], srcs[2],
                                   "-L", LIB, "-lsdpm", "-Wl,-rpath,$ORIGIN"])
    return out


def build_all(force=False):
    return {"sdpm": build_sdpm(force), "synth": build_synth(force), "host_test": build_host_test(force)}


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv))
