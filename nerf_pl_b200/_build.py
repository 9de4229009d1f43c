"""Build libnerf_b200.so (sm_100a only) with nvcc, in-tree, so the .so travels with the repo."""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.environ.get("NRF_LIBDIR") or os.path.join(HERE, "lib")   # NRF_LIBDIR: experiment builds beside the product library
LIB = os.path.join(LIBDIR, "libnerf_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
EXTRA = os.environ.get("NVCC_EXTRA", "").split()
# NRF_DEBUG_BUILD=1 (2: only the light per-warp trace points in the forward kernel) adds csrc/debug/*.cu (streaming / MMA-rate / store-rate microbenchmarks, the experimental CTA-pair
# forward, the forward kernels' cycle counters) and their entry points (include/nerf_b200_debug.h).  The product
# library has none of them and honours no debug environment variable.
DEBUG = os.environ.get("NRF_DEBUG_BUILD", "") not in ("", "0")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-I", CSRC] + ([f"-DNRF_FWD_DEBUG={os.environ['NRF_DEBUG_BUILD']}"] if DEBUG else []) + EXTRA


def _sources():
    src = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    if DEBUG:
        src += sorted(os.path.join("debug", f) for f in os.listdir(os.path.join(CSRC, "debug")) if f.endswith(".cu"))
    return src


def _digest():
    h = hashlib.sha256()
    files = [f for f in sorted(os.listdir(CSRC)) if os.path.isfile(os.path.join(CSRC, f))]
    if DEBUG:
        files += [os.path.join("debug", f) for f in sorted(os.listdir(os.path.join(CSRC, "debug")))]
    for f in files + ["../../include/nerf_b200.h"]:
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(f.encode() + fh.read())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    """Compile every csrc/*.cu and link lib/libnerf_b200.so.  Returns the path."""
    os.makedirs(LIBDIR, exist_ok=True)
    stamp = os.path.join(LIBDIR, "build.sha256")
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == dig:
        return LIB
    objs = []

    def cc(src):
        obj = os.path.join(LIBDIR, os.path.basename(src)[:-3] + ".o")
        cmd = [NVCC, *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr, file=sys.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(cc, _sources()))
    r = subprocess.run([NVCC, "-shared", "-o", LIB, *objs, "-lcudart", "-lcuda"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(stamp, "w") as f:
        f.write(dig)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
