"""Compile the reference's own C++ searchsorted (CPU) into oracle/_ref/.  TEST INFRASTRUCTURE.

Sources are compiled WHERE THEY LIE under /root/reference/torchsearchsorted/src/cpu; nothing
is copied into this repository.  The reference file does not compile unmodified against
torch >= 2.x (`AT_DISPATCH_ALL_TYPES(a.type(), ...)`: DeprecatedTypeProperties no longer
converts to ScalarType, searchsorted_cpu_wrapper.cpp:100).  Instead of patching it we
compile a 6-line translation unit that includes torch's headers, re-points the dispatch
macro at `.scalarType()`, and then #includes the reference .cpp verbatim.

Output: oracle/_ref/tss_ref_cpu.so (git-ignored, travels to the GPU box with gpurun).
Only the build container has /root/reference; elsewhere this is a no-op."""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("NERF_REFERENCE", "/root/reference") + "/torchsearchsorted/src/cpu"


def build(verbose=False):
    out_dir = os.path.join(HERE, "_ref")
    so = os.path.join(out_dir, "tss_ref_cpu.so")
    src = os.path.join(REF_SRC, "searchsorted_cpu_wrapper.cpp")
    if not os.path.exists(src):
        return so if os.path.exists(so) else None
    if os.path.exists(so) and os.path.getmtime(so) > os.path.getmtime(__file__):
        return so
    os.makedirs(out_dir, exist_ok=True)
    import torch
    from torch.utils import cpp_extension
    tu = os.path.join(out_dir, "tss_ref_cpu_tu.cpp")
    with open(tu, "w") as f:
        f.write('#include <torch/extension.h>\n'
                '#undef AT_DISPATCH_ALL_TYPES\n'
                '#define AT_DISPATCH_ALL_TYPES(TYPE, NAME, ...) \\\n'
                '  AT_DISPATCH_SWITCH((TYPE).scalarType(), NAME, AT_DISPATCH_CASE_ALL_TYPES(__VA_ARGS__))\n'
                f'#include "{src}"\n')
    inc = [f"-I{p}" for p in cpp_extension.include_paths()] + [f"-I{sysconfig.get_paths()['include']}", f"-I{REF_SRC}"]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-w", "-DTORCH_EXTENSION_NAME=tss_ref_cpu",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           *inc, tu, "-o", so, f"-L{libdir}", "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_python",
           f"-Wl,-rpath,{libdir}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        if verbose:
            print(r.stderr[-3000:], file=sys.stderr)
        return None
    return so


if __name__ == "__main__":
    print(build(verbose=True))
