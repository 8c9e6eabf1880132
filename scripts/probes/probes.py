"""Builds and loads scripts/probes/libvda_probes.so: measurement probes (tcgen05 descriptor semantics, mma.sync
issue rate) that are deliberately kept out of the product library."""
import ctypes as C
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
LIB = os.path.join(HERE, "libvda_probes.so")


def load():
    from vda_b200 import build as B
    import vda_b200.lib as L

    L.load()  # the probes link against the product library (set_error / check_cuda / encode_bf16_map)
    src = os.path.join(HERE, "umma_probe.cu")
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        subprocess.check_call([nvcc] + B.NVCC_FLAGS + [src, "-o", LIB, "-L" + B.HERE, "-l:" + os.path.basename(B.LIB),
                                                       "-Xlinker", "-rpath," + B.HERE, "-lcudart_static", "-ldl", "-lrt", "-lpthread"])
    lib = C.CDLL(LIB)
    vp, i = C.c_void_p, C.c_int
    lib.vda_debug_umma_shift.argtypes = [vp, vp, vp, i, i, i, vp]
    lib.vda_debug_mma_rate.argtypes = [vp, i, i, i, vp]
    lib.vda_debug_pipe_rate.argtypes = [vp, i, i, i, i, vp]
    return lib
