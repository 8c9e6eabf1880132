"""Builds vda_b200/libvda_b200.so (hand-written sm_100a kernels + C ABI) with nvcc, in-tree."""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, os.environ.get("VDA_LIB_NAME", "libvda_b200.so"))  # VDA_LIB_NAME: A/B builds for tuning
MARK = b"vda-src-digest:"   # followed by the 64 hex digits of _digest() inside the binary (csrc/api.cu: vda_build_digest)
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
] + os.environ.get("VDA_NVCC_EXTRA", "").split()


def _sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest():
    h = hashlib.sha256()
    for f in sorted(os.listdir(CSRC)) + ["../../include/vda_b200.h"]:
        p = os.path.join(CSRC, f)
        if os.path.isfile(p):
            h.update(f.encode())
            h.update(open(p, "rb").read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def binary_digest(path=None):
    """The source digest compiled into an existing library file, read without loading it (None if absent)."""
    try:
        blob = open(path or LIB, "rb").read()
    except OSError:
        return None
    at = blob.find(MARK)
    return blob[at + len(MARK):at + len(MARK) + 64].decode("ascii", "replace") if at >= 0 else None


def build(force=False, verbose=False):
    """Compile every .cu under csrc/ into one shared library.  Returns the library path.
    The digest of the sources + flags is compiled INTO the library, so a library left over from other sources is
    detected from the file itself (there is no side stamp file that could travel without its binary)."""
    digest = _digest()
    if not force and binary_digest() == digest:
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    for src in _sources():
        obj = os.path.join(HERE, "build", os.path.basename(src)[:-3] + ".o")
        cmd = [nvcc] + [f for f in NVCC_FLAGS if f != "-shared"] + ["-c", src, "-o", obj]
        if os.path.basename(src) == "api.cu":
            cmd.append(f'-DVDA_SRC_DIGEST="{digest}"')
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError(f"nvcc failed on {src}")
    subprocess.check_call([nvcc, "-shared", "-o", LIB] + objs + ["-lcudart_static", "-ldl", "-lrt", "-lpthread"])
    if binary_digest() != digest:
        raise RuntimeError("built library does not carry its source digest")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
