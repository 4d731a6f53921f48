"""In-tree build of the CUDA extension (libffb200.so) for sm_100a.

`python -m flexflow_b200.build` or `flexflow_b200.build.build()`.  The .so is git-ignored but travels
to the GPU box with the repo snapshot.  No torch dependency: plain nvcc -shared.
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libffb200.so")
SOURCES = [
    "attn_prep.cu",
    "attn_decode.cu",
    "attn_multi.cu",
    "attn_tc.cu",
    "capi_attn.cu",
    "dequant.cu",
    "linear_stream.cu",
    "linear_gemv.cu",
    "linear_qgemv.cu",
    "linear_tc.cu",
    "linear_tc2.cu",
    "linear_tc4.cu", "linear_sk.cu", "linear_static.cu", "linear_qtc.cu", "linear_w512.cu", "linear_tcc.cu",
    "linear_tcq.cu",
    "capi_linear.cu",
    "comm.cu",
    "peer_allreduce.cu",
    "norm_argmax.cu",
    "host_tree.cu",
    "request_manager.cu",
    "weights_io.cu",
]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _digest(paths):
    h = hashlib.sha256()
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    """Compile every .cu under csrc/ into one shared library. Returns the .so path."""
    objdir = os.path.join(HERE, "_obj")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers += [os.path.join(HERE, "..", "include", f) for f in ("ffb200.h", "ffb200_batch_config.h")]
    objs = []
    procs = []
    for src in SOURCES:
        sp = os.path.join(CSRC, src)
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        stamp = obj + ".sha"
        dig = _digest([sp] + headers)
        objs.append(obj)
        if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
            continue
        cmd = [_nvcc()] + NVCC_FLAGS + ["-c", sp, "-o", obj]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        procs.append((subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT), src, stamp, dig))
    relink = not os.path.exists(OUT)
    for p, src, stamp, dig in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (src, out.decode(errors="replace")))
        with open(stamp, "w") as f:
            f.write(dig)
        relink = True
    if relink or force:
        cmd = [_nvcc(), "-shared", "-o", OUT] + objs + ["-lcudart", "-ldl"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s" % r.stdout.decode(errors="replace"))
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
