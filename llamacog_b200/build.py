"""Build recipes for the native libraries (nvcc, sm_100a only; objects compile in parallel).

  libb200ops.so   kernels + plain C ABI (include/b200_ops.h); no ggml dependency
  libggml-b200.so the ggml backend plugin (include/ggml-b200.h); needs the reference headers under
                  /root/reference (build container only) and links the host's libggml-base.so

Built files stay in-tree (git-ignored) so they travel to the GPU box with the snapshot.
"""
import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OBJ = os.path.join(PKG, "build")
OPS_LIB = os.path.join(PKG, "libb200ops.so")
PLUGIN_LIB = os.path.join(PKG, "libggml-b200.so")
REF = os.environ.get("LLAMACOG_REF", "/root/reference")
HOST = os.path.join(ROOT, "baseline", "_ref")

NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-I", os.path.join(ROOT, "include")]
if os.environ.get("B200_KTRACE", "0") not in ("", "0"):      # debug build with the kernel timeline hooks (common.cuh); rebuild everything
    NVCC_FLAGS.append("-DB200_KTRACE")
if os.environ.get("B200_MMQ_TRACE", "0") not in ("", "0"):   # debug build: clock64 stamps inside mmq_kernel (tools/mmq_debug.py trace)
    NVCC_FLAGS.append("-DB200_MMQ_TRACE")


def _newer(src_files, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in src_files)


def _flags_changed():
    """Objects are reused by mtime; a different flag set (B200_KTRACE / B200_MMQ_TRACE debug builds) must rebuild all of them."""
    stamp = os.path.join(OBJ, "nvcc_flags.txt")
    want = " ".join(NVCC_FLAGS)
    have = open(stamp).read() if os.path.exists(stamp) else None
    if have == want:
        return False
    os.makedirs(OBJ, exist_ok=True)
    for o in glob.glob(os.path.join(OBJ, "*.o")):
        os.remove(o)
    open(stamp, "w").write(want)
    return True


def _compile(src, extra=()):
    obj = os.path.join(OBJ, os.path.basename(src) + ".o")
    deps = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "backend", "*.h")) + glob.glob(os.path.join(CSRC, "backend", "*.inc")) + \
        glob.glob(os.path.join(ROOT, "include", "*.h")) + [src]
    if _newer(deps, obj):
        cmd = [NVCC] + NVCC_FLAGS + list(extra) + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
        if r.stderr.strip():
            sys.stderr.write(r.stderr)
    return obj


def build_ops(verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    _flags_changed()
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    extra = ["-Xptxas", "-v"] if verbose else []
    with cf.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        objs = list(ex.map(lambda s: _compile(s, extra), srcs))
    if _newer(objs, OPS_LIB):
        subprocess.check_call([NVCC, "-shared", "-o", OPS_LIB] + objs + ["-lcudart"])
    return OPS_LIB


def build_plugin():
    """ggml backend plugin; only possible where the reference headers exist."""
    bdir = os.path.join(CSRC, "backend")
    srcs = sorted(glob.glob(os.path.join(bdir, "*.cpp")) + glob.glob(os.path.join(bdir, "*.cu")))
    if not srcs:
        return None
    if not os.path.isdir(os.path.join(REF, "ggml", "include")):
        if os.path.exists(PLUGIN_LIB):
            return PLUGIN_LIB          # GPU box: use the prebuilt plugin that travelled with the snapshot
        raise RuntimeError("reference headers not found and no prebuilt libggml-b200.so")
    if not os.path.exists(os.path.join(HOST, "lib", "libggml-base.so")):
        subprocess.check_call(["make", "-s", "-j", str(os.cpu_count() or 4), "-C", os.path.join(ROOT, "baseline"), "all"])
    inc = ["-I", os.path.join(REF, "ggml", "include"), "-I", os.path.join(REF, "ggml", "src"),
           "-DGGML_BACKEND_SHARED", "-DGGML_BACKEND_BUILD", "-DGGML_SHARED"]
    os.makedirs(OBJ, exist_ok=True)
    with cf.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        objs = list(ex.map(lambda s: _compile(s, ["-x", "cu"] + inc), srcs))
    if _newer(objs + [OPS_LIB], PLUGIN_LIB):
        subprocess.check_call([NVCC, "-shared", "-o", PLUGIN_LIB] + objs +
                              ["-L", PKG, "-lb200ops", "-L", os.path.join(HOST, "lib"), "-lggml-base", "-lcudart",
                               "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN",
                               "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../baseline/_ref/lib"])
    return PLUGIN_LIB


HOST_TEST = os.path.join(ROOT, "tests", "host", "build", "logits_check")


def build_host_tests():
    """tests/host/logits_check: libllama harness for the end-to-end logits parity test (needs the reference headers)."""
    src = os.path.join(ROOT, "tests", "host", "logits_check.cpp")
    if not os.path.isdir(os.path.join(REF, "include")):
        return HOST_TEST if os.path.exists(HOST_TEST) else None
    if _newer([src], HOST_TEST):
        os.makedirs(os.path.dirname(HOST_TEST), exist_ok=True)
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-I", os.path.join(REF, "include"), "-I", os.path.join(REF, "ggml", "include"),
                               src, "-o", HOST_TEST, "-L", os.path.join(HOST, "lib"), "-lllama", "-lggml", "-lggml-base",
                               "-Wl,-rpath,$ORIGIN/../../../baseline/_ref/lib"])
    return HOST_TEST


def build_tools():
    """Stand-alone measurement programs under tools/ (tcgen05 issue-rate peaks for bench.py's frac_int8_peak, NVLink flag round trip)."""
    for name in ("tc_peak", "p2p_pingpong"):
        src, exe = os.path.join(ROOT, "tools", name + ".cu"), os.path.join(ROOT, "tools", name + ".bin")
        if os.path.exists(src) and _newer([src] + glob.glob(os.path.join(CSRC, "*.cuh")), exe):
            subprocess.check_call([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-o", exe, src])


def build_all(verbose=False):
    build_ops(verbose)
    build_plugin()
    build_host_tests()
    build_tools()


if __name__ == "__main__":
    build_all(verbose="-v" in sys.argv)
    print("built", OPS_LIB, PLUGIN_LIB if os.path.exists(PLUGIN_LIB) else "")
