"""Builds npge_b200/libnpge_anchor.so (hand-written sm_100a CUDA + the C ABI)
in-tree with nvcc.  nvcc cross-compiles without a GPU."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SOURCES = ["api.cu", "scan_kernels.cu", "msd_kernels.cu", "bucket_kernels.cu", "fragment_kernels.cu",
           "ingest_kernels.cu", "block_hash_kernels.cu", "multi.cu"]
HEADERS = ["common.cuh", "kernels.h", "rank.cuh", "dist_plan.h", "block_hash_core.h", "block_hash_names.h", "fetch_expand.h", os.path.join("..", "..", "include",
                                                     "npge_anchor.h")]
LIB = os.path.join(HERE, "libnpge_anchor.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3",
              "-std=c++17", "-Xcompiler", "-fPIC", "-shared"]


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t
               for f in SOURCES + HEADERS)


def build_variant(name, defines):
    """libnpge_anchor_<name>.so with extra -D flags: kernel experiments that are
    compared on the GPU box in one session (NPGE_ANCHOR_LIB selects it in _abi)."""
    out = os.path.join(HERE, "libnpge_anchor_%s.so" % name)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-D" + d for d in defines] + ["-o", out] + \
        [os.path.join(CSRC, f) for f in SOURCES]
    subprocess.check_call(cmd)
    return out


def build(force=False, verbose=False):
    if not force and not stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not os.path.exists(nvcc):
        nvcc = "nvcc"
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
        ["-o", LIB] + [os.path.join(CSRC, f) for f in SOURCES]
    subprocess.check_call(cmd)
    return LIB


HOST = os.path.join(HERE, "host")
SHIM_LIB = os.path.join(HOST, "libnpge_gpu_processors.so")
SHIM_TEST = os.path.join(HOST, "test_gpu_anchor_finder")
SHIM_TEST_MOVE = os.path.join(HOST, "test_gpu_move_unchanged")
# the same processors compiled in the reference's own spelling (NPGE_B200_REAL_HEADERS:
# string option rules, 3-argument add_gopt, add_opt_check) against host/ref_api
SHIM_TEST_REFAPI = os.path.join(HOST, "test_gpu_anchor_finder_refapi")
SHIM_SOURCES = ["gpu_anchor_finder.cpp", "gpu_anchor_finder.hpp", "npge_facade.hpp",
                os.path.join("ref_api", "Processor.hpp"),
                "test_gpu_anchor_finder.cpp", "test_gpu_move_unchanged.cpp"]


def build_shim(force=False):
    """The C++ host side: npge::gpu::AnchorFinder (libnpge_gpu_processors.so,
    linked against libnpge_anchor.so) and its unit-test binary."""
    build(force=False)
    srcs = [os.path.join(HOST, f) for f in SHIM_SOURCES] + [LIB]
    outs = [SHIM_LIB, SHIM_TEST, SHIM_TEST_MOVE, SHIM_TEST_REFAPI]
    if (not force and all(os.path.exists(f) for f in outs) and
            all(os.path.getmtime(f) <= min(os.path.getmtime(o) for o in outs) for f in srcs)):
        return SHIM_TEST
    cxx = os.environ.get("CXX", "g++")
    common = ["-std=c++17", "-O2", "-Wall", "-Wextra", "-I", HOST]
    link = ["-L", HERE, "-lnpge_anchor", "-Wl,-rpath,$ORIGIN/..", "-Wl,-rpath,$ORIGIN"]
    subprocess.check_call([cxx] + common + ["-fPIC", "-shared", "-o", SHIM_LIB,
                                            os.path.join(HOST, "gpu_anchor_finder.cpp")] + link)
    subprocess.check_call([cxx] + common + ["-o", SHIM_TEST,
                                            os.path.join(HOST, "test_gpu_anchor_finder.cpp"),
                                            "-L", HOST, "-lnpge_gpu_processors"] + link)
    subprocess.check_call([cxx] + common + ["-o", SHIM_TEST_MOVE,
                                            os.path.join(HOST, "test_gpu_move_unchanged.cpp"),
                                            "-L", HOST, "-lnpge_gpu_processors"] + link)
    subprocess.check_call([cxx] + common + ["-DNPGE_B200_REAL_HEADERS", "-I",
                                            os.path.join(HOST, "ref_api"), "-o", SHIM_TEST_REFAPI,
                                            os.path.join(HOST, "test_gpu_anchor_finder.cpp"),
                                            os.path.join(HOST, "gpu_anchor_finder.cpp")] + link)
    return SHIM_TEST


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_shim(force="--force" in sys.argv))
