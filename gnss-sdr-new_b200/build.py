"""Builds libgnss_b200.so (hand-written sm_100a CUDA + the C-ABI) in-tree with nvcc."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libgnss_b200.so")
SOURCES = ["api_misc.cu", "multicorr.cu", "trk_loop.cu", "resampler.cu", "acq.cu", "acq_group.cu", "acq_fast_lower.cu", "acq_fast_upper.cu", "acq_fast_fwd.cu", "acq_cluster.cu", "fft_plan.cc", "wipeoff_pieces.cc"]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "gnss_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: libgnss_b200.so must be built where the CUDA toolkit is installed")
    flags = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
             "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden"]
    if verbose:
        flags += ["-Xptxas", "-v"]
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(objdir, src + ".o")
        subprocess.run([nvcc] + flags + ["-c", "-o", obj, os.path.join(CSRC, src)], check=True, cwd=CSRC)
        return obj

    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 4)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs, check=True)
    return LIB


HOST_DIR = os.path.join(HERE, "host")
HOST_LIB = os.path.join(HERE, "libgnss_b200_host.so")
HOST_SOURCES = ["b200_correlator_pool.cc", "b200_multicorrelator_real_codes.cc", "b200_multicorrelator.cc",
                "b200_acquisition_bank.cc", "b200_mat_writer.cc", "b200_tracking_dump.cc", "b200_symbol_sync.cc", "pcps_acquisition_b200.cc"]
HOST_TEST_SRC = os.path.join(HERE, "..", "tests", "host", "host_mirror_test.cc")
HOST_TEST_BIN = os.path.join(HERE, "..", "tests", "host", "host_mirror_test")


def build_host(force=False):
    """The host C++ mirror of the reference's block / correlator classes (standalone flavour: stand-in headers for
    Acq_Conf / Gnss_Synchro / ChannelFsm), and the test driver that exercises it. Plain g++; links the C-ABI library."""
    build(force=False)
    srcs = [os.path.join(HOST_DIR, f) for f in HOST_SOURCES]
    deps = srcs + [os.path.join(HOST_DIR, f) for f in os.listdir(HOST_DIR) if f.endswith(".h")] + [LIB, HOST_TEST_SRC]
    deps += [os.path.join(HOST_DIR, "stubs", f) for f in os.listdir(os.path.join(HOST_DIR, "stubs")) if f.endswith(".h")]
    if not force and os.path.exists(HOST_LIB) and os.path.exists(HOST_TEST_BIN) and \
            all(os.path.getmtime(d) <= min(os.path.getmtime(HOST_LIB), os.path.getmtime(HOST_TEST_BIN)) for d in deps):
        return HOST_LIB, HOST_TEST_BIN
    inc = ["-I", os.path.join(HERE, "..", "include"), "-I", HOST_DIR]
    common = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-fPIC", "-pthread", "-DB200_HOST_STANDALONE"] + inc
    subprocess.run(common + ["-shared", "-o", HOST_LIB] + srcs + ["-L", HERE, "-lgnss_b200", "-Wl,-rpath,$ORIGIN"], check=True)
    subprocess.run(common + ["-o", HOST_TEST_BIN, HOST_TEST_SRC, "-L", HERE, "-lgnss_b200_host", "-lgnss_b200",
                             "-Wl,-rpath,$ORIGIN/../../gnss-sdr-new_b200"], check=True)
    return HOST_LIB, HOST_TEST_BIN


INTEGRATION_DIR = os.path.join(HERE, "..", "integration")
ADAPTER_TEST_SRC = os.path.join(HERE, "..", "tests", "host", "adapter_test.cc")
ADAPTER_TEST_BIN = os.path.join(HERE, "..", "tests", "host", "adapter_test")


def build_integration(force=False):
    """The sources a maintainer adds to the GNSS-SDR tree (integration/src: gr::block shell, the three adapters, the factory
    arms), compiled with -Wall -Wextra -Werror against the stand-in headers of integration/stubs/, and the harness that drives
    an adapter through Channel's call order (tests/host/adapter_test.cc)."""
    build_host(force=False)
    srcs = []
    for root, _, files in os.walk(os.path.join(INTEGRATION_DIR, "src")):
        srcs += [os.path.join(root, f) for f in sorted(files) if f.endswith(".cc")]
    srcs.append(os.path.join(INTEGRATION_DIR, "stubs", "acq_conf_stub.cc"))
    deps = list(srcs) + [ADAPTER_TEST_SRC, HOST_LIB]
    for root, _, files in os.walk(INTEGRATION_DIR):
        deps += [os.path.join(root, f) for f in files if f.endswith(".h")]
    if not force and os.path.exists(ADAPTER_TEST_BIN) and all(os.path.getmtime(d) <= os.path.getmtime(ADAPTER_TEST_BIN) for d in deps):
        return ADAPTER_TEST_BIN
    inc = ["-I", os.path.join(HERE, "..", "include"), "-I", HOST_DIR, "-I", os.path.join(HOST_DIR, "stubs"), "-I", os.path.join(INTEGRATION_DIR, "stubs")]
    for sub in ("algorithms/acquisition/gnuradio_blocks", "algorithms/acquisition/adapters", "core/receiver"):
        inc += ["-I", os.path.join(INTEGRATION_DIR, "src", sub)]
    cmd = ["g++", "-std=c++20", "-O2", "-Wall", "-Wextra", "-Werror", "-pthread", "-DB200_HOST_STANDALONE"] + inc + ["-o", ADAPTER_TEST_BIN, ADAPTER_TEST_SRC] + srcs + \
          ["-L", HERE, "-lgnss_b200_host", "-lgnss_b200", "-Wl,-rpath,$ORIGIN/../../gnss-sdr-new_b200"]
    subprocess.run(cmd, check=True)
    return ADAPTER_TEST_BIN


if __name__ == "__main__":
    import sys
    print(build(force=True, verbose="-v" in sys.argv))
    print(build_host(force=True))
    print(build_integration(force=True))
