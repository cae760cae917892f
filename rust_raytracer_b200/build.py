"""In-tree build of the CUDA library (sm_100a) and, separately, of the CPU oracle.

`build_library()` compiles rust_raytracer_b200/csrc into rust_raytracer_b200/libb200rt.so with nvcc
(cross-compiles without a GPU).  `build_oracle()` compiles oracle/oracle.cpp into oracle/liboracle.so; the
oracle is test infrastructure and is never loaded by the package itself.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libb200rt.so")
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "liboracle.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC,-O3,-Wall", "-Xptxas", "-v", "--use_fast_math=false",
]
SOURCES = ["api.cu", "builder.cpp", "loaders.cpp", "scenes.cpp", "scene_compile.cpp", "png_out.cpp"]
COUNTERS_LIB = os.path.join(HERE, "libb200rt_counters.so")  # the same sources with -DB200RT_COUNTERS: per-operation counts (bench.py's algorithmic flops, tools/dbg)
CLI = os.path.join(HERE, "b200rt_render")  # the reference's `main` over the C ABI (csrc/cli_main.cpp)


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _deps(directory, extra=()):
    out = [os.path.join(directory, f) for f in os.listdir(directory) if f.endswith((".cu", ".cuh", ".cpp", ".hpp", ".h"))]
    return out + list(extra)


def build_library(force=False, verbose=False):
    header = os.path.join(ROOT, "include", "b200rt.h")
    if not force and not _newer(LIB, _deps(CSRC, [header])):
        if not os.path.exists(CLI):
            build_cli()
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    flags = [f for f in NVCC_FLAGS if f != "--use_fast_math=false"]
    cmd = [nvcc] + flags + ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    log = os.path.join(HERE, "build.log")
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + res.stdout)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed (see %s)" % log)
    build_cli()
    return LIB


def build_counters(force=False):
    """The instrumented twin of the library (-DB200RT_COUNTERS): counts node visits, primitive tests, scatter calls ... so that
    bench.py can report SURVEY.md section 8(d)'s algorithmic flops of the workload it times.  Never on the timed path."""
    header = os.path.join(ROOT, "include", "b200rt.h")
    if not force and not _newer(COUNTERS_LIB, _deps(CSRC, [header])):
        return COUNTERS_LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    flags = [f for f in NVCC_FLAGS if f not in ("--use_fast_math=false", "-v", "-Xptxas")] + ["-DB200RT_COUNTERS"]
    cmd = [nvcc] + flags + ["-o", COUNTERS_LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("nvcc failed for the counters build")
    return COUNTERS_LIB


def build_cli():
    """g++ build of the command-line driver against the in-tree library (rpath $ORIGIN)."""
    src = os.path.join(CSRC, "cli_main.cpp")
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-Wall", "-pthread", "-o", CLI, src, "-L" + HERE, "-lb200rt", "-Wl,-rpath,$ORIGIN"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("g++ failed for b200rt_render")
    return CLI


def build_variant(name, defines=(), out_dir=None):
    """A/B or debug build of the same sources with extra -D flags (tools/dbg, profiling): loaded via B200RT_LIB."""
    out_dir = out_dir or os.path.join(ROOT, "tools", "dbg")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "libb200rt_%s.so" % name)
    nvcc = os.environ.get("NVCC", "nvcc")
    flags = [f for f in NVCC_FLAGS if f != "--use_fast_math=false"] + ["-D" + d for d in defines]
    cmd = [nvcc] + flags + ["-o", out] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    with open(out + ".log", "w") as f:
        f.write(" ".join(cmd) + "\n" + res.stdout)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("nvcc failed for variant %s" % name)
    return out


def build_oracle(force=False):
    src = os.path.join(ORACLE_DIR, "oracle.cpp")
    header = os.path.join(ROOT, "include", "b200rt.h")
    if not force and not _newer(ORACLE_LIB, [src, header]):
        return ORACLE_LIB
    res = subprocess.run(["make", "-C", ORACLE_DIR, "-B"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("oracle build failed")
    return ORACLE_LIB


if __name__ == "__main__":
    build_library(force="--force" in sys.argv, verbose=True)
    build_counters(force="--force" in sys.argv)
    build_oracle(force="--force" in sys.argv)
    print(LIB)
    print(ORACLE_LIB)
