"""Build libbts_b200.so (hand-written sm_100a CUDA + the C ABI of include/bts_b200.h) in-tree.

    python -m bts_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU. The .so lands in bts_b200/lib/ (git-ignored, travels to the GPU
box with the tree). There is no fallback: if the library is missing the package refuses to run.
"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
SO = os.path.join(LIBDIR, "libbts_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, extra_flags=(), libdir=None):
    """extra_flags / libdir: an A/B variant (-D switches) built into its own directory; BTS_LIB selects it at run time."""
    global LIBDIR, SO
    if libdir is not None:
        LIBDIR = libdir
        SO = os.path.join(LIBDIR, "libbts_b200.so")
    os.makedirs(LIBDIR, exist_ok=True)
    headers = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(HERE, "..", "include", "*.h"))
    objs = []
    log = []
    todo = []
    for src in sources():
        obj = os.path.join(LIBDIR, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            todo.append((src, [NVCC] + FLAGS + list(extra_flags) + ["-c", src, "-o", obj]))
    # translation units are independent: compile them side by side (the big kernel files take minutes each)
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=max(1, min(len(todo), os.cpu_count() or 1))) as pool:
        results = list(pool.map(lambda t: subprocess.run(t[1], capture_output=True, text=True), todo))
    for (src, cmd), r in zip(todo, results):
        log.append("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            sys.stderr.write(log[-1])
            raise RuntimeError("nvcc failed on " + src)
    if force or _stale(SO, objs):
        cmd = [NVCC, "-shared", "-o", SO] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcuda"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log.append("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            sys.stderr.write(log[-1])
            raise RuntimeError("link failed")
    if log:
        with open(os.path.join(LIBDIR, "build.log"), "w") as f:
            f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    return SO


if __name__ == "__main__":
    # python -m bts_b200.build [--force] [--verbose] [--variant NAME -DX=1 ...]  (variant -> bts_b200/lib_NAME/)
    argv = sys.argv[1:]
    libdir, extra = None, []
    if "--variant" in argv:
        i = argv.index("--variant")
        libdir = os.path.join(HERE, "lib_" + argv[i + 1])
        extra = [f for f in argv[i + 2:] if f.startswith("-D")]
    print(build(force="--force" in argv, verbose="--verbose" in argv, extra_flags=extra, libdir=libdir))
