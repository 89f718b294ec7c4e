"""In-tree nvcc build of libwdmpnn.so (sm_100a only; nvcc cross-compiles without a GPU).

Each csrc/*.cu is compiled to an object under csrc/build/ (in parallel, reused while neither it nor any
header changed) and the objects are linked into polymer_chemprop_b200/libwdmpnn.so."""
import glob
import os
import subprocess
import time
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ_DIR = os.path.join(CSRC, "build")
LIB_PATH = os.environ.get("WDMPNN_LIB") or os.path.join(HERE, "libwdmpnn.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _headers():
    return glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(os.path.dirname(HERE), "include", "wdmpnn.h")]


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(p) > t for p in sources() + _headers())


def _run(cmd):
    t0 = time.time()
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    if "-c" in cmd:      # stamp the object with the START of its compile: a source edited meanwhile makes it stale
        os.utime(cmd[cmd.index("-o") + 1], (t0, t0))
    return proc.stderr


def build(force: bool = False, verbose: bool = False) -> str:
    """Compiles csrc/*.cu into polymer_chemprop_b200/libwdmpnn.so; returns its path."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "nvcc")
    extra = os.environ.get("WD_NVCC_EXTRA", "").split()
    flags = NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else [])
    os.makedirs(OBJ_DIR, exist_ok=True)
    tag = os.path.basename(LIB_PATH).replace(".so", "")           # variant builds keep their own objects
    newest_header = max(os.path.getmtime(p) for p in _headers())
    jobs, objs = [], []
    for src in sources():
        obj = os.path.join(OBJ_DIR, f"{tag}_{os.path.basename(src)[:-3]}.o")
        objs.append(obj)
        fresh = (not force and not extra and os.path.exists(obj) and
                 os.path.getmtime(obj) > max(os.path.getmtime(src), newest_header))
        if not fresh:
            jobs.append([nvcc] + flags + ["-c", "-o", obj, src])
    with ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 1))) as ex:
        logs = list(ex.map(_run, jobs))
    _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB_PATH] + objs)
    if verbose:
        print("\n".join(logs))
    return LIB_PATH


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
