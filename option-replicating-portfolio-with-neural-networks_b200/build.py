"""Build liborp_b200.so in-tree with nvcc for sm_100a (no torch headers, no JIT cache).

    python build.py            # rebuild if any source is newer than the library
    python build.py --force
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "liborp_b200.so")
SOURCES = ["orp_sobol_host.cu", "orp_paths.cu", "orp_payoff.cu", "orp_fit.cu", "orp_select.cu", "orp_survivors.cu", "orp_probe.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared", "-cudart", "static", "--threads", "0"]


def nvcc_path() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: liborp_b200.so cannot be built (there is no CPU fallback)")


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h", ".inc"))]
    deps.append(os.path.join(HERE, "..", "include", "orp_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, defines=(), out: str = LIB) -> str:
    """Compiles to a temporary name in the same directory under a file lock and renames it into place, so that several
    ranks started together (torchrun) never load a half-written library; whoever gets the lock second finds it built."""
    if not force and out == LIB and not needs_build():
        return LIB
    import fcntl
    with open(out + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        if not force and out == LIB and not needs_build():
            return LIB
        tmp = f"{out}.{os.getpid()}.tmp"
        cmd = [nvcc_path(), *NVCC_FLAGS, *[f"-D{d}" for d in defines], *(["-Xptxas", "-v"] if verbose else []),
               *[os.path.join(CSRC, s) for s in SOURCES], "-o", tmp]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            if os.path.exists(tmp):
                os.remove(tmp)
            raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
        os.replace(tmp, out)
        if verbose:
            print(r.stderr)
    return out


def build_trace() -> str:
    """The instrumented library of tools/trace_fit.py (phase timestamps inside the persistent fit kernel)."""
    return build(force=True, defines=("ORP_FIT_TRACE",), out=os.path.join(HERE, "liborp_b200_trace.so"))


if __name__ == "__main__":
    print(build_trace() if "--trace" in sys.argv else build(force="--force" in sys.argv, verbose="-v" in sys.argv))
