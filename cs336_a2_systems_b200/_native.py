"""Loader and build driver for the native library ``libb200sys.so``.

The hot ops are plain CUDA C++ (``csrc/*.cu``) compiled for sm_100a ONLY and exported
through a C ABI declared in ``include/b200sys.h``.  Python reaches them with ``ctypes``:
raw device pointers + the current CUDA stream, no torch C++ headers in the build (the
whole library compiles in well under a minute and the ``.so`` lives in-tree so that it
travels with a repository snapshot).

The reference has no native code at all; everything here is new.
"""

from __future__ import annotations

import ctypes
import os
import pathlib
import shutil
import subprocess
import sys
import threading

_PKG = pathlib.Path(__file__).resolve().parent
REPO = _PKG.parent
CSRC = REPO / "csrc"
INCLUDE = REPO / "include"
LIB_DIR = _PKG / "lib"
LIB_PATH = LIB_DIR / "libb200sys.so"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]

_lib = None
_lib_err: str | None = None
_lock = threading.Lock()


def nvcc_path() -> str | None:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    return None


def sources() -> list[pathlib.Path]:
    return sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cc"))


def _stale() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    deps = sources() + sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.h")) + sorted(INCLUDE.glob("*.h"))
    return any(p.stat().st_mtime > t for p in deps)


def build(force: bool = False, verbose: bool = False, jobs: int | None = None) -> pathlib.Path:
    """Compile every ``csrc`` translation unit for sm_100a and link ``libb200sys.so``.

    Objects are compiled in parallel (one nvcc per file) and cached by mtime under
    ``build/``; ``-lineinfo`` keeps ncu's source page usable.
    """
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found; cannot build libb200sys.so")
    if not force and not _stale():
        return LIB_PATH
    obj_dir = REPO / "build" / "obj"
    obj_dir.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    hdrs = sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.h")) + sorted(INCLUDE.glob("*.h"))
    hdr_time = max((h.stat().st_mtime for h in hdrs), default=0.0)
    common = [
        "-O3", "-std=c++17", "-lineinfo", "--use_fast_math", "-Xcompiler", "-fPIC",
        "-I", str(INCLUDE), "-I", str(CSRC), *ARCH_FLAGS,
        "-Xptxas", "-v" if verbose else "-warn-spills",
    ]
    jobs = jobs or min(8, os.cpu_count() or 4)
    todo = []
    objs = []
    for src in sources():
        obj = obj_dir / (src.stem + ".o")
        objs.append(obj)
        if force or not obj.exists() or obj.stat().st_mtime < max(src.stat().st_mtime, hdr_time):
            todo.append((src, obj))
    procs: list[tuple[subprocess.Popen, pathlib.Path]] = []
    failed = []

    def _drain(block_until: int):
        while len(procs) > block_until:
            p, s = procs.pop(0)
            out, _ = p.communicate()
            if verbose and out:
                sys.stderr.write(out.decode(errors="replace"))
            if p.returncode != 0:
                failed.append((s, out.decode(errors="replace")))

    for src, obj in todo:
        cmd = [nvcc, *common, "-c", str(src), "-o", str(obj)]
        procs.append((subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT), src))
        _drain(jobs - 1)
    _drain(0)
    if failed:
        msg = "\n".join(f"--- {s.name} ---\n{o}" for s, o in failed)
        raise RuntimeError(f"nvcc failed:\n{msg}")
    link = [nvcc, "-shared", *ARCH_FLAGS, "-o", str(LIB_PATH), *map(str, objs), "-lcudart"]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout.decode(errors="replace"))
    global _lib, _lib_err
    _lib, _lib_err = None, None
    return LIB_PATH


def _load():
    global _lib, _lib_err
    with _lock:
        if _lib is not None or _lib_err is not None:
            return
        if not LIB_PATH.exists():
            _lib_err = f"{LIB_PATH} not built (run `python -m cs336_a2_systems_b200.build`)"
            return
        try:
            import torch  # noqa: F401  (makes libcudart / libcuda resident first)
            lib = ctypes.CDLL(str(LIB_PATH), mode=ctypes.RTLD_GLOBAL)
            lib.b200_last_error.restype = ctypes.c_char_p
            _lib = lib
        except OSError as e:  # pragma: no cover
            _lib_err = f"cannot load {LIB_PATH}: {e}"


def lib():
    """The loaded library; raises loudly when it is missing (never a silent fallback)."""
    _load()
    if _lib is None:
        raise RuntimeError(f"libb200sys.so unavailable: {_lib_err}")
    return _lib


def have_lib() -> bool:
    _load()
    return _lib is not None


def force_torch() -> bool:
    """``B200SYS_FORCE_TORCH=1`` routes CUDA tensors through the PyTorch oracles (debug/A-B only)."""
    return os.environ.get("B200SYS_FORCE_TORCH", "0") == "1"


def use_native(*tensors) -> bool:
    """True when the op must run on the sm_100a kernels: CUDA tensors and not forced off.

    A CUDA tensor with the library missing is an error, not a fallback.
    """
    if not tensors or not all(t.is_cuda for t in tensors if t is not None):
        return False
    if force_torch():
        return False
    lib()
    return True


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().b200_last_error()
        raise RuntimeError(f"b200sys {what} failed (rc={rc}): {msg.decode() if msg else '?'}")


def stream_ptr() -> ctypes.c_void_p:
    import torch

    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def p(t) -> ctypes.c_void_p:
    """Device pointer of a tensor (None → NULL)."""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


if __name__ == "__main__":  # pragma: no cover
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
