"""Builds libddpb200.so (all sm_100a kernels + their extern "C" launchers) in-tree with nvcc.

The reference has no native code and therefore no build step; this is
the only build the framework needs.  Objects are cached on source mtime, so a rebuild
after touching one .cu recompiles just that file.

    python -m ddpb200.build            # build if stale
    python -m ddpb200.build --force    # rebuild everything
    python -m ddpb200.build --ptxas-v  # print registers / spills / smem per kernel
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
CSRC = ROOT / "csrc"
LIB_DIR = ROOT / "_lib"
OBJ_DIR = LIB_DIR / "obj"
LIB_PATH = LIB_DIR / "libddpb200.so"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default",
              "--expt-relaxed-constexpr", "-I", str(ROOT.parent / "include"), "-I", str(CSRC)]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _headers_mtime() -> float:
    hs = (list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.inl")) + list(CSRC.glob("*.h")) +
          list((ROOT.parent / "include").glob("*.h")))
    return max((h.stat().st_mtime for h in hs), default=0.0)


def _compile_one(nvcc: str, src: Path, obj: Path, extra: list[str]) -> str:
    cmd = [nvcc, *ARCH_FLAGS, *NVCC_FLAGS, *extra, "-c", str(src), "-o", str(obj)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
    return res.stderr


def build(force: bool = False, verbose: bool = False, ptxas_v: bool = False, defines: list[str] | None = None,
          out: Path | None = None) -> Path:
    """``defines``/``out`` build an experiment variant (e.g. a different pipeline depth) into its
    own .so; point DDPB_LIB at it to use it."""
    global OBJ_DIR, LIB_PATH
    nvcc = find_nvcc()
    if out is not None:
        LIB_PATH = Path(out).resolve()
        OBJ_DIR = LIB_PATH.parent / ("obj_" + LIB_PATH.stem)
        force = True
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    hdr_m = _headers_mtime()
    extra = (["-Xptxas", "-v"] if ptxas_v else []) + [f"-D{d}" for d in (defines or [])]
    todo, objs = [], []
    for src in sources():
        obj = OBJ_DIR / (src.stem + ".o")
        objs.append(obj)
        stale = force or ptxas_v or not obj.exists() or obj.stat().st_mtime < max(src.stat().st_mtime, hdr_m)
        if stale:
            todo.append((src, obj))
    if todo:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            futs = {ex.submit(_compile_one, nvcc, s, o, extra): s for s, o in todo}
            for fut in cf.as_completed(futs):
                log = fut.result()
                if verbose or ptxas_v:
                    print(f"== {futs[fut].name}\n{log}", file=sys.stderr)
    if todo or not LIB_PATH.exists():
        cmd = [nvcc, *ARCH_FLAGS, "-shared", "-o", str(LIB_PATH), *map(str, objs), "-cudart", "static"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return LIB_PATH


def is_stale() -> bool:
    if not LIB_PATH.exists():
        return True
    m = LIB_PATH.stat().st_mtime
    return any(s.stat().st_mtime > m for s in sources()) or _headers_mtime() > m


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--ptxas-v", action="store_true")
    ap.add_argument("-D", dest="defines", action="append", default=[], help="extra preprocessor define")
    ap.add_argument("--out", default=None, help="write the library here instead of the default path")
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose, ptxas_v=a.ptxas_v, defines=a.defines, out=a.out))


if __name__ == "__main__":
    main()
