"""In-tree native build (the counterpart of the reference's hw/Makefile + sw/Makefile).

Two shared libraries land in ``fpga_ai_nic_b200/_lib`` so that they travel with the source tree:

* ``libainic_oracle.so`` - CPU oracle / fake NIC backend / mini-MPI (g++, OpenMP); needs no GPU.
* ``libainic_cuda.so``   - every CUDA kernel + the communicator + the NIC command interface,
  compiled for ``sm_100a`` only (``-gencode arch=compute_100a,code=sm_100a -lineinfo``).

``python -m fpga_ai_nic_b200.utils.build`` rebuilds what is stale; ``--force`` rebuilds all;
``--sass`` also dumps SASS listings into ``profiles/sass``.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
LIBDIR = ROOT / "fpga_ai_nic_b200" / "_lib"
OBJDIR = ROOT / "build" / "obj"

NVCC = os.environ.get("NVCC", shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc")


def _pick_cxx() -> tuple[str, bool]:
    """Host compiler + whether it can link OpenMP (the image's $CXX wrapper lacks libgomp.spec)."""
    cands = [os.environ.get("AINIC_CXX"), "/usr/bin/g++", shutil.which("g++"), os.environ.get("CXX")]
    probe = "int main(){return 0;}"
    for c in [c for c in cands if c]:
        for omp in (True, False):
            try:
                p = subprocess.run([c, "-x", "c++", "-", "-o", "/dev/null"] + (["-fopenmp"] if omp else []),
                                   input=probe, capture_output=True, text=True, timeout=60)
            except Exception:
                continue
            if p.returncode == 0:
                return c, omp
    return "g++", False


CXX, HAVE_OMP = _pick_cxx()

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr", "--extended-lambda", "-ccbin", CXX, "-Xcompiler", "-fPIC,-O3", "-Xptxas", "-v",
    "-I", str(ROOT / "include"), "-I", str(CSRC),
]
CXX_FLAGS = ["-O3", "-march=x86-64-v3", "-std=c++17", "-fPIC", *(["-fopenmp"] if HAVE_OMP else []), "-Wall", "-Wno-unknown-pragmas", "-I", str(ROOT / "include"), "-I", str(CSRC)]

CUDA_SOURCES = [
    "codec/codec_kernels.cu",
    "optim/sgd.cu",
    "nn/softmax.cu",
    "gemm/gemm_tcgen05.cu",
    "gemm/gemm_tcgen05_2cta.cu",
    "gemm/gemm_kinds.cu",
    "comm/comm.cu",
    "comm/nvls.cu",
    "comm/emulate_noalloc.cu",
    "api/nic.cpp",
    "api/mlp.cu",
    "oracle/oracle.cpp",   # the fake NIC backend inside nic.cpp computes with the CPU oracle
]
ORACLE_SOURCES = [
    "oracle/oracle.cpp",
    "oracle/mini_mpi.cpp",
    "api/nic.cpp",
]
APPS = {"mlp_mpi_example_f32": ["apps/mlp_mpi_example_f32.cpp"]}


def _deps_mtime() -> float:
    hs = list(CSRC.rglob("*.h")) + list(CSRC.rglob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    return max((p.stat().st_mtime for p in hs), default=0.0)


def _stale(out: Path, srcs: list[Path], hdr_mtime: float) -> bool:
    if not out.exists():
        return True
    t = out.stat().st_mtime
    return any(s.stat().st_mtime > t for s in srcs) or hdr_mtime > t


def _run(cmd: list[str], log: Path | None = None) -> None:
    p = subprocess.run(cmd, capture_output=True, text=True)
    if log is not None:
        log.parent.mkdir(parents=True, exist_ok=True)
        log.write_text(p.stdout + p.stderr)
    if p.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + p.stdout + p.stderr)
        raise RuntimeError(f"build step failed: {cmd[0]} ... {cmd[-1]}")


def _content_hash(srcs: list[Path]) -> str:
    """Hash of everything a library is built from (its sources, every header, the flags).  Staleness is
    decided by content, not by mtime: a copied tree (the GPU boxes get a snapshot) has arbitrary mtimes."""
    import hashlib
    h = hashlib.sha1()
    hs = sorted(list(CSRC.rglob("*.h")) + list(CSRC.rglob("*.cuh")) + list((ROOT / "include").glob("*.h")))
    for f in sorted(srcs) + hs:
        h.update(str(f.relative_to(ROOT)).encode())
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS[:-6] + CXX_FLAGS[:-4]).encode())
    return h.hexdigest()


def _lib_current(out: Path, srcs: list[Path]) -> bool:
    side = out.with_suffix(out.suffix + ".srchash")
    try:
        return out.exists() and out.stat().st_size > 0 and side.read_text().strip() == _content_hash(srcs)
    except OSError:
        return False


class _BuildLock:
    """One builder at a time per tree (N ranks importing the package at once must not all compile and
    overwrite the same .so): an advisory lock on a file in the library directory."""

    def __enter__(self):
        import fcntl
        LIBDIR.mkdir(parents=True, exist_ok=True)
        self.fh = open(LIBDIR / ".build.lock", "w")
        fcntl.flock(self.fh, fcntl.LOCK_EX)
        return self

    def __exit__(self, *a):
        import fcntl
        fcntl.flock(self.fh, fcntl.LOCK_UN)
        self.fh.close()


def _existing(rel: list[str], base: Path) -> list[Path]:
    return [base / r for r in rel if (base / r).exists()]


def build_oracle(force: bool = False) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    out = LIBDIR / "libainic_oracle.so"
    srcs = _existing(ORACLE_SOURCES, CSRC)
    if not force and _lib_current(out, srcs):
        return out
    with _BuildLock():
        if force or not _lib_current(out, srcs):      # somebody else may have built it while we waited
            tmp = out.with_suffix(f".tmp{os.getpid()}.so")
            _run([CXX, *CXX_FLAGS, "-shared", "-o", str(tmp), *map(str, srcs), "-lpthread"])
            os.replace(tmp, out)
            out.with_suffix(out.suffix + ".srchash").write_text(_content_hash(srcs))
    return out


def build_cuda(force: bool = False, jobs: int | None = None) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    OBJDIR.mkdir(parents=True, exist_ok=True)
    out = LIBDIR / "libainic_cuda.so"
    srcs = _existing(CUDA_SOURCES, CSRC)
    if not force and _lib_current(out, srcs):
        return out
    with _BuildLock():
        if not force and _lib_current(out, srcs):
            return out
        return _build_cuda_locked(out, srcs, force, jobs)


def _build_cuda_locked(out: Path, srcs: list[Path], force: bool, jobs: int | None) -> Path:
    hdr = _deps_mtime()
    objs, todo = [], []
    for s in srcs:
        o = OBJDIR / (s.relative_to(CSRC).as_posix().replace("/", "_") + ".o")
        objs.append(o)
        if force or _stale(o, [s], hdr):
            todo.append((s, o))

    def compile_one(so):
        s, o = so
        extra = ["-x", "cu", "-DAINIC_WITH_CUDA"] if s.suffix == ".cpp" else []
        _run([NVCC, *NVCC_FLAGS, *extra, "-c", str(s), "-o", str(o)], log=OBJDIR / (o.name + ".ptxas.log"))

    if todo:
        with cf.ThreadPoolExecutor(max_workers=jobs or min(8, os.cpu_count() or 4)) as ex:
            list(ex.map(compile_one, todo))
    tmp = out.with_suffix(f".tmp{os.getpid()}.so")
    _run([NVCC, "-shared", "-o", str(tmp), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a",
          "--cudart=static", "-ccbin", CXX, "-lpthread", "-ldl", "-lrt"])
    os.replace(tmp, out)          # atomic: a concurrent dlopen sees the old or the new library, never half of one
    out.with_suffix(out.suffix + ".srchash").write_text(_content_hash(srcs))
    return out


def build_apps(force: bool = False) -> list[Path]:
    bindir = ROOT / "build" / "bin"
    bindir.mkdir(parents=True, exist_ok=True)
    outs = []
    oracle = build_oracle(force)
    for name, rel in APPS.items():
        srcs = _existing(rel, ROOT)
        if not srcs:
            continue
        out = bindir / name
        if force or _stale(out, srcs + [oracle], _deps_mtime()):
            _run([CXX, *CXX_FLAGS, "-o", str(out), *map(str, srcs), f"-L{LIBDIR}", "-lainic_oracle",
                  f"-Wl,-rpath,{LIBDIR}", "-lpthread"])
        outs.append(out)
    return outs


def dump_sass() -> Path:
    out = ROOT / "profiles" / "sass"
    out.mkdir(parents=True, exist_ok=True)
    lib = LIBDIR / "libainic_cuda.so"
    p = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True)
    (out / "libainic_cuda.sass").write_text(p.stdout)
    return out


def build_all(force: bool = False) -> dict:
    return {"oracle": build_oracle(force), "cuda": build_cuda(force), "apps": build_apps(force)}


if __name__ == "__main__":
    force = "--force" in sys.argv
    res = build_all(force)
    for k, v in res.items():
        print(k, v)
    if "--sass" in sys.argv:
        print("sass", dump_sass())
