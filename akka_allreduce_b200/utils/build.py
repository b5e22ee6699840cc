"""In-tree native build: host runtime (`_core`, g++/pybind11), CUDA data plane (`libaar_cuda.so`, nvcc
sm_100a) and the two CLI binaries.

The reference builds with sbt (build.sbt:1-34); here everything is a handful of direct compiler
invocations so the built artefacts live next to the package (they travel to the GPU box with the
snapshot) and no JIT cache is involved.

    python -m akka_allreduce_b200.utils.build [--force] [--only core|cuda|cli]
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
PKG = ROOT / "akka_allreduce_b200"
CSRC = ROOT / "csrc"
BIN = ROOT / "bin"

CORE_SOURCES = [
    "core/worker.cpp",
    "core/master.cpp",
    "core/config.cpp",
    "core/checkpoint.cpp",
    "core/local_cluster.cpp",
    "core/tcp_transport.cpp",
    "core/device_loop.cpp",
]
CUDA_SOURCES = ["cuda/aar_kernels.cu", "cuda/aar_bulk.cu", "cuda/aar_symm.cu", "cuda/aar_bench.cu", "cuda/aar_engine.cu"]
CUDA_ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]


def core_path() -> Path:
    return PKG / ("_core" + sysconfig.get_config_var("EXT_SUFFIX"))


def cuda_path() -> Path:
    return PKG / "libaar_cuda.so"


def _digest(deps: list[Path]) -> str:
    import hashlib

    h = hashlib.sha1()
    for d in sorted(set(deps)):
        if d.exists():
            h.update(str(d.relative_to(ROOT)).encode())
            h.update(d.read_bytes())
    return h.hexdigest()


def _stale(out: Path, deps: list[Path]) -> bool:
    """Stale = the output is missing or was built from different source CONTENT.  Modification times are useless here:
    the tree is snapshotted onto fresh GPU boxes, where every file looks new and a time-based check would rebuild
    everything on paid GPU time."""
    if not out.exists():
        return True
    stamp = out.with_name(out.name + ".srchash")
    return (not stamp.exists()) or stamp.read_text().strip() != _digest(deps)


def _stamp(out: Path, deps: list[Path]) -> None:
    out.with_name(out.name + ".srchash").write_text(_digest(deps))


def _run(cmd: list[str]) -> None:
    print("+ " + " ".join(str(c) for c in cmd), flush=True)
    subprocess.run([str(c) for c in cmd], check=True, cwd=ROOT)


def _headers() -> list[Path]:
    return list(CSRC.rglob("*.h")) + list(CSRC.rglob("*.cuh")) + list((ROOT / "include").glob("*.h"))


def _compile_objects(srcs: list[Path], flags: list[str], tag: str, force: bool) -> list[Path]:
    """Compile each translation unit to build/<tag>/<name>.o in parallel; only stale ones are rebuilt."""
    from concurrent.futures import ThreadPoolExecutor

    objdir = ROOT / "build" / tag
    objdir.mkdir(parents=True, exist_ok=True)
    hdrs = _headers()
    jobs = []
    objs = []
    for src in srcs:
        obj = objdir / (src.stem + ".o")
        objs.append(obj)
        if force or not obj.exists():
            jobs.append(["g++", *flags, "-c", src, "-o", obj])
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(_run, jobs))
    return objs


def build_core(force: bool = False) -> Path:
    import pybind11

    out = core_path()
    srcs = [CSRC / s for s in CORE_SOURCES] + [CSRC / "bindings/core_pybind.cpp"]
    if force or _stale(out, srcs + _headers()):
        flags = ["-O2", "-g1", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-Wall", "-Wno-unused-function",
                 f"-I{pybind11.get_include()}", f"-I{sysconfig.get_paths()['include']}", f"-I{ROOT / 'include'}"]
        objs = _compile_objects(srcs, flags, "core", True)
        _run(["g++", "-shared", *objs, "-o", out, "-lpthread"])
        _stamp(out, srcs + _headers())
    return out


def nvcc() -> str | None:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    return None


def build_cuda(force: bool = False) -> Path | None:
    out = cuda_path()
    srcs = [CSRC / s for s in CUDA_SOURCES if (CSRC / s).exists()]
    if not srcs:
        return None
    if force or _stale(out, srcs + _headers()):
        cc = nvcc()
        if cc is None:
            raise RuntimeError("nvcc not found; cannot build the sm_100a data plane")
        cmd = [cc, "-O3", "-std=c++17", "-lineinfo", *CUDA_ARCH_FLAGS, "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared",
               "-Xptxas", "-v", f"-I{ROOT / 'include'}", f"-I{CSRC}"]
        cmd += srcs + ["-o", out, "-lpthread", "-ldl"]
        _run(cmd)
        _stamp(out, srcs + _headers())
    return out


def build_cli(force: bool = False) -> list[Path]:
    BIN.mkdir(exist_ok=True)
    outs = []
    flags = ["-O2", "-g1", "-std=c++17", "-fPIC", "-Wall", "-Wno-unused-function", f"-I{ROOT / 'include'}"]
    core_srcs = [CSRC / s for s in CORE_SOURCES]
    for name in ("aar_master", "aar_worker"):
        src = CSRC / "cli" / (name + ".cpp")
        if not src.exists():
            continue
        out = BIN / name
        if force or _stale(out, core_srcs + [src] + _headers()):
            objs = _compile_objects(core_srcs + [src], flags, "cli", True)
            _run(["g++", *objs, "-o", out, "-lpthread"])
            _stamp(out, core_srcs + [src] + _headers())
        outs.append(out)
    return outs


def build_all(force: bool = False, only: str | None = None) -> None:
    if only in (None, "core"):
        build_core(force)
    if only in (None, "cuda"):
        build_cuda(force)
    if only in (None, "cli"):
        build_cli(force)


if __name__ == "__main__":
    args = sys.argv[1:]
    only = None
    if "--only" in args:
        only = args[args.index("--only") + 1]
    build_all(force="--force" in args, only=only)
