"""In-tree build of the two native libraries.

* ``lib/libpdb200_rt.so``  — host runtime (g++): reducer, bucket planner, NUMA binding.
* ``lib/libpdb200.so``     — sm_100a CUDA kernels (nvcc), C ABI in ``include/pdb200.h``.

Both are plain shared objects loaded with ctypes: no torch C++ ABI coupling, nvcc compiles
each ``.cu`` in seconds, and the built ``.so`` travels with the source tree to the GPU box.
``python -m pytorch_distributed_b200.build`` builds everything; ``--force`` rebuilds.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import List

ROOT = Path(__file__).resolve().parent.parent
PKG = Path(__file__).resolve().parent
LIB = PKG / "lib"
OBJ = ROOT / "build" / "obj"
INCLUDE = ROOT / "include"
CSRC = ROOT / "csrc"

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# No --use_fast_math: the numerics oracles are fp32 PyTorch, and flushing denormals to zero would eat fp16
# loss-scaled gradients.  The hot paths name their approximate instructions explicitly (rsqrtf, exp2f, fmaf).
# -ftz=true alone IS kept: it flushes fp32 denormals only (< 1.2e-38, far below anything a loss-scaled fp16 gradient
# turns into when widened to fp32; fp16 subnormals survive the round trip, see test_fp16_subnormals_survive_bn_backward)
# and without it the BN backward apply kernel that also writes the residual gradient is 8 % slower (0.7 % of the step,
# same-box A/B with bench_tools/ab_bn.py).  Division, sqrt, exp and log stay IEEE / accurate.
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-ftz=true", "-Xcompiler", "-fPIC",
              "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def _nvcc() -> str:
    for c in (os.environ.get("CUDA_HOME", "") + "/bin/nvcc", "/usr/local/cuda/bin/nvcc", shutil.which("nvcc") or ""):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found (set CUDA_HOME)")


def _stamp(paths: List[Path], extra: str = "") -> str:
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        h.update(str(p.relative_to(ROOT)).encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _headers() -> List[Path]:
    return sorted(list(INCLUDE.glob("*.h")) + list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.h")))


def build_runtime(force: bool = False, verbose: bool = False) -> Path:
    out = LIB / "libpdb200_rt.so"
    srcs = sorted((CSRC / "runtime").glob("*.cpp"))
    stamp = _stamp(srcs + _headers(), "rt")
    sf = LIB / ".rt.stamp"
    if out.exists() and not force and sf.exists() and sf.read_text() == stamp:
        return out
    LIB.mkdir(parents=True, exist_ok=True)
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", f"-I{INCLUDE}", *map(str, srcs), "-o", str(out), "-pthread"]
    if verbose:
        print(" ".join(cmd))
    subprocess.run(cmd, check=True)
    sf.write_text(stamp)
    return out


def cuda_sources() -> List[Path]:
    return sorted(list((CSRC / "ops").glob("*.cu")) + list((CSRC / "comm").glob("*.cu")))


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    out = LIB / "libpdb200.so"
    srcs = cuda_sources()
    if not srcs:
        raise RuntimeError("no CUDA sources found")
    hdr_stamp = _stamp(_headers(), "cu" + " ".join(NVCC_FLAGS + NVCC_ARCH))
    OBJ.mkdir(parents=True, exist_ok=True)
    LIB.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()
    jobs = []
    for s in srcs:
        o = OBJ / (s.stem + ".o")
        st = OBJ / (s.stem + ".stamp")
        want = hashlib.sha256((hdr_stamp + s.read_text()).encode()).hexdigest()
        if force or not o.exists() or not st.exists() or st.read_text() != want:
            jobs.append((s, o, st, want))

    def compile_one(job):
        s, o, st, want = job
        cmd = [nvcc, *NVCC_ARCH, *NVCC_FLAGS, f"-I{INCLUDE}", f"-I{CSRC}", "-c", str(s), "-o", str(o)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = OBJ / (s.stem + ".ptxas.log")
        log.write_text(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {s.name}:\n{r.stdout}\n{r.stderr}")
        st.write_text(want)
        return s.name

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for name in ex.map(compile_one, jobs):
                if verbose:
                    print("compiled", name)
    if jobs or not out.exists() or force:
        objs = [str(OBJ / (s.stem + ".o")) for s in srcs]
        cuda_lib = str(Path(nvcc).resolve().parent.parent / "lib64")
        cmd = [nvcc, "-shared", *NVCC_ARCH, "-o", str(out), *objs, f"-L{cuda_lib}", "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return out


def build_all(force: bool = False, verbose: bool = False):
    return build_runtime(force, verbose), build_cuda(force, verbose)


if __name__ == "__main__":
    force = "--force" in sys.argv
    rt, cu = build_all(force=force, verbose=True)
    print("built", rt)
    print("built", cu)
