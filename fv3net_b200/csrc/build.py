"""Builds fv3net_b200/libb200cs.so with nvcc for sm_100a (cross-compiles without a GPU).

One object per translation unit, compiled in parallel; the fused pressure-coarsening kernel is
instantiated per (window width, arithmetic mode): EXACT units use -fmad=false (bit parity with
the no-FMA fp32 restatement of mappm.f90), FAST units allow FMA contraction.
"""
import concurrent.futures
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
OUT = os.path.join(PKG, "libb200cs.so")
OBJ = os.path.join(HERE, "build")
NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
HOSTCXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else (shutil.which("g++") or "g++")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-ccbin", HOSTCXX,
          "-Xcompiler", "-ffp-contract=off", "-Xptxas", "-v"]

UNITS = [("status", "status.cu", ["-fmad=false"]),
         ("block_ops", "block_ops.cu", ["-fmad=false"]),
         ("stagger", "stagger.cu", ["-fmad=false"]),
         ("mappm", "mappm.cu", ["-fmad=false"]),
         ("mappm_fast", "mappm_fast.cu", ["-fmad=true"]),
         ("column_ops", "column_ops.cu", ["-fmad=false"]),
         ("elementwise", "elementwise.cu", ["-fmad=false"]),
         ("pc_api", "pressure_coarsen_api.cu", ["-fmad=false"]),
         ("host_stage", "host_stage.cu", ["-fmad=false"]),
         ("pc_fast", "pressure_coarsen_fast.cu", ["-fmad=true"])]
for fx in (2, 4, 8):
    UNITS.append((f"pc_fx{fx}_m0", "pressure_coarsen_inst.cu", ["-fmad=false", f"-DB200CS_FX={fx}", "-DB200CS_MODE=0"]))
    UNITS.append((f"pc_fx{fx}_m1", "pressure_coarsen_inst.cu", ["-fmad=true", f"-DB200CS_FX={fx}", "-DB200CS_MODE=1"]))

HEADERS = ["common.cuh", "ppm_stream.cuh", "ppm_fast.cuh", "pressure_coarsen.cuh", "pressure_coarsen_fast.cuh", "pressure_coarsen_sweep.cuh", "ppm_sweep.cuh", "pressure_coarsen_tma.cuh", os.path.join("experimental", "pressure_coarsen_ws.cuh"), "ppm_pack.cuh", os.path.join("..", "..", "include", "b200cs.h")]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(unit):
    name, src, flags = unit
    obj = os.path.join(OBJ, name + ".o")
    deps = [os.path.join(HERE, src), os.path.abspath(__file__)] + [os.path.join(HERE, h) for h in HEADERS]
    if not _stale(obj, deps):
        return name, "", False
    extra = os.environ.get("B200CS_DEV_NVCC_FLAGS", "").split()  # developer experiments only (e.g. -DB200CS_TMA_HALVES=1)
    cmd = [NVCC] + ARCH + COMMON + flags + extra + ["-c", os.path.join(HERE, src), "-o", obj]
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError(f"nvcc failed for {name}:\n{' '.join(cmd)}\n{p.stdout}\n{p.stderr}")
    return name, p.stderr, True


def build(verbose=False, force=False):
    os.makedirs(OBJ, exist_ok=True)
    if force:
        for f in os.listdir(OBJ):
            os.remove(os.path.join(OBJ, f))
    logs = {}
    rebuilt = False
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        for name, log, did in ex.map(_compile, UNITS):
            logs[name] = log
            rebuilt |= did
    objs = [os.path.join(OBJ, u[0] + ".o") for u in UNITS]
    if rebuilt or _stale(OUT, objs):
        cmd = [NVCC] + ARCH + ["-shared", "-ccbin", HOSTCXX, "-o", OUT] + objs + ["-lcudart", "-lpthread"]
        p = subprocess.run(cmd, capture_output=True, text=True)
        if p.returncode != 0:
            raise RuntimeError(f"link failed:\n{p.stdout}\n{p.stderr}")
    if verbose:
        with open(os.path.join(OBJ, "ptxas.log"), "w") as f:
            for name, log in logs.items():
                if log:
                    f.write(f"==== {name}\n{log}\n")
    return OUT


if __name__ == "__main__":
    print(build(verbose=True, force="--force" in sys.argv))
