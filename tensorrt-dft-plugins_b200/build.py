"""Builds libtrt_dft_plugins.so in-tree for sm_100a (nvcc cross-compiles without a GPU).

    python tensorrt-dft-plugins_b200/build.py [--force] [--verbose]

Output: tensorrt-dft-plugins_b200/trt_dft_plugins/libtrt_dft_plugins.so -- the same file name,
next to the Python loader, as the reference (CMakeLists.txt:26, src/dft_plugins/CMakeLists.txt:38).
"""
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
OUT = HERE / "trt_dft_plugins" / "libtrt_dft_plugins.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17", "-t", "0",
    "-shared", "-Xcompiler", "-fPIC",
    "-I", str(HERE.parent / "include"),
]


def tensorrt_include():
    """The real <NvInfer.h> wins when available (TENSORRT_INCLUDE_DIR or a standard location);
    otherwise the plugin shell compiles against the stand-in in csrc/plugin/trt_min (test aid)."""
    import os
    cands = [os.environ.get("TENSORRT_INCLUDE_DIR"), "/usr/include/x86_64-linux-gnu", "/usr/include", "/usr/local/tensorrt/include"]
    for c in cands:
        if c and (Path(c) / "NvInfer.h").exists():
            return c
    return str(CSRC / "plugin" / "trt_min")


def sources():
    cu = sorted(CSRC.glob("*.cu")) + sorted((CSRC / "plugin").glob("*.cpp"))
    deps = list(CSRC.rglob("*.h")) + list(CSRC.rglob("*.cuh")) + list((HERE.parent / "include").glob("*.h")) + cu
    return cu, deps


# A/B switch, off by default: TRT_DFT_BUILD_F32X2=1 compiles the butterflies with packed fp32 arithmetic (add / sub / fma.f32x2 ->
# FADD2 / FFMA2).  Measured on B200: 18 % fewer instructions in the headline forward kernel but 0.138 -> 0.191 ms -- the
# packed forms have the scalar forms' lane throughput (tools/dev/ffma2_bench.cu) and evidently a longer dependent-issue
# latency, which these latency-bound kernels cannot hide.
import os as _os
if _os.environ.get("TRT_DFT_BUILD_F32X2"):
    NVCC_FLAGS.append("-DDFT_F32X2")


def build(force=False, verbose=False):
    cu, deps = sources()
    if not force and OUT.exists() and all(OUT.stat().st_mtime >= d.stat().st_mtime for d in deps + [Path(__file__)]):
        return OUT
    trt_inc = tensorrt_include()
    print(f"[build] plugin shell compiled against {trt_inc}/NvInfer.h"
          + (" (STAND-IN: creators register with the test double only)" if "trt_min" in trt_inc else ""), file=sys.stderr)
    cmd = ["nvcc"] + NVCC_FLAGS + ["-I", trt_inc] + (["-Xptxas", "-v"] if verbose else []) + ["-o", str(OUT)] + [str(c) for c in cu]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
