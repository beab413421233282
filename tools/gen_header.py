"""Regenerates include/dcp_b200.h from the extern "C" definitions under csrc/ (run from the repo root).
Hand-written additions go below the marker line `/* ---- hand-maintained additions ---- */` and are preserved."""
import glob
import re

PROLOGUE = """// Public C ABI of the native libraries (libdcp_host.so: host plumbing, g++; libdcp_cuda.so: sm_100a kernels, nvcc).
// Everything is plain extern "C": Python binds it with ctypes (_native.py, _cuda_sigs.py, ops/cuda.py, ops/tc.py) and
// C/C++ hosts can link it directly.  Device pointers are raw addresses; `stream` arguments are cudaStream_t.
// All int-returning entry points return 0 on success and -1 on failure (message via dcp_last_error()).
// GENERATED from the sources by tools/gen_header.py - the comments are the ones next to each definition.
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __CUDACC__
#include <cuda_runtime.h>
#else
typedef struct CUstream_st* cudaStream_t;
#endif

#ifdef __cplusplus
extern "C" {
#endif

struct DcpDevCtx;                   /* csrc/common.cuh */
typedef struct ConvP { int N, H, W, C, K, R, S, stride, pad, Ho, Wo; } ConvP;
"""
MARK = "/* ---- hand-maintained additions ---- */"


def main():
    try:
        old = open("include/dcp_b200.h").read()
        extra = old[old.index(MARK):old.rindex("#ifdef __cplusplus")] if MARK in old else MARK + "\n\n"
    except FileNotFoundError:
        extra = MARK + "\n\n"
    hdr = [PROLOGUE]
    for f in sorted(glob.glob("csrc/**/*.cu", recursive=True) + glob.glob("csrc/**/*.cpp", recursive=True)):
        src = open(f).read()
        decls = []
        for m in re.finditer(r'^(?:extern "C" )?((?:const )?(?:int|void\*?|size_t|int64_t|long|const char\*|const DcpDevCtx\*)\s+\*?)(dcp_\w+)\(([^)]*)\)\s*\{', src, re.M | re.S):
            if MARK_SKIP in src[max(0, m.start() - 200):m.start()]:
                continue
            ret, name, args = m.group(1).strip(), m.group(2), " ".join(m.group(3).split())
            pre = src[:m.start()].rstrip("\n").split("\n")
            com = []
            while pre and pre[-1].strip().startswith("//"):
                com.insert(0, pre.pop().strip())
            decls.append((com, ret.replace("const DcpDevCtx*", "const struct DcpDevCtx*"), name, args))
        if decls:
            hdr.append(f"/* ---- {f} ---- */")
            for com, ret, name, args in decls:
                hdr.extend(com)
                hdr.append(f"{ret} {name}({args});")
            hdr.append("")
    hdr.append(extra.rstrip("\n") + "\n")
    hdr.append('#ifdef __cplusplus\n}  /* extern "C" */\n#endif\n')
    open("include/dcp_b200.h", "w").write("\n".join(hdr))


MARK_SKIP = "gen_header: skip"

if __name__ == "__main__":
    main()
