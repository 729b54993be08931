#!/usr/bin/env python
"""Builds the native layer for sm_100a with nvcc (cross-compiles without a GPU).

    python build.py            -> distributed_dot_product/_lib/libddp_b200.so
    python build.py harness    -> also build/gemm_harness (standalone CUDA validation binary)
    python build.py sass       -> also profiles/sass/*.sass.txt (SASS mnemonic evidence)

Objects are cached by mtime under build/obj.
"""
import concurrent.futures as cf
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ARCH + ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
                "-Xptxas", "-v"]
LIB_SOURCES = [
    "csrc/runtime/runtime.cu",
    "csrc/runtime/vmm_heap.cu",
    "csrc/gemm/gemm_api.cu",
    "csrc/softmax/softmax.cu",
    "csrc/reduce/colsum.cu",
    "csrc/attn/attn_api.cu",
]
HEADERS = [
    "include/ddp_b200.h",
    "csrc/common/ptx.cuh",
    "csrc/common/host_util.h",
    "csrc/gemm/gemm_sm100.cuh",
    "csrc/gemm/tma_map.h",
    "csrc/attn/attn_fwd_sm100.cuh",
    "csrc/attn/attn_bwd_sm100.cuh",
    "csrc/attn/attn_fwd2_sm100.cuh",
    "csrc/attn/attn_bwd2_sm100.cuh",
]
LIB_OUT = "distributed_dot_product/_lib/libddp_b200.so"
# `python build.py trace`: separate objects + build/attn_harness_trace with the pipeline time stamps of the
# attention forward compiled in (ddp_attn_debug_trace); the production library never carries them.
TRACE = False


def _newer(src_list, out):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(os.path.join(ROOT, s)) > t for s in src_list if os.path.exists(os.path.join(ROOT, s)))


def _run(cmd, log=None):
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log:
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("command failed: " + " ".join(cmd))
    return r.stdout


def compile_obj(src):
    obj = os.path.join(ROOT, "build/obj_trace" if TRACE else "build/obj", src.replace("/", "_") + ".o")
    os.makedirs(os.path.dirname(obj), exist_ok=True)
    if _newer([src] + HEADERS, obj):
        _run([NVCC] + FLAGS + (["-DDDP_ATTN_TRACE"] if TRACE else []) + ["-c", src, "-o", obj], log=obj + ".log")
    return obj


def build_lib(verbose=True):
    srcs = [s for s in LIB_SOURCES if os.path.exists(os.path.join(ROOT, s))]
    with cf.ThreadPoolExecutor(max_workers=4) as ex:
        objs = list(ex.map(compile_obj, srcs))
    out = os.path.join(ROOT, LIB_OUT)
    os.makedirs(os.path.dirname(out), exist_ok=True)
    if _newer([os.path.relpath(o, ROOT) for o in objs], out):
        _run([NVCC] + ARCH + ["-shared", "-o", out] + objs + ["-lcudart_static", "-ldl", "-lrt", "-lpthread"])
    if verbose:
        print("built", LIB_OUT)
    return out


def build_harness():
    srcs = [s for s in LIB_SOURCES if os.path.exists(os.path.join(ROOT, s))]
    objs = [compile_obj(s) for s in srcs]
    os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
    for name in ("gemm_harness", "attn_harness"):
        src = "csrc/tests/%s.cu" % name
        if not os.path.exists(os.path.join(ROOT, src)):
            continue
        out = os.path.join(ROOT, "build", name + ("_trace" if TRACE else ""))
        if _newer([src] + HEADERS + [os.path.relpath(o, ROOT) for o in objs], out):
            _run([NVCC] + ARCH + ["-O2", "-std=c++17", src, "-o", out] + objs)
        print("built build/" + name)


def dump_sass():
    out_dir = os.path.join(ROOT, "profiles/sass")
    os.makedirs(out_dir, exist_ok=True)
    lib = os.path.join(ROOT, LIB_OUT)
    sass = _run(["/usr/local/cuda/bin/cuobjdump", "-sass", lib])
    import collections
    import re
    per_kernel = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            per_kernel[cur] = collections.Counter()
            continue
        if cur:
            m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
            if m:
                per_kernel[cur][m.group(1).split(".")[0]] += 1
    keys = ["UTCHMMA", "UTCQMMA", "UTCMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "HMMA", "LDG",
            "STG", "ATOMG", "RED", "MUFU"]
    with open(os.path.join(out_dir, "mnemonics.txt"), "w") as f:
        f.write("# SASS mnemonic counts per kernel in %s (cuobjdump -sass)\n" % LIB_OUT)
        for k, cnt in per_kernel.items():
            tc = {m: n for m, n in cnt.items() if any(m.startswith(p) for p in keys)}
            f.write("%s\n    %s\n" % (k, " ".join("%s=%d" % kv for kv in sorted(tc.items()))))
    print("wrote profiles/sass/mnemonics.txt (%d kernels)" % len(per_kernel))
    # trimmed listings: every tcgen05 / TMA / mbarrier instruction of the hot kernels with its neighbourhood
    want = {"attn_fwd2_kernelILi1ELi4ELb0": "attn_fwd2_bf16", "attn_bwd2_rows_kernelILi1": "attn_bwd2_rows_bf16",
            "attn_bwd2_cols_kernelILi1": "attn_bwd2_cols_bf16", "gemm_kernelILi0ELi1ELi256ELi4ELi0ELi4": "gemm_nt_bf16_bn256",
            "gemm_kernelILi1ELi1ELi256ELi4ELi0ELi4": "gemm_nn_bf16_bn256", "gemm_kernelILi2ELi1ELi128ELi6ELi0ELi4": "gemm_tn_bf16_bn128",
            "gemm_kernelILi0ELi2ELi256": "gemm_nt_tf32_bn256"}
    hot = re.compile(r"UTC|LDTM|STTM|UTMA|UBLKCP|SYNCS|MUFU|FFMA2|FADD2|FMUL2|ELECT|RED|ATOM|MEMBAR|FENCE")
    cur, buf = None, []

    def flush():
        if not cur:
            return
        tag = next((v for k, v in want.items() if k in cur), None)
        if tag is None or os.path.exists(os.path.join(out_dir, tag + ".sass.txt.done")):
            return
        keep = set()
        for i, ln in enumerate(buf):
            if hot.search(ln):
                keep.update(range(max(0, i - 2), min(len(buf), i + 3)))
        with open(os.path.join(out_dir, tag + ".sass.txt"), "w") as f:
            f.write("# %s\n# cuobjdump -sass of %s, trimmed to the tcgen05 / TMA / mbarrier / MUFU instructions and two\n"
                    "# neighbours each ('...' = omitted lines); %d of %d instructions shown\n" % (cur, LIB_OUT, len(keep), len(buf)))
            last = -1
            for i in sorted(keep):
                if i != last + 1:
                    f.write("        ...\n")
                f.write(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", buf[i]) + "\n")
                last = i
        print("wrote profiles/sass/%s.sass.txt" % tag)

    seen = set()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            flush()
            cur, buf = m.group(1), []
            tag = next((v for k, v in want.items() if k in cur), None)
            if tag in seen:
                cur = None
            elif tag:
                seen.add(tag)
            continue
        if cur and re.match(r"^\s+/\*[0-9a-f]{4,}\*/", line):
            buf.append(line.rstrip())
    flush()


if __name__ == "__main__":
    if "trace" in sys.argv[1:]:
        TRACE = True
        build_harness()
        sys.exit(0)
    build_lib()
    if "harness" in sys.argv[1:]:
        build_harness()
    if "sass" in sys.argv[1:]:
        dump_sass()
