#!/usr/bin/env python
"""TEST INFRASTRUCTURE ONLY: build tests/emu/_build/libnhstep_emu.so — the sources of oceananigans.jl_b200/csrc/, rewritten just
enough for g++ (kernel launches, dynamic shared memory, inline PTX) and compiled against the CUDA stand-in emu_cuda.h.

    python tests/emu/build_emu.py            # → tests/emu/_build/libnhstep_emu.so

The shipped sources are not modified; the rewritten copies live under tests/emu/_build/ (git-ignored).  See README.md here."""
import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SRC = os.path.join(ROOT, "oceananigans.jl_b200", "csrc")
OUT = os.environ.get("NHSTEP_EMU_BUILD", os.path.join(HERE, "_build"))
UNITS = ["api.cu", "model.cu", "dist.cu", "tendency_fma.cu", "solver_generic.cu", "no_alloc.cu", "grid_host.cpp"]


def _matching(text, at, open_ch, close_ch):
    """Index just past the bracket that closes the one at text[at]."""
    depth = 0
    for i in range(at, len(text)):
        if text[i] == open_ch:
            depth += 1
        elif text[i] == close_ch:
            depth -= 1
            if depth == 0:
                return i + 1
    raise ValueError("unbalanced " + open_ch)


def _split_top(text):
    parts, depth, cur = [], 0, ""
    for ch in text:
        if ch in "([{<":
            depth += 1
        elif ch in ")]}>":
            depth -= 1
        if ch == "," and depth == 0:
            parts.append(cur.strip())
            cur = ""
        else:
            cur += ch
    parts.append(cur.strip())
    return parts


def rewrite_launches(text, name):
    out, pos = "", 0
    while True:
        at = text.find("<<<", pos)
        if at < 0:
            return out + text[pos:]
        # kernel expression: identifier, optionally with template arguments, right before <<<
        k = at
        if text[k - 1] == ">":
            depth = 0
            while True:
                k -= 1
                if text[k] == ">":
                    depth += 1
                elif text[k] == "<":
                    depth -= 1
                    if depth == 0:
                        break
        while k > 0 and (text[k - 1].isalnum() or text[k - 1] in "_:"):
            k -= 1
        kernel = text[k:at]
        end_cfg = text.index(">>>", at)
        cfg = _split_top(text[at + 3:end_cfg])
        if not 2 <= len(cfg) <= 4:
            raise ValueError(f"{name}: launch configuration {cfg}")
        lp = end_cfg + 3
        while text[lp].isspace():
            lp += 1
        assert text[lp] == "(", (name, text[at - 40:at + 80])
        rp = _matching(text, lp, "(", ")")
        args = text[lp + 1:rp - 1]
        shared = cfg[2] if len(cfg) > 2 else "0"
        out += text[pos:k] + (f"emu::launch(dim3({cfg[0]}), dim3({cfg[1]}), (size_t)({shared}), [&]() {{ {kernel}({args}); }}, "
                              f"(const void*)({kernel}), \"{kernel}\")")
        pos = rp


def rewrite_asm(text, name):
    out, pos = "", 0
    for m in re.finditer(r"\basm\s*(?:volatile\s*)?\(", text):
        if m.start() < pos:
            continue
        end = _matching(text, m.end() - 1, "(", ")")
        body = text[m.end():end - 1]
        semi = end
        while text[semi].isspace():
            semi += 1
        assert text[semi] == ";", (name, body[:60])
        ops = re.findall(r'"[=+]?[a-z]"\s*\(((?:[^()]|\([^()]*\))*)\)', body)
        if "rcp.approx.ftz.f64" in body:
            new = f"{ops[0]} = 1.0 / ({ops[1]});"
        elif "prefetch" in body:
            new = ";"
        elif "globaltimer" in body:
            new = f"{ops[0]} = 0;"
        elif "st.release.sys.global.u32" in body:
            new = f"*(volatile unsigned*)({ops[0]}) = ({ops[1]});"
        elif "ld.acquire.sys.global.u32" in body:
            new = f"{ops[0]} = *(volatile unsigned*)({ops[1]}); emu_relax();"      # a polling load of the flag barrier
        elif "mbarrier.init" in body:
            new = f"emu::mbar_init({ops[0]}, {ops[1]});"
        elif "mbarrier.arrive.expect_tx" in body:
            new = f"emu::mbar_expect_tx({ops[0]}, {ops[1]});"
        elif "mbarrier.try_wait.parity" in body:
            new = f"emu::mbar_wait({ops[0]}, {ops[1]});"
        elif "mbarrier.arrive.shared" in body:
            new = f"emu::mbar_arrive({ops[0]});"
        elif "cp.async.bulk.tensor.3d" in body:
            new = f"emu::tma_load_3d({ops[0]}, {ops[1]}, {ops[2]}, {ops[3]}, {ops[4]}, {ops[5]});"
        elif body.lstrip().startswith('"fence.'):
            new = ";"
        else:                                                   # anything else is not emulated
            new = 'emu_unsupported("' + re.sub(r"[^A-Za-z0-9_.: ]", "", body.split(";")[0])[:60] + '");'
        out += text[pos:m.start()] + new
        pos = semi + 1
    return out + text[pos:]


def rewrite(text, name):
    text = re.sub(r'#include\s*<(cuda_runtime|cuda|cudaTypedefs)\.h>', "", text)
    text = re.sub(r"extern\s+__shared__\s+(?:__align__\(\d+\)\s+)?unsigned char\s+(\w+)\[\];", r"unsigned char* \1 = emu::dynamic_shared();", text)
    text = rewrite_asm(text, name)
    return rewrite_launches(text, name)


def build(verbose=True):
    csrc = os.path.join(OUT, "oceananigans.jl_b200", "csrc")
    os.makedirs(csrc, exist_ok=True)
    os.makedirs(os.path.join(OUT, "include"), exist_ok=True)
    shutil.copy(os.path.join(ROOT, "include", "nhstep.h"), os.path.join(OUT, "include", "nhstep.h"))
    for fn in sorted(os.listdir(SRC)):
        if fn.endswith((".cu", ".cuh", ".cpp", ".h")):
            with open(os.path.join(SRC, fn)) as fh:
                text = fh.read()
            new = rewrite(text, fn) if fn.endswith((".cu", ".cuh")) else text
            dst = os.path.join(csrc, fn + (".cpp" if fn.endswith(".cu") else ""))
            if not os.path.exists(dst) or open(dst).read() != new:
                with open(dst, "w") as fh:
                    fh.write(new)
    objs = []
    flags = ["-std=c++17", "-O1", "-g", "-fPIC", "-D__CUDACC__", "-DNH_EMU=1", "-ffp-contract=off", "-fno-strict-aliasing", "-Wno-unknown-pragmas", "-Wno-attributes",
             "-I", HERE, "-include", os.path.join(HERE, "emu_cuda.h"), "-include", os.path.join(HERE, "emu_extra.h")]
    jobs = []
    for unit in UNITS + ["emu_runtime.cpp"]:
        src = os.path.join(HERE, unit) if unit == "emu_runtime.cpp" else os.path.join(csrc, unit + (".cpp" if unit.endswith(".cu") else ""))
        obj = os.path.join(OUT, unit + ".o")
        objs.append(obj)
        deps = [src] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cuh", ".h"))] + \
               [os.path.join(HERE, "emu_cuda.h"), os.path.join(HERE, "emu_extra.h")]
        if os.path.exists(obj) and all(os.path.getmtime(obj) >= os.path.getmtime(d) for d in deps):
            continue
        jobs.append((unit, subprocess.Popen(["g++"] + flags + ["-c", src, "-o", obj], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    failed = False
    for unit, p in jobs:
        log = p.communicate()[0].decode()
        if p.returncode != 0:
            failed = True
            print(f"---- {unit}: g++ failed\n{log[:6000]}")
        elif verbose:
            print(f"compiled {unit}")
    if failed:
        raise SystemExit(1)
    lib = os.path.join(OUT, "libnhstep_emu.so")
    if not os.path.exists(lib) or any(os.path.getmtime(o) > os.path.getmtime(lib) for o in objs):
        tmp = lib + f".{os.getpid()}.tmp"                         # linked aside and renamed: a concurrent loader never sees half a file
        subprocess.check_call(["g++", "-shared", "-o", tmp] + objs)
        os.replace(tmp, lib)
    return lib


if __name__ == "__main__":
    print(build())
    sys.exit(0)
