"""TEST INFRASTRUCTURE -- see cuda_runtime.h in this directory.

Rewrites the CUDA-only syntax of lammps-fugaku-threadpool_b200/csrc/*.cu (kernel launches, `extern __shared__`
arrays, the two inline-PTX helpers) into plain C++ and compiles the result with g++ against the emulation shim into
tests/cudaemu/_build/libb200md_emu.so, which exports the same C ABI as libb200md.so.  Only tests load it."""
from __future__ import annotations

import re
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parents[1]
CSRC = ROOT / "lammps-fugaku-threadpool_b200" / "csrc"
OUT = HERE / "_build"
LIB = OUT / "libb200md_emu.so"
FAKE_NCCL = OUT / "libnccl_emu.so"


def _match_back(text: str, end: int) -> int:
    """index of the `<` that opens the template argument list whose closing `>` is at text[end]"""
    depth = 0
    i = end
    while i >= 0:
        ch = text[i]
        if ch == ">":
            depth += 1
        elif ch == "<":
            depth -= 1
            if depth == 0:
                return i
        i -= 1
    raise ValueError("unbalanced template brackets before <<<")


def _match_fwd(text: str, start: int, open_ch: str, close_ch: str) -> int:
    depth = 0
    i = start
    while i < len(text):
        ch = text[i]
        if ch == open_ch:
            depth += 1
        elif ch == close_ch:
            depth -= 1
            if depth == 0:
                return i
        i += 1
    raise ValueError("unbalanced brackets after >>>")


def _split_top(s: str) -> list[str]:
    parts, depth, cur = [], 0, ""
    for ch in s:
        if ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            parts.append(cur.strip())
            cur = ""
        else:
            cur += ch
    parts.append(cur.strip())
    return parts


def rewrite_launches(text: str) -> str:
    out, pos = [], 0
    while True:
        k = text.find("<<<", pos)
        if k < 0:
            out.append(text[pos:])
            break
        # kernel expression: identifier, optionally followed by <template args>
        e = k - 1
        while text[e].isspace():
            e -= 1
        if text[e] == ">":
            e = _match_back(text, e) - 1
        while text[e].isspace():
            e -= 1
        b = e
        while b >= 0 and (text[b].isalnum() or text[b] in "_:"):
            b -= 1
        b += 1
        kernel = text[b:k].strip()
        close = text.index(">>>", k)
        cfg = _split_top(text[k + 3:close])
        grid, block = cfg[0], cfg[1]
        smem = cfg[2] if len(cfg) > 2 else "0"
        a0 = close + 3
        while text[a0].isspace():
            a0 += 1
        assert text[a0] == "(", (kernel, text[a0:a0 + 20])
        a1 = _match_fwd(text, a0, "(", ")")
        args = text[a0 + 1:a1]
        out.append(text[pos:b])
        out.append(f"emu::launch(dim3({grid}), dim3({block}), (size_t) ({smem}), [&]() {{ {kernel}({args}); }})")
        pos = a1 + 1
    return "".join(out)


def rewrite(text: str, name: str) -> str:
    # dynamic shared memory: one arena per launch
    text = re.sub(r"extern\s+__shared__\s+(\w+)\s+(\w+)\[\];", r"\1 *\2 = (\1 *) emu::dyn_smem;", text)
    # the two inline-PTX helpers of ctx.cuh
    text = re.sub(r'asm\("ld\.global\.nc\.v4\.f64[^;]*;"[^;]*;', "r = *p;", text)
    text = re.sub(r'asm\("rcp\.approx\.ftz\.f64[^;]*;"[^;]*;', "y = 1.0 / x;", text)
    # sm_100-only paths (TMA bulk copies + mbarrier) sit in `#ifndef B200MD_CUDAEMU ... #endif` blocks with a plain-C
    # equivalent beside them; the emulation compiles with -DB200MD_CUDAEMU=1, drop the blocks before the PTX check
    text = re.sub(r"#ifndef B200MD_CUDAEMU\n.*?\n#endif\n", "", text, flags=re.S)
    if "asm(" in text or "asm volatile" in text:
        raise RuntimeError(f"{name}: inline PTX the emulation does not know")
    # comm.cu binds NCCL by dlopen: hand it the in-process stand-in for emulated ranks (threads) instead
    text = text.replace('{"libnccl.so.2", "libnccl.so"}', '{"%s", "%s"}' % (FAKE_NCCL, FAKE_NCCL))
    return rewrite_launches(text)


def build(verbose: bool = False) -> Path:
    OUT.mkdir(exist_ok=True)
    srcs = sorted(CSRC.glob("*.cu"))
    deps = srcs + sorted(CSRC.glob("*.cuh")) + [HERE / "cuda_runtime.h", HERE / "emu.cpp", HERE / "nccl.h", HERE / "fake_nccl.cpp", Path(__file__),
                                                ROOT / "include" / "b200md.h"]
    if LIB.exists() and all(LIB.stat().st_mtime > d.stat().st_mtime for d in deps):
        return LIB
    gen = OUT / "csrc"
    gen.mkdir(exist_ok=True)
    for h in CSRC.glob("*.cuh"):
        (gen / h.name).write_text(rewrite(h.read_text(), h.name).replace('"../../include/b200md.h"', '"b200md.h"'))
    objs = []
    flags = ["-O2", "-g", "-std=c++17", "-fPIC", "-ffp-contract=off", "-fno-strict-aliasing", "-w", f"-I{HERE}", f"-I{gen}",
             f"-I{ROOT / 'include'}", "-DB200MD_CUDAEMU=1"]
    procs = []
    for s in srcs:
        cpp = gen / (s.stem + ".cpp")
        # the sources include "../../include/b200md.h" relative to csrc/: keep that path valid from _build/csrc/
        cpp.write_text(rewrite(s.read_text(), s.name).replace('"../../include/b200md.h"', '"b200md.h"'))
        obj = OUT / (s.stem + ".o")
        objs.append(obj)
        procs.append((s.name, subprocess.Popen(["g++", *flags, "-c", str(cpp), "-o", str(obj)], stdout=subprocess.PIPE,
                                               stderr=subprocess.STDOUT, text=True)))
    emu_o = OUT / "emu.o"
    procs.append(("emu.cpp", subprocess.Popen(["g++", *flags, "-c", str(HERE / "emu.cpp"), "-o", str(emu_o)],
                                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    bad = False
    for name, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            bad = True
            sys.stderr.write(f"--- cudaemu: g++ {name} ---\n{out[-6000:]}\n")
        elif verbose and out.strip():
            print(out)
    if bad:
        raise RuntimeError("cudaemu build failed")
    subprocess.check_call(["g++", "-O1", "-g", "-std=c++17", "-fPIC", "-shared", "-pthread", str(HERE / "fake_nccl.cpp"), "-o", str(FAKE_NCCL)])
    subprocess.check_call(["g++", "-shared", "-pthread", "-o", str(LIB), *map(str, objs), str(emu_o), "-ldl"])
    return LIB


if __name__ == "__main__":
    print(build(verbose=True))
