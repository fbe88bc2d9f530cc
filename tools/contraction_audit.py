#!/usr/bin/env python
"""Which product of `a*b + c*d + e*f` stays rounded?  nvcc decides per call site and per template instantiation.

Compiles the reference's kernels (oracle/ref_harness.cu, i.e. /root/reference/src/ACMMP.cu where it lies) and this
repository's kernels.cu to PTX with their own flags and prints, for Vec3DotVec3(prior, plane) - ACMMP.cu:95, inlined at
:765, :1110, :1121 - and for prior_dot (sweep_kernel.cu.inc), the order in which the x / y / z products enter the chain:
`mul` = the product that is rounded on its own, `fma` = fused.  Components are recovered from the vector loads that define
the operands.  Needs /root/reference (build container only); output: profiles/r2_contraction_audit.txt."""
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/src"


def ptx(cmd):
    with tempfile.NamedTemporaryFile(suffix=".ptx") as f:
        subprocess.run(cmd + ["-ptx", "-o", f.name], check=True, cwd=ROOT, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        return open(f.name).read()


def sites(txt, want):
    files = {m.group(1): m.group(2).split("/")[-1] for m in re.finditer(r'\.file\s+(\d+)\s+"([^"]+)"', txt)}
    for e in re.split(r"\n\.visible \.entry ", txt)[1:]:
        name = e.split("(")[0]
        comp = {}
        for m in re.finditer(r"ld\.\w+(?:\.\w+)*\.v4\.f32\s+\{(%f\d+), (%f\d+), (%f\d+), (%f\d+)\}", e):
            for r, c in zip(m.groups(), "xyzw"):
                comp[r] = c
        loc, out = None, []
        for l in e.split("\n"):
            m = re.match(r"\s*\.loc\s+(\d+)\s+(\d+)\s+\d+(?:.*inlined_at\s+\d+\s+(\d+))?", l)
            if m:
                loc = (files.get(m.group(1), "?"), int(m.group(2)), int(m.group(3)) if m.group(3) else None)
                continue
            if loc and want(loc):
                m = re.match(r"\s+(mul\.ftz\.f32|fma\.rn\.ftz\.f32|mul\.rn\.ftz\.f32)\s+(%f\d+), (%f\d+), (%f\d+)", l)
                if m:
                    out.append((loc[2] or loc[1], m.group(1)[:3], comp.get(m.group(3), "?") + comp.get(m.group(4), "?")))
        if out:
            yield name, out


def chains(ops):
    """group (site, op, comps) into chains mul fma fma"""
    res, cur = [], None
    for site, op, c in ops:
        if op == "mul":
            cur = [site, [c]]
            res.append(cur)
        elif cur is not None and cur[0] == site:
            cur[1].append(c)
    return res


def main():
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference")
    ref = ptx(["nvcc", "-O3", "--use_fast_math", "--maxrregcount=128", "-std=c++14", "-gencode", "arch=compute_100,code=compute_100",
               "-lineinfo", "-w", "-Ioracle/ref_stub", "-I" + REF, "-Iinclude", "oracle/ref_harness.cu"])
    print("reference build (ACMMP.cu:95 Vec3DotVec3, by inlining site):")
    for name, ops in sites(ref, lambda loc: loc[0] == "ACMMP.cu" and loc[1] == 95 and loc[2] in (765, 1110, 1121)):
        seen = {}
        for site, c in chains(ops):
            seen.setdefault((site, " ".join(c)), 0)
            seen[(site, " ".join(c))] += 1
        for (site, c), n in sorted(seen.items()):
            print("  %-18s ACMMP.cu:%d  rounded product, then fused: %s   (%d chains)" % (name[3:19], site, c, n))
    ours = ptx(["nvcc", "-O3", "--use_fast_math", "-std=c++17", "-gencode", "arch=compute_100a,code=compute_100a", "-lineinfo", "-Iinclude",
                "acmmp_b200/csrc/kernels.cu"])
    src = open(os.path.join(ROOT, "acmmp_b200/csrc/sweep_kernel.cu.inc")).read().split("\n")
    line = 1 + next(i for i, l in enumerate(src) if l.startswith("ACMMP_DEV float prior_dot("))
    print("this repository (prior_dot, sweep_kernel.cu.inc:%d, explicit fmul / ffma), per instantiation of sweep_kernel:" % line)
    # the intrinsics live in common.cuh; their .loc carries `inlined_at <sweep_kernel.cu.inc> <line of prior_dot>`
    for name, ops in sites(ours, lambda loc: loc[0] == "common.cuh" and loc[2] == line):
        if "sweep_kernel" not in name:
            continue
        ops = [(s, "mul" if o == "mul" else "fma", c) for s, o, c in ops]
        seen = {}
        for site, c in chains(ops):
            seen[" ".join(c)] = seen.get(" ".join(c), 0) + 1
        inst = re.search(r"sweep_kernelI(\w+?)EEv", name)
        print("  sweep_kernel<%s>  %s" % (inst.group(1) if inst else name, ", ".join("%s (%d)" % kv for kv in seen.items())))


if __name__ == "__main__":
    main()
