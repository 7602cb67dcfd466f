#!/usr/bin/env python
"""Opcode histogram of the kernels the library SELECTS BY DEFAULT (tuning.cu) at 2^20, 2^25 and 2^29 elements, fp32 and fp64,
plus the two async-proxy copy variants -- from `cuobjdump -sass` of the objects `make lib` leaves in build/obj.

    python tools/sass_summary.py > profiles/r03/sass_summary.txt        (or: make sass)

What to look for: LDG.E...128 / ENL2.256 (vector loads and their cache policy), STG, DFMA/FFMA (one per element in
triad/nstream/dot with contraction on), SHFL + ATOMG/RED (dot's reduction and tickets), and for the copy variants
UBLKCP (cp.async.bulk), UTMALDG/UTMASTG (tensor-map loads/stores) and SYNCS (mbarrier).  No GPU needed.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import babelstream_b200 as bs  # noqa: E402

OBJ = os.path.join(ROOT, "build", "obj")
OP = {"copy": 0, "mul": 1, "add": 2, "triad": 3, "nstream": 4, "init": 6}
INTEREST = ("LDG", "STG", "DFMA", "DMUL", "DADD", "FFMA", "FMUL", "FADD", "SHFL", "ATOMG", "ATOM", "RED", "MEMBAR", "BAR",
            "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "LDS", "STS", "CCTL", "ACQBULK", "GRIDDEPCONTROL", "LDGDEPBAR", "DEPBAR")


def functions(obj):
    """{demangled name: [opcode, ...]} of one object file"""
    text = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
    names = re.findall(r"Function : (\S+)", text)
    dem = subprocess.run(["cu++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout.splitlines()
    out, cur = {}, None
    it = iter(dem)
    for line in text.splitlines():
        if "Function :" in line:
            cur = out.setdefault(next(it), [])
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and cur is not None:
            cur.append(m.group(1))
    return out


def histogram(ops):
    h = collections.Counter()
    for o in ops:
        base = o.split(".")[0]
        if base in INTEREST:
            h[o] += 1
    return h


def show(title, name, ops):
    print(f"## {title}")
    print(f"   {name}")
    h = histogram(ops)
    print(f"   {len(ops)} instructions; " + ", ".join(f"{k} x{v}" for k, v in sorted(h.items())))
    print()


def main():
    cache = {}

    def funcs(obj):
        if obj not in cache:
            cache[obj] = functions(os.path.join(OBJ, obj))
        return cache[obj]

    print("# SASS summary of the default-selected instantiations (sm_100a, cuobjdump -sass build/obj/*.o)")
    print(f"# library: {bs.lib().b200_version().decode()}\n")
    for log2n in (20, 25, 29):
        for tname, eb in (("double", 8), ("float", 4)):
            n = 1 << log2n
            for k in ("copy", "mul", "add", "triad", "nstream"):
                cfg, grid = bs.resolve_cfg(k, n, eb)
                want = f"b200::ew_kernel<{tname}, (int){OP[k]}, (int){cfg.vec_bytes}, (int){cfg.unroll}, (int){cfg.cache_hint}, (bool){1 if cfg.contract != 2 else 0}>"
                hit = [(nm, ops) for nm, ops in funcs(f"ew_{k}.o").items() if want in nm]
                title = f"{k} {tname} 2^{log2n}: vec {cfg.vec_bytes} B x unroll {cfg.unroll}, block {cfg.block}, hint {cfg.cache_hint}, grid {grid}"
                if hit:
                    show(title, hit[0][0], hit[0][1])
                else:
                    print(f"## {title}\n   NOT FOUND: {want}\n")
            cfg, grid = bs.resolve_cfg("dot", n, eb)
            mode = 1 if cfg.variant == 2 else 0
            want = f"b200::dot_kernel<{tname}, (int){cfg.vec_bytes}, (int){cfg.unroll}, (int){cfg.cache_hint}, (bool){1 if cfg.contract != 2 else 0}, (int){mode}>"
            hit = [(nm, ops) for nm, ops in funcs("dot.o").items() if want in nm]
            title = f"dot {tname} 2^{log2n}: vec {cfg.vec_bytes} B x unroll {cfg.unroll}, block {cfg.block}, hint {cfg.cache_hint}, variant {cfg.variant}, grid {grid}"
            if hit:
                show(title, hit[0][0], hit[0][1])
            else:
                print(f"## {title}\n   NOT FOUND: {want}\n")
    for nm, ops in funcs("copy_async.o").items():
        show("copy variant through the async proxy (opt-in: copy_variant=bulk|tma)", nm, ops)
    for nm, ops in funcs("dot.o").items():
        if "dot_final_kernel" in nm or "dot_collect_kernel" in nm:
            show("dot: second pass / peer-memory collect", nm, ops)
    total = sum(len(funcs(o)) for o in cache)
    print(f"# {total} kernels in the objects read; library size {os.path.getsize(bs.lib_path()) / 1e6:.1f} MB")


if __name__ == "__main__":
    main()
