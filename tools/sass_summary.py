"""Per-kernel opcode summary of libminipic_b200.so (commit SASS evidence instead of asking the reader to run
cuobjdump).  For every kernel in the in-tree library: registers, spill bytes and shared memory from the ptxas logs the Makefile
keeps (csrc/*.ptxas.log), and from `cuobjdump -sass` the number of instructions per family --

  UTMALDG       TMA tensor loads (cp.async.bulk.tensor)          LDGSTS        cp.async (global -> shared without registers)
  SYNCS         mbarrier operations                              RED / ATOM    reductions / atomics, .64 counted apart
  DFMA DADD DMUL  FP64 pipe                                      F2I I2F FRND  conversions (XU pipe)
  LDS STS LDG STG  shared / global loads and stores              SHFL BAR      warp shuffles, block barriers

Static counts of the binary, not execution counts (those are in the ncu source-page summaries under profiles/).
usage: python tools/sass_summary.py [out.txt]      (needs cuobjdump + c++filt on PATH; no GPU)
"""
import collections
import glob
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "mini-pic--2d-pic-code-with-tiling-load-balance_b200")
LIB = os.path.join(PKG, "libminipic_b200.so")
FAMILIES = ["UTMALDG", "LDGSTS", "SYNCS", "RED", "RED.64", "ATOM", "ATOM.64", "DFMA", "DADD", "DMUL", "MUFU", "F2I", "I2F", "FRND",
            "LDS", "STS", "LDG", "STG", "SHFL", "BAR", "total"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    short = []
    for d in out[:len(names)]:
        d = re.sub(r"\(anonymous namespace\)::", "", d)
        d = re.sub(r"^void ", "", d)
        short.append(re.sub(r"\(.*$", "", d))
    return short


def ptxas_info():
    """mangled name -> (registers, spill bytes, shared bytes) from csrc/*.ptxas.log"""
    info = {}
    for log in glob.glob(os.path.join(PKG, "csrc", "*.ptxas.log")):
        cur = None
        spill = 0
        for line in open(log):
            m = re.search(r"Compiling entry function '([^']+)'", line)
            if m:
                cur, spill = m.group(1), 0
            m = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m and cur:
                spill = int(m.group(1)) + int(m.group(2))
            m = re.search(r"Used (\d+) registers", line)
            if m and cur:
                sm = re.search(r"(\d+) bytes smem", line)
                info[cur] = (int(m.group(1)), spill, int(sm.group(1)) if sm else 0)
    return info


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for line in sass.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = counts.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Za-z0-9_]+)*)", line)
        if not m or cur is None:
            continue
        op, mods = m.group(1), m.group(2)
        cur["total"] += 1
        if op in ("RED", "ATOM", "ATOMG", "ATOMS", "REDG", "REDS", "REDUX"):
            fam = "RED" if op.startswith("RED") and op != "REDUX" else "ATOM"
            if op == "REDUX":
                continue
            cur[fam + (".64" if ".64" in mods else "")] += 1
        elif op in ("UTMALDG", "LDGSTS", "SYNCS", "DFMA", "DADD", "DMUL", "MUFU", "F2I", "I2F", "FRND", "LDS", "STS", "LDG", "STG", "SHFL", "BAR"):
            cur[op] += 1
    names = list(counts)
    short = demangle(names)
    pt = ptxas_info()
    lines = ["# static SASS opcode counts per kernel of libminipic_b200.so (sm_100a) -- tools/sass_summary.py",
             "# " + subprocess.run(["nvcc", "--version"], capture_output=True, text=True).stdout.strip().split("\n")[-1],
             f"{'kernel':58s} {'regs':>4s} {'spill':>5s} {'smem':>6s} " + " ".join(f"{f:>7s}" for f in FAMILIES)]
    for n, s in sorted(zip(names, short), key=lambda t: t[1]):
        r = pt.get(n, ("?", "?", "?"))
        c = counts[n]
        lines.append(f"{s[:58]:58s} {r[0]!s:>4s} {r[1]!s:>5s} {r[2]!s:>6s} " + " ".join(f"{c.get(f, 0):7d}" for f in FAMILIES))
    tot = collections.Counter()
    for c in counts.values():
        tot.update(c)
    lines.append(f"{'ALL KERNELS (' + str(len(names)) + ')':58s} {'':>4s} {'':>5s} {'':>6s} " + " ".join(f"{tot.get(f, 0):7d}" for f in FAMILIES))
    lines.append("# no tcgen05 / UTCMMA / HMMA anywhere: the path has no contraction (DESIGN.md section 3); tensor-memory opcodes found: "
                 + str(len(re.findall(r"\b(UTCMMA|UTCHMMA|HMMA|IMMA|DMMA)\b", sass))))
    text = "\n".join(lines) + "\n"
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(text)
    else:
        sys.stdout.write(text)


if __name__ == "__main__":
    main()
