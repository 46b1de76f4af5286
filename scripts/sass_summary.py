"""SASS evidence for the library as built: per kernel, how many tensor-core / tensor-memory / TMA / wide-load instructions
the sm_100a machine code contains, with registers and spills from ptxas.

    python scripts/sass_summary.py [out.md]        (run by __graft_entry__.build(); default profiles/r02_sass_summary.md)

Sources: `cuobjdump -sass cl-waffe2_b200/lib/libwaffe2cuda.so` (mnemonics per `Function :` block) and the `-Xptxas -v`
logs build.py keeps in cl-waffe2_b200/_obj/*.log.  Mnemonics (B200_PROFILING.md): UTCHMMA = tcgen05.mma kind::tf32/f16, 2CTA = any instruction in its
cta_group::2 form (UTCHMMA.2CTA, UTMALDG...2CTA, UTCBAR.2CTA: the CTA-pair kernel),
LDTM / STTM = tcgen05.ld / st (tensor memory), UTMALDG / UTMASTG = cp.async.bulk.tensor load / store (TMA),
UBLKCP = cp.async.bulk (1-D bulk copy), UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, LDG.E.128 / STG.E.128 = 16-byte
global accesses, HMMA/IMMA would be legacy mma.sync (none expected)."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cl-waffe2_b200", "lib", "libwaffe2cuda.so")
OBJ = os.path.join(ROOT, "cl-waffe2_b200", "_obj")
CUDA_BIN = os.path.dirname(os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc"))
COLUMNS = ["UTCHMMA", "2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "LDG.E.128", "STG.E.128", "HMMA"]


def demangle(names):
    try:
        r = subprocess.run([os.path.join(CUDA_BIN, "cu++filt")], input="\n".join(names), capture_output=True, text=True, check=True)
        out = r.stdout.splitlines()
        if len(out) == len(names):
            return dict(zip(names, out))
    except (OSError, subprocess.CalledProcessError):
        pass
    return {n: n for n in names}


def short(name):
    """`void wf2::kernel<Args...>(params)` -> `kernel<Args...>` with long template lists elided."""
    name = re.sub(r"^void\s+", "", name)
    depth, cut = 0, len(name)
    for i, ch in enumerate(name):                      # drop the parameter list: the '(' at template depth 0
        depth += ch == "<"
        depth -= ch == ">"
        if ch == "(" and depth == 0:
            cut = i
            break
    name = name[:cut].replace("(anonymous namespace)::", "")
    return name if len(name) <= 110 else name[:107] + "..."


def ptxas_stats():
    """mangled entry -> (registers, spill stores, spill loads, smem bytes)."""
    stats = {}
    for f in sorted(os.listdir(OBJ)) if os.path.isdir(OBJ) else []:
        if not f.endswith(".log"):
            continue
        entry = None
        for line in open(os.path.join(OBJ, f), errors="replace"):
            m = re.search(r"Compiling entry function '([^']+)' for 'sm_100a'", line)
            if m:
                entry, spill = m.group(1), (0, 0)
                continue
            m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m and entry:
                spill = (int(m.group(2)), int(m.group(3)))
            m = re.search(r"Used (\d+) registers(?:, used \d+ barriers)?(?:, (\d+) bytes smem)?", line)
            if m and entry:
                stats[entry] = (int(m.group(1)), spill[0], spill[1], int(m.group(2) or 0))
                entry = None
    return stats


def sass_counts():
    r = subprocess.run([os.path.join(CUDA_BIN, "cuobjdump"), "-sass", LIB], capture_output=True, text=True, check=True)
    counts, cur = {}, None
    for line in r.stdout.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = dict.fromkeys(COLUMNS, 0)
            counts[cur]["n"] = 0
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        counts[cur]["n"] += 1
        if ".2CTA" in op:                                   # cta_group::2 forms (UTCHMMA.2CTA, UTMALDG...2CTA, UTCBAR.2CTA...)
            counts[cur]["2CTA"] += 1
        for col in COLUMNS:
            if op == col or op.startswith(col + ".") or (col in ("LDG.E.128", "STG.E.128") and op.startswith(col.split(".")[0] + ".E.") and ".128" in op):
                counts[cur][col] += 1
    return counts


def main(out_path):
    counts, stats = sass_counts(), ptxas_stats()
    names = demangle(list(counts))
    rows = []
    for mangled, c in counts.items():
        reg, ss, sl, smem = stats.get(mangled, (None, None, None, None))
        rows.append((short(names[mangled]), c, reg, ss, sl, smem))
    rows.sort(key=lambda r: (-(r[1]["UTCHMMA"] > 0), -(r[1]["UTMALDG"] + r[1]["UTMASTG"] + r[1]["UBLKCP"] > 0), r[0]))
    tc = [r for r in rows if r[1]["UTCHMMA"]]
    tma = [r for r in rows if not r[1]["UTCHMMA"] and (r[1]["UTMALDG"] or r[1]["UTMASTG"] or r[1]["UBLKCP"])]
    rest = [r for r in rows if r not in tc and r not in tma]
    lines = ["# SASS summary of `libwaffe2cuda.so` (sm_100a), written by `scripts/sass_summary.py` from `build()`", "",
             f"{len(rows)} kernels; {len(tc)} issue `tcgen05.mma` (UTCHMMA), {len(tma)} more use TMA / bulk copies, "
             f"{sum(1 for r in rows if r[1]['HMMA'])} use legacy `mma.sync` (HMMA).  "
             f"Kernels with register spills: {sum(1 for r in rows if (r[3] or 0) + (r[4] or 0) > 0)}.", "",
             "Columns: instruction counts in the machine code (static, not executed counts); `regs` / `spill st/ld` (bytes) / "
             "static `smem` from `ptxas -v`.", ""]

    def table(title, rs):
        if not rs:
            return
        lines.extend([f"## {title}", "", "| kernel | " + " | ".join(COLUMNS[:-1]) + " | SASS instr | regs | spill st/ld | smem |",
                      "|---|" + "---:|" * (len(COLUMNS) - 1 + 4)])
        for name, c, reg, ss, sl, smem in rs:
            lines.append(f"| `{name}` | " + " | ".join(str(c[k]) if c[k] else "" for k in COLUMNS[:-1]) +
                         f" | {c['n']} | {reg if reg is not None else '?'} | {ss or 0}/{sl or 0} | {smem or 0} |")
        lines.append("")

    table("Tensor-core kernels (tcgen05.mma, accumulators in tensor memory)", tc)
    table("Other kernels that use TMA / bulk copies", tma)
    table("Streaming / SIMT kernels", rest)
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    with open(out_path, "w") as f:
        f.write("\n".join(lines))
    return len(rows), len(tc)


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_summary.md")
    n, ntc = main(out)
    print(f"{out}: {n} kernels, {ntc} with UTCHMMA")
