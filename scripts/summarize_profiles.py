#!/usr/bin/env python
"""Turn what `scripts/gpu.sh <tag> ...` left in gpurun_out/ into the tracked summaries under profiles/.

    python scripts/summarize_profiles.py <tag> [--traffic key=<rep basename>:<kernel substring> ...]

For every gpurun_out/<tag>_*.ncu-rep : profiles/<tag>_ncu_<name>.md   (one row per captured launch: time, DRAM bytes, tensor
                                        pipe, shared-memory pipe, occupancy, registers; read with `ncu -i ... --page raw --csv`)
For gpurun_out/<tag>_launches.csv    : profiles/<tag>_launches.csv + <tag>_launches_summary.md (per-kernel shares)
For <tag>_bench.json / _pytest.log   : copied as profiles/<tag>_bench_n1.json / <tag>_pytest_gpu.log
--traffic writes dram read + write bytes per launch of the named kernel into profiles/ncu_traffic.json (bench.py's
`roofline.traffic`)."""
import collections, csv, glob, io, json, os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out"); P = os.path.join(ROOT, "profiles")


def launch_list(path):
    rows = []
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    for row in csv.DictReader(lines):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            t = float(row["Metric Value"]); t = t / 1000 if row["Metric Unit"] in ("ns", "nsecond") else t
            rows.append((int(row["ID"]), row["Kernel Name"], row["Grid Size"], row["Block Size"], t))
    return rows


def table(rows, title, cmd, out, note=""):
    agg = collections.OrderedDict()
    for _, k, g, b, t in rows:
        a = agg.setdefault((k, g, b), [0, 0.0]); a[0] += 1; a[1] += t
    tot = sum(v[1] for v in agg.values())
    with open(out, "w") as f:
        f.write(f"# {title}\n\nCommand: `{cmd}` (cold-cache, serialised launches: compare SHARES, not absolutes).\n{note}\n")
        f.write("| kernel | grid | block | launches | mean us | share of listed time |\n|---|---|---|---|---|---|\n")
        for (k, g, b), (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k[:100]}` | {g} | {b} | {c} | {t / c:.2f} | {100 * t / tot:.1f}% |\n")
        f.write(f"\nTotal listed: {tot:.1f} us over {len(rows)} launches.\n")


COLS = [("time us", "gpu__time_duration.sum", "time"),
        ("dram read MB", "dram__bytes_read.sum", "bytes"), ("dram write MB", "dram__bytes_write.sum", "bytes"),
        ("dram % of peak", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "num"),
        ("tensor pipe % (while SM active)", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "num"),
        ("tensor pipe % (of elapsed)", "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "num"),
        ("shared-memory pipe % (LSU wavefronts)", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "num"),
        ("L2 throughput %", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "num"),
        ("SM busy cycles / elapsed", None, "busy"),
        ("occupancy %", "sm__warps_active.avg.pct_of_peak_sustained_active", "num"),
        ("regs", "launch__registers_per_thread", "str")]


def raw_rows(rep):
    r = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True)
    rows = list(csv.reader(io.StringIO(r.stdout)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    out = []
    for r_ in rows[2:]:
        def num(k):
            try: return float(r_[ix[k]].replace(",", ""))
            except (KeyError, ValueError): return float("nan")
        d = {"kernel": r_[ix["Kernel Name"]], "grid": r_[ix["Grid Size"]]}
        for label, key, kind in COLS:
            if kind == "time":
                t, u = num(key), units[ix[key]]
                d[label] = t / 1000 if u in ("ns", "nsecond") else (t * 1000 if u in ("ms", "msecond") else t)
            elif kind == "bytes":
                d[label] = num(key) * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(units[ix[key]], 1e-6)
            elif kind == "num":
                d[label] = num(key)
            elif kind == "busy":
                d[label] = num("smsp__cycles_active.avg") / num("sm__cycles_elapsed.max")
            else:
                d[label] = r_[ix[key]] if key in ix else ""
        out.append(d)
    return out


def ncu_summary(rep, out, title):
    rows = raw_rows(rep)
    with open(out, "w") as f:
        f.write(f"# {title}\n\n`ncu --set full --clock-control none --import-source on`, one row per captured launch (cold-cache, serialised; "
                f"the report itself is `gpurun_out/{os.path.basename(rep)}`, not tracked).  Traffic = dram__bytes_read.sum + dram__bytes_write.sum.\n\n")
        f.write("| kernel | grid | " + " | ".join(c[0] for c in COLS) + " |\n|---|---|" + "---:|" * len(COLS) + "\n")
        for d in rows:
            cells = []
            for label, _, kind in COLS:
                v = d[label]
                cells.append(v if kind == "str" else (f"{v:.2f}" if kind in ("time", "busy") else f"{v:.1f}"))
            f.write(f"| `{d['kernel'][:90]}` | {d['grid']} | " + " | ".join(cells) + " |\n")
    return rows


def main():
    tag = sys.argv[1]
    traffic = [a.split("=", 1) for a in sys.argv[3:]] if len(sys.argv) > 2 and sys.argv[2] == "--traffic" else []
    os.makedirs(P, exist_ok=True)
    captured = {}
    for rep in sorted(glob.glob(os.path.join(G, f"{tag}_*.ncu-rep"))):
        name = os.path.basename(rep)[len(tag) + 1:-len(".ncu-rep")]
        captured[name] = ncu_summary(rep, os.path.join(P, f"{tag}_ncu_{name}.md"), f"{tag}: ncu --set full capture while running `{name}`")
        print("wrote", f"profiles/{tag}_ncu_{name}.md", len(captured[name]), "launches")
    ll = os.path.join(G, f"{tag}_launches.csv")
    if os.path.exists(ll):
        shutil.copy(ll, os.path.join(P, f"{tag}_launches.csv"))
        table(launch_list(ll), f"{tag}: ncu launch list of the element-wise / reduction sweep and the fused chains",
              "ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv python bench.py --steps 2 --warmup 1 --gemm 0 --gpt2 0 --cnn 0 --mlp 0 --no-cpu",
              os.path.join(P, f"{tag}_launches_summary.md"), f"Raw CSV: `profiles/{tag}_launches.csv`.\n")
        print("wrote", f"profiles/{tag}_launches_summary.md")
    for src, dst in [(f"{tag}_bench.json", f"{tag}_bench_n1.json"), (f"{tag}_bench_ref.json", f"{tag}_bench_reference_arm.json"),
                     (f"{tag}_pytest.log", f"{tag}_pytest_gpu.log")]:
        if os.path.exists(os.path.join(G, src)):
            shutil.copy(os.path.join(G, src), os.path.join(P, dst))
    if traffic:
        path = os.path.join(P, "ncu_traffic.json")
        try:
            tj = json.load(open(path))
        except (OSError, ValueError):
            tj = {}
        for key, spec in traffic:
            name, sub = spec.split(":", 1)
            hits = [d for d in captured.get(name, []) if sub in d["kernel"]]
            if hits:
                tj[key] = int(1e6 * sum(d["dram read MB"] + d["dram write MB"] for d in hits) / len(hits))
        tj["_source"] = (f"scripts/summarize_profiles.py {tag} --traffic ...: dram__bytes_read.sum + dram__bytes_write.sum per launch (bytes), "
                         f"mean over the captured launches; tables in profiles/{tag}_ncu_*.md")
        json.dump(tj, open(path, "w"), indent=2)
        print("wrote profiles/ncu_traffic.json", {k: v for k, v in tj.items() if not k.startswith("_")})


if __name__ == "__main__":
    main()
