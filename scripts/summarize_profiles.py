#!/usr/bin/env python
"""Turn the ncu exports of scripts/gpu_round1_final.sh (gpurun_out/f_*) into the tracked summaries under profiles/."""
import collections, csv, json, os, shutil, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out"); P = os.path.join(ROOT, "profiles")
TAG = sys.argv[1] if len(sys.argv) > 1 else "r01b"


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


def raw_summary(path, title, out, note=""):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    def get(r, k, d=""):
        return r[ix[k]] if k in ix else d
    with open(out, "w") as f:
        f.write(f"# {title}\n\n`ncu --set full --clock-control none`, per launch; traffic = dram__bytes_read.sum + dram__bytes_write.sum.\n{note}\n")
        f.write("| kernel | grid | time us | dram read MB | dram write MB | traffic MB | dram % of peak | achieved occupancy % | regs | tensor pipe active % | L2 throughput % |\n|---|---|---|---|---|---|---|---|---|---|---|\n")
        for r in rows[2:]:
            def num(k):
                try: return float(get(r, k, "nan").replace(",", ""))
                except ValueError: return float("nan")
            def scaled(k):      # bytes metrics come with a unit column
                v = num(k); u = units[ix[k]] if k in ix else ""
                return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
            t = num("gpu__time_duration.sum"); tu = units[ix["gpu__time_duration.sum"]]
            t = t / 1000 if tu in ("ns", "nsecond") else (t * 1000 if tu in ("ms", "msecond") else t)
            rd, wr = scaled("dram__bytes_read.sum"), scaled("dram__bytes_write.sum")
            f.write(f"| `{get(r, 'Kernel Name')[:90]}` | {get(r, 'Grid Size')} | {t:.2f} | {rd:.1f} | {wr:.1f} | {rd + wr:.1f} | "
                    f"{num('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | {num('sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} | "
                    f"{get(r, 'launch__registers_per_thread')} | {num('TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed'):.1f} | {num('lts__throughput.avg.pct_of_peak_sustained_elapsed'):.1f} |\n")


os.makedirs(P, exist_ok=True)
sweep = launch_list(os.path.join(G, "f_launches.csv"))
table(sweep, "Round 1 (final) — ncu launch list of the C2 sweep",
      "ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv python bench.py --steps 3 --warmup 3 --no-cpu --mlp 0 --gemm 0 --gpt2 0",
      os.path.join(P, f"{TAG}_launches_summary.md"), f"Raw CSV: `profiles/{TAG}_launches.csv`.\n")
shutil.copy(os.path.join(G, "f_launches.csv"), os.path.join(P, f"{TAG}_launches.csv"))
mlp = launch_list(os.path.join(G, "f_mlp_launches.csv"))
table(mlp[-15:], "Round 1 (final) — ncu launch list of ONE compiled MNIST-MLP training step (C4, batch 1024)",
      "ncu --metrics gpu__time_duration.sum --clock-control none --csv python scripts/mlp_profile.py",
      os.path.join(P, f"{TAG}_mlp_launches_summary.md"),
      f"Last step of the run (15 launches, in issue order in `profiles/{TAG}_mlp_launches.csv`).  Inside the replayed graph the "
      "weight-gradient / bias-gradient kernels run on side branches beside the input-gradient GEMMs, so the step takes less than this sum.\n")
shutil.copy(os.path.join(G, "f_mlp_launches.csv"), os.path.join(P, f"{TAG}_mlp_launches.csv"))
raw_summary(os.path.join(G, "f_prof_raw.csv"), "Round 1 (final) — full capture of one C2 sweep step (12 launches)", os.path.join(P, f"{TAG}_ncu_full_sweep.md"))
raw_summary(os.path.join(G, "f_mlp_prof_raw.csv"), "Round 1 (final) — full capture of one MLP training step (15 launches)", os.path.join(P, f"{TAG}_ncu_full_mlp.md"))
raw_summary(os.path.join(G, "f_gemm_prof_raw.csv"), "Round 1 (final) — full capture of MatmulNode N=4096 (split planes x2 + tcgen05 kernel)", os.path.join(P, f"{TAG}_ncu_full_gemm.md"))
for src, dst in [("f_bench.json", f"{TAG}_bench_n1.json"), ("f_bench_ref.json", f"{TAG}_bench_reference_arm.json"), ("f_pytest.log", f"{TAG}_pytest_gpu.log")]:
    shutil.copy(os.path.join(G, src), os.path.join(P, dst))
print("ok")
