"""Builds the committed evidence under profiles/ and sass/ from the built library and from whatever
ncu artefacts gpurun brought back into gpurun_out/ (scratch, not committed)."""
import collections
import csv
import glob
import gzip
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "build", "obj")
PROF = os.path.join(ROOT, "profiles")
SASS = os.path.join(ROOT, "sass")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UTMACCTL", "UBLKCP",
        "SYNCS", "LDGMC", "STG.SYS", "RED", "ATOMG", "HMMA", "UTCATOMSWS", "UCGABAR"]


def sass_evidence():
    os.makedirs(SASS, exist_ok=True)
    lines = ["# SASS evidence (cuobjdump -sass of build/obj/*.o, sm_100a)", "",
             "Counts of the instructions that identify a Blackwell-native kernel (B200_PROFILING.md): `UTC*MMA` = "
             "`tcgen05.mma`, `LDTM`/`STTM` = `tcgen05.ld/st`, `UTMALDG`/`UTMASTG` = TMA tensor copies, `SYNCS` = mbarrier, "
             "`LDGMC` = `multimem.ld_reduce`, `RED`/`ATOMG` = global reductions/atomics, `UCGABAR` = cluster barrier. "
             "`HMMA` (legacy `mma.sync`) must be absent. `multimem.st` has no mnemonic of its own: it is an ordinary "
             "`STG.E.{64,128}.STRONG.SYS` whose address lies in the multicast mapping (column `STG.SYS`), next to the "
             "system-scope flag stores. Full listings: `sass/<file>.sass.gz`.", ""]
    for obj in sorted(glob.glob(os.path.join(OBJ, "*.o"))):
        name = os.path.basename(obj)[:-2]
        txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
        with gzip.open(os.path.join(SASS, name + ".sass.gz"), "wt") as f:
            f.write(txt)
        lines += [f"## {name}.cu", "", "| kernel | " + " | ".join(KEYS) + " | instrs |", "|---|" + "---|" * (len(KEYS) + 1)]
        cur, counts, n = None, None, 0
        def flush():
            if cur is not None:
                lines.append(f"| `{cur}` | " + " | ".join(str(counts.get(k, 0)) for k in KEYS) + f" | {n} |")
        for ln in txt.splitlines():
            m = re.search(r"Function : (\S+)", ln)
            if m:
                flush()
                cur = subprocess.run(["c++filt", "-p", m.group(1)], capture_output=True, text=True).stdout.strip()[:90]
                counts, n = collections.Counter(), 0
                continue
            m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
            if m and cur is not None:
                n += 1
                op = m.group(1)
                for k in KEYS:
                    if op.startswith(k):
                        counts[k] += 1
                if op.startswith("STG") and op.endswith("STRONG.SYS"):
                    counts["STG.SYS"] += 1
        flush()
        lines.append("")
    open(os.path.join(PROF, "sass_summary.md"), "w").write("\n".join(lines))


def launch_lists():
    for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "launches_*.csv"))):
        rows = list(csv.DictReader([l for l in open(path) if not l.startswith("==")]))
        agg, tot = collections.defaultdict(lambda: [0, 0.0]), 0.0
        for r in rows:
            nm = re.sub(r"\(.*", "", r["Kernel Name"])[:80]
            v = float(r["Metric Value"]) / 1000.0
            agg[nm][0] += 1
            agg[nm][1] += v
            tot += v
        out = [f"# {os.path.basename(path)}: per-kernel device time (ncu gpu__time_duration.sum, --clock-control none)",
               "# cold-cache, serialised launches: compare shares, not absolutes", "kernel,launches,total_us,share_pct,avg_us"]
        for nm, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            out.append(f"\"{nm}\",{c},{t:.1f},{100 * t / tot:.1f},{t / c:.1f}")
        open(os.path.join(PROF, os.path.basename(path)), "w").write("\n".join(out) + "\n")


def ncu_reports():
    want = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed", "launch__grid_size", "launch__block_size",
            "launch__cluster_size", "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
    for rep in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "*.ncu-rep"))):
        r = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True)
        rows = list(csv.reader(r.stdout.splitlines()))
        if len(rows) < 3:
            continue
        hdr = rows[0]
        out = [f"# {os.path.basename(rep)} (ncu --set full --clock-control none --import-source on), selected raw metrics",
               "metric,unit," + ",".join(f"launch{i}" for i in range(len(rows) - 2))]
        kcol = hdr.index("Kernel Name") if "Kernel Name" in hdr else None
        if kcol is not None:
            out.append("kernel,," + ",".join('"' + rows[i][kcol][:100] + '"' for i in range(2, len(rows))))
        for i, h in enumerate(hdr):
            if h in want:
                out.append(f"{h},{rows[1][i]}," + ",".join(rows[j][i] for j in range(2, len(rows))))
        open(os.path.join(PROF, os.path.basename(rep).replace(".ncu-rep", ".csv")), "w").write("\n".join(out) + "\n")


if __name__ == "__main__":
    os.makedirs(PROF, exist_ok=True)
    sass_evidence()
    launch_lists()
    ncu_reports()
    print(sorted(os.listdir(PROF)))
