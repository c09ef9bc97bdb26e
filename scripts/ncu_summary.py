"""Summarise an .ncu-rep into markdown: per-kernel headline metrics (raw page) and the hottest SASS
lines by warp-stall samples (source page).  Usage: python scripts/ncu_summary.py rep.ncu-rep [out.md]"""
import csv
import io
import subprocess
import sys

RAW = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__registers_per_thread",
       "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
       "sm__inst_executed_pipe_tensor.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
       "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
       "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
       "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
       "smsp__cycles_active.avg", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
       "smsp__inst_executed.sum"]


def run(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    out = []
    raw = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = raw[0], raw[1]
    ki = hdr.index("Kernel Name")
    out.append(f"# ncu summary of `{rep}`\n")
    out.append("(captured with `ncu --set full --clock-control none --import-source on`; durations are cold-cache, "
               "serialised replays - compare shares, not absolutes)\n")
    for r in raw[2:]:
        out.append(f"## {r[ki][:110]}\n")
        out.append("| metric | value | unit |\n|---|---|---|")
        for m in RAW:
            if m in hdr:
                i = hdr.index(m)
                out.append(f"| {m} | {r[i]} | {units[i]} |")
        out.append("")
    src = run(["-i", rep, "--page", "source", "--csv"])
    blocks = src.split('"Kernel Name",')
    seen = set()
    for b in blocks[1:]:
        lines = b.splitlines()
        name = lines[0].strip('",')[:110]
        if name in seen:
            continue
        seen.add(name)
        rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
        if not rows:
            continue
        h = rows[0]
        try:
            cs, ca = h.index("Source"), h.index("Warp Stall Sampling (All Samples)")
        except ValueError:
            continue
        stall_cols = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
        body = [r for r in rows[1:] if len(r) > ca]
        tot = sum(float(r[ca] or 0) for r in body) or 1.0
        body.sort(key=lambda r: -float(r[ca] or 0))
        out.append(f"### hottest SASS by stall samples: {name}\n")
        out.append("| samples | share | top stall reasons | SASS |\n|---|---|---|---|")
        for r in body[:14]:
            reasons = sorted(((float(r[i] or 0), h[i]) for i in stall_cols), reverse=True)[:2]
            rs = ", ".join(f"{n}={int(v)}" for v, n in reasons if v > 0)
            out.append(f"| {r[ca]} | {100 * float(r[ca] or 0) / tot:.1f}% | {rs} | `{r[cs].strip()[:90]}` |")
        agg = {h[i]: sum(float(r[i] or 0) for r in body) for i in stall_cols}
        top = sorted(agg.items(), key=lambda kv: -kv[1])[:6]
        out.append("\nstall totals: " + ", ".join(f"{k}={int(v)}" for k, v in top) + "\n")
    text = "\n".join(out)
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(text)
    else:
        print(text)


if __name__ == "__main__":
    main()
