"""Turns the ncu outputs of tools/refresh_profiles.sh into the committed summaries:
  profiles/<tag>_launches.csv (copied), profiles/<tag>_ncu_summary.md, profiles/ncu_traffic.json
Usage: python tools/summarize_profiles.py <tag> <launches.csv> <full.ncu-rep>"""
import collections, csv, json, os, shutil, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OURS = ("nnrt_b200::",)


def short(name):
    name = name.split("(")[0].strip()
    for prefix in ("void ", "nnrt_b200::"):
        name = name.replace(prefix, "")
    return name.replace("(int)", "")


def launch_table(path):
    rows = list(csv.reader(open(path)))
    start = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr = rows[start]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[start + 1:]:
        if len(r) <= mv:
            continue
        v = float(r[mv].replace(",", ""))
        v = v / 1e3 if r[mu] == "ns" else v * 1e3 if r[mu] == "ms" else v
        agg.setdefault(r[kn], []).append(v)
    return agg


def full_metrics(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    result = []
    for r in rows[2:]:
        def get(name, r=r):
            return (r[col[name]], units[col[name]]) if name in col else ("n/a", "")
        stalls = sorted(((float(r[i]), h.split("issue_stalled_")[1].replace("_per_issue_active.ratio", "")) for h, i in col.items()
                         if "smsp__average_warps_issue_stalled" in h and h.endswith("_per_issue_active.ratio") and r[i] not in ("", "n/a")), reverse=True)
        result.append(dict(name=r[col["Kernel Name"]], get=get, stalls=stalls[:5], grid=r[col["Grid Size"]], block=r[col["Block Size"]]))
    return result


def to_bytes(value, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    return float(value.replace(",", "")) * scale.get(unit, 1)


def main():
    tag, launches, rep = sys.argv[1:4]
    shutil.copy(launches, os.path.join(ROOT, "profiles", f"{tag}_launches.csv"))
    agg = launch_table(launches)
    ours = {k: v for k, v in agg.items() if any(p in k for p in OURS)}
    total = sum(sum(v) for v in ours.values())
    lines = [f"# {tag} — ncu launch list and per-kernel captures",
             "", f"Launch list: `profiles/{tag}_launches.csv` (`ncu --metrics gpu__time_duration.sum --clock-control none` over `python bench.py --steps 1 "
             "--warmup 3 --no-cpu-baseline --no-configs`, tools/refresh_profiles.sh). Times under ncu are cold-cache and serialised: only the SHARE per kernel is meaningful.",
             "", "| kernel | launches | total us | avg us | share of our kernels |", "|---|---|---|---|---|"]
    for k, v in sorted(ours.items(), key=lambda kv: -sum(kv[1])):
        lines.append(f"| `{short(k)}` | {len(v)} | {sum(v):.1f} | {sum(v) / len(v):.2f} | {100 * sum(v) / total:.1f} % |")
    other = sum(sum(v) for k, v in agg.items() if k not in ours)
    lines += ["", f"Library kernels in the same run (torch fills / copies of the bench harness): {other:.1f} us in total.", "",
              f"## `ncu --set full --clock-control none --import-source on`, one launch per kernel of the iteration (`gpurun_out/{os.path.basename(rep)}`)", ""]
    traffic = {}
    for m in full_metrics(rep):
        g = m["get"]
        rd, wr = g("dram__bytes_read.sum"), g("dram__bytes_write.sum")
        name = short(m["name"])
        traffic[name] = {"dram_bytes_per_launch": to_bytes(*rd) + to_bytes(*wr), "source": f"ncu --set full, profiles/{tag}_ncu_summary.md"}
        lines += [f"### `{name}`", "",
                  f"* time: {g('gpu__time_duration.sum')[0]} {g('gpu__time_duration.sum')[1]}; grid {m['grid']} x block {m['block']}; "
                  f"{g('launch__registers_per_thread')[0]} registers/thread",
                  f"* DRAM read {rd[0]} {rd[1]}, write {wr[0]} {wr[1]}; DRAM throughput {g('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')[0]} % of peak",
                  f"* issue slots busy (active cycles): {g('smsp__issue_active.avg.pct_of_peak_sustained_active')[0]} %; achieved occupancy "
                  f"{g('sm__warps_active.avg.pct_of_peak_sustained_active')[0]} %; executed warp instructions {g('smsp__inst_executed.sum')[0]}",
                  f"* warp latency per instruction: {g('smsp__average_warp_latency_per_inst_issued.ratio')[0]} cycles; shared-memory bank conflicts "
                  f"{g('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum')[0]}",
                  "* top stalls (warps per issue): " + ", ".join(f"{n} {v:.2f}" for v, n in m["stalls"]), ""]
    open(os.path.join(ROOT, "profiles", f"{tag}_ncu_summary.md"), "w").write("\n".join(lines))
    json.dump(traffic, open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
