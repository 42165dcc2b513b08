#!/usr/bin/env python
"""Turn what a gpurun call brought back (gpurun_out/) into the tracked summaries under profiles/."""
import csv, json, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out"); P = os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"

def short(n):
    m = re.search(r"(\w+_kernel)", n)
    return m.group(1) if m else n[:40]

# 1. launch list (ncu --metrics gpu__time_duration.sum)
lp = os.path.join(G, "launches.csv")
if os.path.exists(lp):
    rows = list(csv.DictReader([l for l in open(lp) if not l.startswith("==")]))
    with open(os.path.join(P, f"launches_{tag}.md"), "w") as f:
        f.write("# Kernel launch list of training steps (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised: compare shares)\n\n")
        f.write("| # | kernel | grid | block | ns |\n|---|---|---|---|---|\n")
        for r in rows[-30:]:
            f.write(f"| {r['ID']} | {short(r['Kernel Name'])} | {r['Grid Size']} | {r['Block Size']} | {r['Metric Value']} |\n")

# 2. full-set metrics per kernel
rp = os.path.join(G, "prof_step.ncu-rep")
if os.path.exists(rp):
    raw = subprocess.run(["ncu", "-i", rp, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
            "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor", "l1tex__m_xbar2l1tex_read_bytes.sum"]
    kn = hdr.index("Kernel Name")
    with open(os.path.join(P, f"ncu_full_{tag}.md"), "w") as f:
        f.write("# ncu --set full --clock-control none, one eager training step (B=128, 1 GPU)\n\n| kernel | " + " | ".join(w.split(".")[0].replace("launch__", "") for w in want) + " |\n")
        f.write("|---|" + "---|" * len(want) + "\n")
        for d in data:
            f.write(f"| {short(d[kn])} | " + " | ".join((d[hdr.index(w)] + " " + units[hdr.index(w)]) if w in hdr else "-" for w in want) + " |\n")

# 3. SASS evidence
so = os.path.join(ROOT, "b200dp", "lib", "libb200dp.so")
if os.path.exists(so):
    sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    cur, counts = None, {}
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = short(m.group(1)); counts.setdefault(cur, {}); continue
        m = re.search(r"\s(UTCHMMA|UTCQMMA|UTMALDG\.\dD|UTMASTG\.\dD|LDTM\S*|STTM\S*|UTCBAR|UTCATOMSWS\S*|UBLKRED\S*|UBLKCP\S*|LDGMC\S*|STGMC\S*|REDGMC\S*|SYNCS\.\S+|RED\.E\.ADD\.F32\S*|HMMA\S*|LDG\.E\.128\S*|ST\.E\.STRONG\.SYS|LD\.E\.STRONG\.SYS|STG\.E\.STRONG\.SYS|LDG\.E\.STRONG\.SYS)", line)
        if m and cur:
            k = m.group(1).split(".")[0] + ("." + ".".join(m.group(1).split(".")[1:3]) if "." in m.group(1) else "")
            counts[cur][k] = counts[cur].get(k, 0) + 1
    with open(os.path.join(P, f"sass_evidence_{tag}.md"), "w") as f:
        f.write("# SASS mnemonics per kernel (cuobjdump -sass b200dp/lib/libb200dp.so)\n\nUTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA tensor load, SYNCS = mbarrier, *.STRONG.SYS = system-scope peer flag traffic\n\n")
        for k, v in sorted(counts.items()):
            if v:
                f.write(f"- `{k}`: " + ", ".join(f"{a}×{b}" for a, b in sorted(v.items())) + "\n")

# 4. bench lines + kernel bench
with open(os.path.join(P, f"bench_{tag}.md"), "w") as f:
    f.write("# bench.py lines returned by gpurun\n\n")
    for fn in sorted(os.listdir(G)):
        if fn.startswith("bench") and fn.endswith(".log"):
            lines = [l for l in open(os.path.join(G, fn)) if l.startswith("{")]
            if lines:
                f.write(f"## {fn}\n```json\n{lines[-1].strip()}\n```\n")
    kb = os.path.join(G, "kernel_bench.json")
    if os.path.exists(kb):
        d = json.load(open(kb))
        f.write(f"\n## tools/kernel_bench.py (B={d['batch']}, graph-captured back-to-back launches, warm L2)\n\n| kernel | us/launch |\n|---|---|\n")
        for k, v in d["us_per_launch"].items():
            f.write(f"| {k} | {v:.2f} |\n")
print("profiles written:", sorted(os.listdir(P)))
