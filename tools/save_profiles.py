#!/usr/bin/env python
"""tools/save_profiles.py TAG: copy what tools/round.sh TAG left under gpurun_out/TAG into profiles/ (bench lines,
launch list, ncu summaries of the two --set full captures) and refresh profiles/traffic.json from the summaries."""
import csv, json, os, re, shutil, subprocess, sys
tag = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src, dst = os.path.join(root, "gpurun_out", tag), os.path.join(root, "profiles")
for w in ("c1", "c2", "c3", "c2_3d", "c2_reference"):
    f = os.path.join(src, f"bench_{w}.json")
    if os.path.exists(f):
        shutil.copy(f, os.path.join(dst, f"r1_bench_{tag}_{w}.json"))
if os.path.exists(os.path.join(src, "launches_c2.csv")):
    shutil.copy(os.path.join(src, "launches_c2.csv"), os.path.join(dst, f"r1_launches_{tag}.csv"))
traffic = {}
for w in ("c2", "c3"):
    rep = os.path.join(src, f"march_{w}.ncu-rep")
    if not os.path.exists(rep):
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    kern = rows[2][rows[0].index("Kernel Name")]
    pick = subprocess.run([sys.executable, os.path.join(root, "tools", "ncu_pick.py"), rep], capture_output=True, text=True).stdout
    keep = [l for l in pick.splitlines() if "Triage" not in l and ".max." not in l and ".min." not in l]
    out = os.path.join(dst, f"r1_march_{tag}_{w}_ncu.txt")
    with open(out, "w") as fh:
        fh.write(f"# ncu --set full --clock-control none, one k_march launch of: bench.py --workload {w} (round {tag}); kernel:\n# {kern}\n")
        fh.write("\n".join(keep) + "\n")
    txt = "\n".join(keep)
    def grab(key):
        m = re.search(re.escape(key) + r"\s+([\d.,]+)\s+(\S+)", txt)
        return float(m.group(1).replace(",", "")) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}.get(m.group(2), 1)
    traffic[w] = {"dram_bytes_per_launch": grab("dram__bytes_read.sum ") + grab("dram__bytes_write.sum "),
                  "source": f"profiles/r1_march_{tag}_{w}_ncu.txt (ncu --set full, one k_march launch, orbit frame 1)"}
if traffic:
    json.dump(traffic, open(os.path.join(dst, "traffic.json"), "w"), indent=1)
print(json.dumps(traffic, indent=1))
