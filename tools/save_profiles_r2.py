#!/usr/bin/env python
"""tools/save_profiles_r2.py: copy the round-2 evidence from gpurun_out/ (scratch) into profiles/ (tracked):
experiment texts, bench lines, launch lists, and text summaries of the ncu --set full captures (the .ncu-rep files stay
in gpurun_out/).  Refreshes profiles/traffic.json with the DRAM bytes per launch of the captured kernel version."""
import csv, glob, json, os, re, shutil, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
go, dst = os.path.join(root, "gpurun_out"), os.path.join(root, "profiles")
copied = []


def cp(src, name):
    if os.path.exists(src) and os.path.getsize(src) > 0:
        shutil.copy(src, os.path.join(dst, name)); copied.append(name)


for f in glob.glob(os.path.join(go, "r2_evidence", "exp_*.txt")):
    cp(f, "r2_" + os.path.basename(f))
cp(os.path.join(go, "r2_evidence", "bench_c2.json"), "r2_bench_c2_dirty_clear_on.json")
cp(os.path.join(go, "r2_evidence", "bench_c2_fullclear.json"), "r2_bench_c2_dirty_clear_off.json")
for tag in sorted(glob.glob(os.path.join(go, "r2_mg*"))):
    t = os.path.basename(tag)
    for f in glob.glob(os.path.join(tag, "bench_n*.json")) + glob.glob(os.path.join(tag, "exp_d2h_n*.txt")) + glob.glob(os.path.join(tag, "topo_n*.txt")) + glob.glob(os.path.join(tag, "pytest_multi_n*.log")):
        cp(f, f"{t}_{os.path.basename(f)}")
for tag in sorted(glob.glob(os.path.join(go, "r2_sl*"))):
    for f in glob.glob(os.path.join(tag, "sl_n*.json")):
        cp(f, f"r2_sortlast_{os.path.basename(f)}")
fin = os.path.join(go, "r2_final")
for w in ("c1", "c2", "c3", "c2_reference"):
    cp(os.path.join(fin, f"bench_{w}.json"), f"r2_bench_final_{w}.json")
for f in ("pytest_gpu.log", "smoke.log", "launches_c2.csv", "launches_c3.csv"):
    cp(os.path.join(fin, f), "r2_" + f)
traffic_path = os.path.join(dst, "traffic.json")
traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
for name, key in (("march_c2", "c2"), ("march_c3", "c3"), ("texstream", None), ("tile_mask_c3", None)):
    rep = os.path.join(fin, name + ".ncu-rep")
    if not os.path.exists(rep):
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    kern = rows[2][rows[0].index("Kernel Name")]
    pick = subprocess.run([sys.executable, os.path.join(root, "tools", "ncu_pick.py"), rep], capture_output=True, text=True).stdout
    keep = [l for l in pick.splitlines() if "Triage" not in l and ".max." not in l and ".min." not in l and "surface_op" not in l]
    lines = subprocess.run([sys.executable, os.path.join(root, "tools", "ncu_lines.py"), rep, "40"], capture_output=True, text=True).stdout
    version = open(os.path.join(fin, "kernel_version.txt")).read().strip() if os.path.exists(os.path.join(fin, "kernel_version.txt")) else "?"
    out = f"r2_{name}_ncu.txt"
    with open(os.path.join(dst, out), "w") as fh:
        fh.write(f"# ncu --set full --clock-control none --import-source on, one launch (tools/r2_final.sh); kernel version {version}; kernel:\n# {kern}\n")
        fh.write("\n".join(keep) + "\n\n# instructions executed / stall samples per source line (tools/ncu_lines.py)\n" + lines)
    copied.append(out)
    if key:
        txt = "\n".join(keep)

        def grab(k):
            m = re.search(re.escape(k) + r"\s+([\d.,]+)\s+(\S+)", txt)
            return float(m.group(1).replace(",", "")) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}.get(m.group(2), 1)
        traffic[key] = {"dram_bytes_per_launch": grab("dram__bytes_read.sum ") + grab("dram__bytes_write.sum "), "kernel_version": version,
                        "source": f"profiles/{out} (ncu --set full, one k_march launch of bench.py --workload {key})"}
json.dump(traffic, open(traffic_path, "w"), indent=1)
print("\n".join(copied)); print(json.dumps(traffic, indent=1))
