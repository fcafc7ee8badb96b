#!/usr/bin/env python
"""Turn the .ncu-rep files of tools/prof_all.sh into profiles/<tag>_<kernel>_ncu.txt summaries and profiles/r2_traffic.json."""
import json, os, subprocess, sys
tag = sys.argv[1]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GO, PR = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
kernels = ["logmel_kernel", "decode_scan_kernel", "decode_tail_kernel", "no_speech_kernel", "kv_reorder_kernel", "proj_kernel"]
mangled = {"logmel_kernel": "logmel_kernelILb1", "decode_scan_kernel": "decode_scan_kernel", "decode_tail_kernel": "decode_tail_kernelILi0",
           "no_speech_kernel": "no_speech_kernel", "kv_reorder_kernel": "kv_reorder_kernel", "proj_kernel": "proj_kernelILi6E"}
traffic = {"source": f"ncu --set full --clock-control none, one launch per kernel of `python bench.py --no-cpu-baseline --only decode,kv,project --steps 8 "
                     f"--warmup 3` (profiles/{tag}_<kernel>_ncu.txt): dram__bytes_read.sum + dram__bytes_write.sum", "kernels": {}}
def num(v, unit):
    x = float(v.replace(",", ""))
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
for k in kernels:
    rep = os.path.join(GO, f"{tag}_{k}.ncu-rep")
    if not os.path.exists(rep):
        print("missing", rep); continue
    raw = os.path.join(GO, f"{tag}_{k}_raw.csv")
    sass = os.path.join(GO, f"{tag}_{k}_sass.csv")
    subprocess.run(f"ncu -i {rep} --page raw --csv > {raw} 2>/dev/null", shell=True)
    subprocess.run(f"ncu -i {rep} --page source --csv --print-source sass > {sass} 2>/dev/null", shell=True)
    keys = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_keys.py"), raw], capture_output=True, text=True).stdout
    lines = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), sass, os.path.join(ROOT, "whisper_rs_b200", "lib", "libwrs_b200.so"),
                            mangled[k], "40"], capture_output=True, text=True).stdout
    with open(os.path.join(PR, f"{tag}_{k}_ncu.txt"), "w") as f:
        f.write(keys.replace("----\n", "") + "\ntop source lines by warp-stall samples (tools/ncu_lines.py):\n" + lines)
    import csv
    rows = list(csv.reader(open(raw)))
    hdr, units, r = rows[0], rows[1], rows[2]
    rd, wr = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    traffic["kernels"][k] = {"dram_bytes": num(r[rd], units[rd]) + num(r[wr], units[wr]),
                             "dram_bytes_read": num(r[rd], units[rd]), "dram_bytes_write": num(r[wr], units[wr]),
                             "duration_us_under_ncu": float(r[hdr.index("gpu__time_duration.sum")].replace(",", ""))}
    print(k, traffic["kernels"][k])
for f in (f"{tag}_launches.csv", f"{tag}_plain.json"):
    src = os.path.join(GO, f)
    if os.path.exists(src):
        dst = os.path.join(PR, f.replace("_plain.json", "_bench_for_profiles.json"))
        txt = open(src).read()
        if f.endswith(".csv"):
            txt = "\n".join(l for l in txt.splitlines() if not l.startswith("==PROF=="))
        open(dst, "w").write(txt)
json.dump(traffic, open(os.path.join(PR, "r2_traffic.json"), "w"), indent=1)
print("wrote profiles/r2_traffic.json")
