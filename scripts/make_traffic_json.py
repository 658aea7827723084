"""profiles/traffic.json from an ncu --metrics dram__bytes_{read,write}.sum capture of ONE launch of
the search kernel at the bench size: bench.py reports it as roofline.traffic (bytes per launch).
usage: python scripts/make_traffic_json.py gpurun_out/traffic_<tag>.csv vq_assign_tc_K2048"""
import csv, json, os, sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ik, im, iv, iu = (hdr.index(n) for n in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit"))
tot, info = 0.0, {}
for r in rows[1:]:
    v = float(r[iv].replace(",", ""))
    u = r[iu]
    if r[im].startswith("dram__bytes"):
        mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        tot += v * mul
        info[r[im]] = v * mul
    else:
        info[r[im] + " [" + u + "]"] = v
    info["kernel"] = r[ik][:80]
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
d = json.load(open(path)) if os.path.exists(path) else {}
d[sys.argv[2]] = tot
d[sys.argv[2] + "_detail"] = info
json.dump(d, open(path, "w"), indent=1, sort_keys=True)
print(sys.argv[2], tot, info)
