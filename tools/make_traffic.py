"""profiles/traffic.json: DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of the hot
kernels from `ncu --set full` captures of tools/run_kernels.py (65,536 uniform images).

    python tools/make_traffic.py gpurun_out/a.ncu-rep [...]  > profiles/traffic.json
"""
import csv, io, json, subprocess, sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out = {"images": 65536, "profile": "uniform", "source": [p.split("/")[-1] for p in sys.argv[1:]], "kernels": {}}
for path in sys.argv[1:]:
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d, u = dict(zip(hdr, r)), dict(zip(hdr, units))
        name = d["Kernel Name"].split("(")[0].replace("void ", "").replace("yh::", "")
        tot = sum(float(d[m].replace(",", "")) * UNIT[u[m]] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        out["kernels"][name] = tot
print(json.dumps(out, indent=1))
