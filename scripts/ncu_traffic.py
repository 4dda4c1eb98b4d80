"""profiles/ncu_traffic.json from an `ncu --set full` capture: dram__bytes_read.sum + dram__bytes_write.sum per launch of
the kernels bench.py reports a roofline for.  python scripts/ncu_traffic.py <capture.ncu-rep> <config> [out.json]
(run where ncu is installed; no GPU needed).  bench.py reads the file for `roofline*.traffic`."""
import csv, io, json, os, subprocess, sys

GROUPS = {  # bench key -> kernel name fragments whose launches make up one pass (summed)
    "rj_form": ["rj_kernel<1, 1>("],
    "rj": ["rj_kernel<1, 0>("],
    "dense_schur": ["dense_schur_kernel<1>("],
    "schur_rhs": ["schur_pass_kernel<0>(", "cam_gather6_kernel", "cam_reduce6_kernel"],
    "schur_matvec": ["schur_pass_kernel<1>(", "cam_gather6_kernel", "cam_reduce6_kernel"],
    "backsub": ["schur_pass_kernel<2>("],
    "cost": ["cost_kernel("],
}


def main():
    rep, cfg = sys.argv[1], sys.argv[2]
    out_path = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json")
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = {}
    for r in rows[2:]:
        b = float(r[ir].replace(",", "")) * scale[units[ir]] + float(r[iw].replace(",", "")) * scale[units[iw]]
        per.setdefault(r[ik], []).append(b)
    res = {}
    for key, frags in GROUPS.items():
        tot, ok = 0.0, True
        for f in frags:
            hit = [v for k, v in per.items() if f in k]
            if not hit:
                ok = False
                break
            tot += sum(hit[0]) / len(hit[0])
        if ok:
            res[key] = tot
    data = json.load(open(out_path)) if os.path.exists(out_path) else {}
    data[cfg] = res
    data.setdefault("_source", {})[cfg] = os.path.basename(rep)
    json.dump(data, open(out_path, "w"), indent=1, sort_keys=True)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
