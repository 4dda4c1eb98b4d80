"""One line per kernel launch of an .ncu-rep: time, DRAM bytes and rate, L1 wavefronts, issue/occupancy, top stalls.
python scripts/ncu_table.py report.ncu-rep [kernel-substring]   (needs ncu; no GPU)"""
import csv, io, subprocess, sys

out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
flt = sys.argv[2] if len(sys.argv) > 2 else ''
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]


def get(d, name):
    for k in hdr:
        if k == name or k.endswith('.' + name):
            try:
                return float(d[k])
            except ValueError:
                pass
    return float('nan')


print(f"{'kernel':42s} {'us':>7s} {'MBrd':>7s} {'MBwr':>7s} {'GB/s':>6s} {'dram%':>5s} {'l1%':>5s} {'lts%':>5s} {'wf_lg':>7s} {'wf_sh':>7s} {'cyc':>7s} "
      f"{'iss%':>5s} {'occ%':>5s} {'reg':>4s} {'fp64%':>5s}  stalls")
for r in rows[2:]:
    d = dict(zip(hdr, r))
    name = d['Kernel Name'].replace('void ', '').replace('hsfm::', '').replace('<unnamed>::', '').replace('unnamed>::', '')
    if flt not in name:
        continue
    name = name.split('(')[0][:42]
    t = get(d, 'gpu__time_duration.sum')
    unit = rows[1][hdr.index('gpu__time_duration.sum')]
    t_us = t * {'ms': 1e3, 'us': 1.0, 'ns': 1e-3, 's': 1e6}.get(unit, 1.0)

    def bytes_of(k):
        v = get(d, k)
        u = rows[1][hdr.index(k)] if k in hdr else 'byte'
        return v * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0}.get(u, 1.0)
    rd, wr = bytes_of('dram__bytes_read.sum'), bytes_of('dram__bytes_write.sum')
    stalls = []
    for k in hdr:
        if 'issue_stalled' in k and k.endswith('per_issue_active.ratio') and 'not_issued' not in k:
            try:
                v = float(d[k])
            except ValueError:
                continue
            nm = k.split('issue_stalled_')[1].replace('_per_issue_active.ratio', '')
            if nm not in ('selected',):
                stalls.append((v, nm))
    stalls.sort(reverse=True)
    st = ' '.join(f"{n}={v:.1f}" for v, n in stalls[:4])
    print(f"{name:42s} {t_us:7.1f} {rd / 1e6:7.1f} {wr / 1e6:7.1f} {(rd + wr) / t_us / 1e3:6.0f} {get(d, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):5.1f} "
          f"{get(d, 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed'):5.1f} {get(d, 'lts__throughput.avg.pct_of_peak_sustained_elapsed'):5.1f} "
          f"{get(d, 'l1tex__data_pipe_lsu_wavefronts_mem_lgds.avg') / 1e3:6.0f}k {get(d, 'l1tex__data_pipe_lsu_wavefronts_mem_shared.avg') / 1e3:6.0f}k "
          f"{get(d, 'sm__cycles_elapsed.avg') / 1e3:6.0f}k {get(d, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):5.1f} "
          f"{get(d, 'sm__warps_active.avg.pct_of_peak_sustained_active'):5.1f} {get(d, 'launch__registers_per_thread'):4.0f} "
          f"{get(d, 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active'):5.1f}  {st}")
