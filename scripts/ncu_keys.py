"""Print the handful of ncu raw-page metrics we read for the SpMM kernels: python scripts/ncu_keys.py report.ncu-rep"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.max',
        'launch__registers_per_thread', 'smsp__average_warps_issue_stalled', 'smsp__issue_active.avg.pct',
        'l1tex__m_xbar2l1tex_read_bytes.sum', 'sm__warps_active.avg.pct', 'lts__t_sector_hit_rate', 'l1tex__throughput.avg.pct',
        'lts__throughput.avg.pct', 'smsp__warps_issue_stalled', 'l1tex__m_l1tex2xbar_write_bytes.sum']
for r in rows[2:]:
    print("==", r[hdr.index('Kernel Name')] if 'Kernel Name' in hdr else '')
    for i, h in enumerate(hdr):
        if any(h.startswith(w) for w in want) and not h.endswith('.peak_sustained') and '.per_second' not in h and 'pct_of_peak_sustained_active' not in h:
            try:
                v = float(r[i].replace(',', ''))
            except ValueError:
                continue
            if 'stalled' in h and v < 0.3:
                continue
            print("  %-95s %-10s %s" % (h, units[i], r[i]))
