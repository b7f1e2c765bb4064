"""Summarise an `ncu --page raw --csv` dump: python tools/ncu_summary.py raw.csv [regex]"""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else
                 r"Kernel Name|Block Size|Grid Size|gpu__time_duration.sum|dram__bytes_(read|write).sum$|registers_per_thread|"
                 r"sm__warps_active.avg.pct|sm__throughput.avg.pct|smsp__inst_executed.sum$|inst_executed_pipe_(fma|alu|xu|fp64|lsu)[a-z_]*.sum$|"
                 r"pipe_(fma|alu|xu|fp64)[a-z_]*cycles_active.avg.pct|issue_active.avg.pct|warp_issue_stalled.*per_warp_active.pct$|"
                 r"thread_inst_executed_per_inst_executed.ratio|achieved_occupancy|occupancy_limit|shared_mem_per_block|l1tex__data_bank_conflicts.*shared.*sum$|"
                 r"sm__cycles_elapsed.max|sm__cycles_active.avg$|smsp__cycles_active.avg$|dram__throughput.avg.pct")
for d in data:
    for h, u, v in zip(hdr, units, d):
        if pat.search(h):
            print(f"{h.split('.', 2)[-1] if h.count('.') > 3 else h:90s} {v:>16s} {u}")
    print('-' * 60)
