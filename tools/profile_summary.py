"""profiles/prof_<name>_summary.txt from one `ncu --set full --import-source on` capture:
  python tools/profile_summary.py gpurun_out/ncu_c2_r2.ncu-rep 'step_kernelIfLi128ELb0' > profiles/prof_step_c2_r2_summary.txt
Sections: launch / time / DRAM bytes, per-pipe executed-instruction table (sm__inst_executed_pipe_*), pipe activity, stall
reasons (source-page samples), hottest source lines, SASS opcode histogram + an excerpt of the shipped cubin
(cuobjdump -sass of libbsgcore.so, function name filter = argv[2])."""
import collections, csv, io, os, re, subprocess, sys
rep, fn = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else None)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2]
M = {h: (v, u) for h, u, v in zip(hdr, units, data)}
def show(pat, title):
    print(f"== {title}")
    for h in hdr:
        if re.search(pat, h):
            v, u = M[h]
            print(f"  {h:86s} {v:>18s} {u}")
show(r"^Kernel Name$|^Block Size$|^Grid Size$|gpu__time_duration.sum|launch__registers_per_thread$|launch__shared_mem_per_block$|"
     r"launch__occupancy_limit_(registers|shared_mem|warps)|sm__warps_active.avg.pct|dram__bytes_(read|write).sum$|sm__cycles_(active.avg|elapsed.max)$",
     "launch, time, DRAM traffic")
show(r"sm__inst_executed_pipe_[a-z0-9_]+\.sum$|sm__inst_executed_pipe_[a-z0-9_]+\.avg\.pct_of_peak_sustained_active$|smsp__inst_executed.sum$|sm__inst_executed.sum$|"
     r"smsp__thread_inst_executed_per_inst_executed.ratio|sass_thread_inst_executed_op_ffma_pred_on_x2|smsp__sass_thread_inst_executed_op_(fadd|fmul|ffma)_pred_on.sum$",
     "executed warp instructions per pipe (share of that pipe's issue peak while the SM is active; counts: tools/gpu_pipes.sh)")
show(r"sm__pipe_[a-z0-9_]+_cycles_active.avg.pct_of_peak_sustained_active|sm__issue_active.avg.pct|smsp__issue_active.avg.pct|sm__throughput.avg.pct|"
     r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared(_op_(ld|st|atom))?.sum$|dram__throughput.avg.pct", "pipe activity, issue, shared-memory conflicts")
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True, text=True).stdout
cur, h2, agg, tot = None, None, [], collections.Counter()
for r in csv.reader(io.StringIO(src)):
    if len(r) == 2 and r[0] == 'File Path':
        cur = r[1].split('/')[-1]; continue
    if r and r[0] == 'Line No':
        h2 = r; continue
    if h2 and len(r) == len(h2) and r[0].isdigit():
        d = dict(zip(h2, r))
        agg.append((int(d['Instructions Executed']), int(d['# Samples']), cur, r[0], r[1].strip()[:96]))
        for k, v in d.items():
            if k.startswith('stall_') and '(' not in k and v.isdigit():
                tot[k] += int(v)
ts = sum(tot.values()) or 1
print("== warp stall reasons (share of the source-page samples)")
print("  " + ", ".join(f"{k[6:]} {100 * v / ts:.1f}%" for k, v in tot.most_common(10)))
ti = sum(a[0] for a in agg) or 1; sm = sum(a[1] for a in agg) or 1
print(f"== hottest source lines (share of {ti} executed warp instructions | share of samples)")
for a in sorted(agg, reverse=True)[:24]:
    print(f"  {100 * a[0] / ti:5.1f}% | {100 * a[1] / sm:5.1f}%  {a[2]}:{a[3]:>4s}  {a[4]}")
if fn:
    so = os.path.join(ROOT, 'bluesky_gym_hallucination_b200', 'libbsgcore.so')
    sass = subprocess.run(['cuobjdump', '-sass', so], capture_output=True, text=True).stdout
    blocks = sass.split('Function : ')
    blk = next((b for b in blocks if fn in b.split('\n', 1)[0]), None)
    if blk:
        ops = collections.Counter()
        lines = []
        for l in blk.split('\n'):
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
            if m:
                ops[m.group(3).split('.')[0]] += 1; lines.append(l.split('/*', 2)[1].split('*/')[1].rstrip(' ;') if False else l[:100].rstrip())
        print(f"== SASS of the shipped cubin: {blk.split(chr(10), 1)[0][:110]}  ({sum(ops.values())} instructions)")
        print("  static opcode histogram: " + ", ".join(f"{k} {v}" for k, v in ops.most_common(28)))
        key = next((i for i, l in enumerate(lines) if 'FMNMX3' in l or 'ATOMS' in l or 'UBLKCP' in l), None)
        if key is not None:
            print("  excerpt around the first FMNMX3 / ATOMS / UBLKCP:")
            for l in lines[max(0, key - 28):key + 12]:
                print("   " + l)
