"""Summarise an .ncu-rep: key raw metrics per kernel + dynamic opcode mix / stall samples (source page)."""
import csv, collections, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__cycles_elapsed.avg', 'smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed',
        'smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed',
        'smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed',
        'launch__grid_size', 'launch__block_size', 'l1tex__t_bytes.sum', 'lts__t_bytes.sum',
        'smsp__inst_executed_pipe_alu.sum','smsp__inst_executed_pipe_fma.sum','smsp__inst_executed_pipe_fp64.sum','smsp__inst_executed_pipe_xu.sum','smsp__inst_executed_pipe_lsu.sum', 'smsp__inst_executed_pipe_cbu.sum','smsp__inst_executed_pipe_adu.sum','smsp__inst_executed_pipe_uniform.sum']
for r in rows[2:]:
    print('=====', r[hdr.index('Kernel Name')][:50])
    for k in keys:
        if k in hdr:
            print(f"  {k} = {r[hdr.index(k)]} {units[hdr.index(k)]}")
    stalls = {h.replace('smsp__pcsamp_warps_issue_stalled_', ''): int(r[i]) for i, h in enumerate(hdr)
              if h.startswith('smsp__pcsamp_warps_issue_stalled_') and not h.endswith('not_issued') and r[i].isdigit()}
    tot = sum(stalls.values())
    print("  stalls:", ", ".join(f"{k} {v/tot*100:.0f}%" for k, v in sorted(stalls.items(), key=lambda x: -x[1])[:7]))
for kern in sys.argv[2:]:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    h = rows[1]
    si, sa, ie = h.index('Source'), h.index('# Samples'), h.index('Instructions Executed')
    data = [(r[si].strip(), int(r[sa]), int(r[ie])) for r in rows[2:] if len(r) > ie and r[sa].isdigit()]
    n = len(data) // 2 if len(data) > 2 and data[0][0] == data[len(data)//2][0] else len(data)
    data = data[:n]
    ti, ts = sum(d[2] for d in data), sum(d[1] for d in data)
    byop = collections.defaultdict(lambda: [0, 0])
    for s_, sm, i in data:
        tok = s_.split()
        op = (tok[1] if tok[0].startswith('@') else tok[0]).split('.')[0]
        byop[op][0] += sm; byop[op][1] += i
    print('=====', kern, 'warp-instr', ti, 'sass lines', n)
    print("  " + ", ".join(f"{k} {v[1]/ti*100:.1f}%/{v[0]/ts*100:.0f}%s" for k, v in sorted(byop.items(), key=lambda x: -x[1][1])[:24]))
