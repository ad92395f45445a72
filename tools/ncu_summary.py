"""Summarise an `ncu --set full` capture (.ncu-rep) into the small tracked files under profiles/.

    python tools/ncu_summary.py gpurun_out/prof_X.ncu-rep profiles/X [--top 40] [--all-sass]

Writes  <out>_summary.json   per-launch key metrics (duration, DRAM bytes, instruction mix, issue / FMA-pipe
                             utilisation, occupancy limits, warp stall breakdown)
        <out>_hot_sass.csv   the --top instructions by executed count with their source line and stall samples
Needs the `ncu` CLI (reads the report only; no GPU).
"""
import csv
import io
import json
import re
import subprocess
import sys

KEYS = [
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum',
    'smsp__inst_executed.sum', 'smsp__thread_inst_executed.sum', 'sm__cycles_elapsed.max', 'smsp__cycles_active.avg',
    'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed.avg.per_cycle_active',
    'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_xu.sum',
    'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_fmaheavy.sum', 'sm__inst_executed_pipe_fmalite.sum',
    'smsp__sass_thread_inst_executed_op_ffma_pred_on.sum', 'smsp__sass_thread_inst_executed_op_fadd_pred_on.sum',
    'smsp__sass_thread_inst_executed_op_fmul_pred_on.sum', 'smsp__sass_thread_inst_executed_op_fp32_pred_on.sum',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__maximum_warps_per_active_cycle_pct',
    'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
    'launch__occupancy_limit_warps', 'launch__waves_per_multiprocessor', 'launch__shared_mem_per_block_dynamic',
    'launch__grid_size', 'launch__block_size', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed_op_shared_ld.sum',
    'smsp__inst_executed_op_shared_st.sum',
]


def page(rep, name, extra=()):
    out = subprocess.run(['ncu', '-i', rep, '--page', name, '--csv', *extra], capture_output=True, text=True, check=True).stdout
    start = out.find('"')
    return list(csv.reader(io.StringIO(out[start:])))


def num(x):
    try:
        return float(x.replace(',', ''))
    except ValueError:
        return x


def main():
    rep, out = sys.argv[1], sys.argv[2]
    top = int(sys.argv[sys.argv.index('--top') + 1]) if '--top' in sys.argv else 40
    rows = page(rep, 'raw')
    hdr, units = rows[0], rows[1]
    launches = []
    for r in rows[2:]:
        d = {'kernel': r[hdr.index('Kernel Name')], 'grid': r[hdr.index('Grid Size')], 'block': r[hdr.index('Block Size')]}
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                d[k] = {'value': num(r[i]), 'unit': units[i]}
        stalls = {}
        for i, h in enumerate(hdr):
            if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio') or \
               h.startswith('smsp__average_warp_latency_issue_stalled_') and h.endswith('.ratio'):
                v = num(r[i])
                if isinstance(v, float) and v > 0.005:
                    stalls[h.replace('smsp__average_warps_issue_stalled_', '').replace('smsp__average_warp_latency_issue_stalled_', '')
                            .replace('_per_issue_active.ratio', '').replace('.ratio', '')] = round(v, 4)
        d['warp_stalls_per_issue'] = dict(sorted(stalls.items(), key=lambda kv: -kv[1]))
        rd, wr = d.get('dram__bytes_read.sum'), d.get('dram__bytes_write.sum')
        if rd and wr:
            scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
            d['dram_bytes'] = rd['value'] * scale.get(rd['unit'], 1) + wr['value'] * scale.get(wr['unit'], 1)
        launches.append(d)
    summary = {'report': rep.split('/')[-1], 'launches': launches,
               'dram_bytes_per_launch': (sum(l.get('dram_bytes', 0) for l in launches) / len(launches)) if launches else None}
    with open(out + '_summary.json', 'w') as f:
        json.dump(summary, f, indent=1)
    # hot instructions
    try:
        src = page(rep, 'source', ('--print-source', 'sass'))
    except subprocess.CalledProcessError:
        src = []
    if src:
        # sections: ["Kernel Name", name] / ["Address", "Source", ...] / instruction rows ; keep the first launch only
        h, body = None, []
        for r in src:
            if not r:
                continue
            if r[0] == 'Kernel Name':
                if h is not None:
                    break
                continue
            if r[0] == 'Address':
                h = r
                continue
            if h is not None:
                body.append(r)
        ci, cs, csmp = h.index('Instructions Executed'), h.index('Source'), h.index('# Samples')
        stall_cols = [i for i, x in enumerate(h) if x.startswith('stall_') and 'Not Issued' not in x]
        total_i = sum(num(r[ci]) for r in body)
        total_s = sum(num(r[csmp]) for r in body)
        by_inst = sorted(body, key=lambda r: -num(r[ci]))[:top]
        by_stall = sorted(body, key=lambda r: -num(r[csmp]))[:top]
        with open(out + '_hot_sass.csv', 'w', newline='') as f:
            w = csv.writer(f)
            w.writerow(['# total warp instructions', total_i, 'total stall samples', total_s, 'sass instructions', len(body)])
            w.writerow(['rank_by', 'inst_executed', 'stall_samples', 'top_stall', 'sass'])
            for tag, lst in (('inst', by_inst), ('stall', by_stall)):
                for r in lst:
                    st = max(stall_cols, key=lambda i: num(r[i]))
                    w.writerow([tag, r[ci], r[csmp], h[st] if num(r[st]) > 0 else '', r[cs].strip()])
        if '--all-sass' in sys.argv:           # every instruction in address order (regions of a kernel: see tools/sass_regions.py)
            with open(out + '_all_sass.csv', 'w', newline='') as f:
                w = csv.writer(f)
                w.writerow(['address', 'inst_executed', 'stall_samples', 'top_stall', 'sass'])
                for r in body:
                    st = max(stall_cols, key=lambda i: num(r[i]))
                    w.writerow([r[0], r[ci], r[csmp], h[st] if num(r[st]) > 0 else '', r[cs].strip()])
        # opcode histogram (executed warp instructions per SASS mnemonic)
        ops = {}
        for r in body:
            m = re.match(r'\s*(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[cs])
            if m:
                op = m.group(1).split('.')[0]
                ops[op] = ops.get(op, 0) + num(r[ci])
        summary['sass_opcode_mix'] = {k: v for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:25]}
        summary['sass_total_warp_inst'] = total_i
        with open(out + '_summary.json', 'w') as f:
            json.dump(summary, f, indent=1)
    print(json.dumps({k: v for k, v in summary.items() if k != 'launches'}))
    for l in launches:
        print(l['kernel'][:60], l['grid'], l['block'], l.get('gpu__time_duration.sum'), l['warp_stalls_per_issue'])


if __name__ == '__main__':
    main()
