"""Prints the judged subset of an `ncu --page raw --csv` export, one block per kernel launch.
Usage: ncu -i X.ncu-rep --page raw --csv > X_raw.csv ; python profiles/ncu_extract.py X_raw.csv"""
import csv
import sys

WANT = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'launch__occupancy_limit_registers', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__inst_executed.avg.per_cycle_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_fmaheavy.sum', 'sm__inst_executed_pipe_fmalite.sum',
        'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_xu.sum', 'sm__inst_executed_pipe_lsu.sum',
        'sm__inst_executed_pipe_uniform.sum', 'sm__inst_executed_pipe_cbu.sum', 'sm__inst_executed_pipe_adu.sum',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__sass_thread_inst_executed_op_ffma_pred_on.sum', 'smsp__sass_thread_inst_executed_op_fadd_pred_on.sum',
        'smsp__sass_thread_inst_executed_op_fmul_pred_on.sum',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_bytes.sum', 'lts__t_bytes.sum', 'sm__cycles_elapsed.avg', 'sm__cycles_active.avg',
        'smsp__cycles_active.avg', 'smsp__thread_inst_executed_per_inst_executed.ratio']


def main(path, extra=()):
    rows = list(csv.reader(open(path)))
    h, units, data = rows[0], rows[1], rows[2:]
    for r in data:
        print('-----')
        for w in list(WANT) + list(extra):
            if w in h:
                i = h.index(w)
                print(f"{w:72s} {r[i]:>22s} {units[i]}")
        # warp stall breakdown (issue-slot samples)
        st = [(float(r[i].replace(',', '') or 0), h[i]) for i in range(len(h))
              if h[i].startswith('smsp__average_warps_issue_stalled') and h[i].endswith('per_issue_active.ratio')
              or h[i].startswith('smsp__average_warp_latency_issue_stalled') and h[i].endswith('.ratio')]
        for v, n in sorted(st, reverse=True)[:8]:
            print(f"   stall {n:90s} {v:8.3f}")


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2:])
