"""Turns gpurun_out/*.ncu-rep and launch-list CSVs into the markdown summaries kept in
profiles/. Usage: python profiles/summarize.py <report.ncu-rep> <out.md> "<title>" """
import csv
import subprocess
import sys

KEEP = ['Kernel Name', 'Block Size', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum',
        'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'smsp__inst_executed.sum',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_st.sum']


def main():
    rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(out, 'w') as fh:
        fh.write('# %s\n\nSource: `%s` (`ncu --set full --clock-control none`), one table per captured launch.\n\n' % (title, rep))
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            fh.write('| metric | value | unit |\n|---|---|---|\n')
            for k in KEEP:
                if k in d:
                    fh.write('| %s | %s | %s |\n' % (k, d[k], units[hdr.index(k)]))
            fh.write('\n')


if __name__ == '__main__':
    main()
