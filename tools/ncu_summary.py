"""Print the metrics we track from an .ncu-rep (read here, no GPU):  python tools/ncu_summary.py file.ncu-rep [--json out.json]
--json also writes them as a list of {kernel, grid, block, metrics{name: value}, units{name: unit}}: what bench.py reads for the
measured-traffic fields of its roofline objects (profiles/r02_stage_metrics.json)."""
import csv, io, json, subprocess, sys
KEEP = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__occupancy_limit_registers',
        'sm__warps_active.avg.per_cycle_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'sm__inst_executed.sum.per_cycle_active', 'sm__inst_executed.sum.per_cycle_elapsed', 'sm__cycles_active.avg', 'sm__cycles_active.max',
        'sm__cycles_active.min', 'sm__cycles_elapsed.avg', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'smsp__pcsamp_sample_count', 'smsp__pcsamp_warps_issue_stalled_wait', 'smsp__pcsamp_warps_issue_stalled_long_scoreboard',
        'smsp__pcsamp_warps_issue_stalled_short_scoreboard', 'smsp__pcsamp_warps_issue_stalled_branch_resolving',
        'smsp__pcsamp_warps_issue_stalled_not_selected', 'smsp__pcsamp_warps_issue_stalled_selected', 'smsp__pcsamp_warps_issue_stalled_no_instructions',
        'smsp__pcsamp_warps_issue_stalled_math_pipe_throttle', 'smsp__pcsamp_warps_issue_stalled_mio_throttle',
        'smsp__pcsamp_warps_issue_stalled_dispatch_stall', 'lts__t_bytes.sum', 'l1tex__t_bytes.sum',
        'sm__sass_thread_inst_executed_op_fadd_pred_on.sum', 'sm__sass_thread_inst_executed_op_fmul_pred_on.sum',
        'sm__sass_thread_inst_executed_op_ffma_pred_on.sum', 'smsp__sass_thread_inst_executed_op_fadd_pred_on.sum',
        'smsp__sass_thread_inst_executed_op_fmul_pred_on.sum', 'smsp__sass_thread_inst_executed_op_ffma_pred_on.sum']
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
KEEP += ['l1tex__m_xbar2l1tex_read_bytes.sum', 'l1tex__m_l1tex2xbar_write_bytes.sum', 'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum',
         'lts__t_sectors_srcunit_tex.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.avg.per_cycle_active',
         'smsp__cycles_active.avg', 'launch__shared_mem_per_block_dynamic', 'sm__maximum_warps_per_active_cycle_pct', 'sm__warps_active.avg.pct_of_peak_sustained_active',
         # the L1 / LSU side (round 2, late): data-pipe wavefronts are the second-busiest unit of the shadow pass
         'l1tex__throughput.avg.pct_of_peak_sustained_active', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
         'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed',
         'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum',
         'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active']
kernels = []
for r in rows[2:]:
    print('==', r[ix['Kernel Name']][:90], 'grid', r[ix['Grid Size']], 'block', r[ix['Block Size']])
    k_out = {'kernel': r[ix['Kernel Name']], 'grid': r[ix['Grid Size']], 'block': r[ix['Block Size']], 'metrics': {}, 'units': {}}
    for k in KEEP:
        if k in ix:
            print(f'  {k} [{units[ix[k]]}] = {r[ix[k]]}')
            try:                                # JSON values are normalised: bytes, milliseconds
                v, u = float(r[ix[k]].replace(',', '')), units[ix[k]]
                scale = {'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'Tbyte': 1e12}.get(u)
                if scale: v, u = v * scale, 'byte'
                tscale = {'s': 1e3, 'us': 1e-3, 'ns': 1e-6, 'msecond': 1.0, 'usecond': 1e-3, 'nsecond': 1e-6, 'second': 1e3}.get(u)
                if tscale: v, u = v * tscale, 'ms'
                k_out['metrics'][k] = v
                k_out['units'][k] = u
            except ValueError:
                pass
    kernels.append(k_out)
if '--json' in sys.argv:
    with open(sys.argv[sys.argv.index('--json') + 1], 'w') as f:
        json.dump({'source': sys.argv[1], 'kernels': kernels}, f, indent=1)
