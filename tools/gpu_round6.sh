set -x
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.avg.per_second,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,launch__registers_per_thread,launch__grid_size,launch__block_size,launch__shared_mem_per_block_dynamic,launch__cluster_size,dram__bytes_read.sum,lts__t_sector_hit_rate.pct,sm__warps_active.avg.pct_of_peak_sustained_active \
  --clock-control none -k regex:'attn_sm100_v3|fmha|Fmha|flash_fwd' -c 12 --csv --log-file gpurun_out/compare_dense_ncu.csv python tools/compare_dense.py --n 131072 > gpurun_out/compare_dense_ncu.log 2>&1
tail -3 gpurun_out/compare_dense_ncu.log
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/compare_dense_ncu.csv')) if len(r)>10]
h=rows[0]; ix={k:i for i,k in enumerate(h)}
seen=set()
for r in rows[1:]:
    key=(r[ix['Kernel Name']][:60], r[ix['Metric Name']])
    if key in seen: continue
    seen.add(key)
    print(r[ix['Kernel Name']][:60], '|', r[ix['Metric Name']], '|', r[ix['Metric Value']], r[ix['Metric Unit']])
PY
