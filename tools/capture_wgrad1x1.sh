# ncu --set full of the first-generation weight-gradient kernel on a 1x1 layer (32->32 @512^2, batch 32) - after the plain run.
set -e
O=gpurun_out
MODE=1x1 ONLY=512,32,32 python tools/bench_wgrad.py > $O/w1x1_plain.log 2>&1
MODE=1x1 ONLY=512,64,64 python tools/bench_wgrad.py >> $O/w1x1_plain.log 2>&1
cat $O/w1x1_plain.log
MODE=1x1 ONLY=512,32,32 ncu --set full --clock-control none --import-source on -k regex:igemm_wgrad_kernel -s 3 -c 1 -o $O/w1x1_32 python tools/bench_wgrad.py > $O/w1x1_ncu.log 2>&1
ncu -i $O/w1x1_32.ncu-rep --page raw --csv > $O/w1x1_32_raw.csv 2>/dev/null
python - <<'PY'
import csv
rows = list(csv.reader(open("gpurun_out/w1x1_32_raw.csv")))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ("gpu__time_duration.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "lts__throughput.avg.pct",
        "l1tex__throughput.avg.pct", "sm__throughput.avg.pct", "sm__inst_executed_pipe_tmem", "tensor", "smsp__warp_issue_stalled",
        "l1tex__data_pipe", "sm__warps_active", "lts__t_sectors_srcunit_tex_op_read", "tma", "sm__pipe_tensor")
for h, u, v in zip(hdr, units, vals):
    if any(w in h for w in want) and ("pct" in h or "sum" in h and "bytes" in h or "duration" in h):
        print(h, v, u)
PY
