# ncu --set full of the FAST N = 64 kernel on conv3x3 32->32 @512^2 batch 8 (nested_unet level 0) - after the plain run.
set -e
O=gpurun_out
BATCH=8 SPEC=c3,512,32,32 python tools/bench_conv.py > $O/r02_cb_plain.log 2>&1
BATCH=8 SPEC=c3,512,32,32 ncu --set full --clock-control none --import-source on -k regex:igemm_fwd2 -s 3 -c 1 -o $O/r02_cb_fwd2_fast32 python tools/bench_conv.py > $O/r02_cb_ncu.log 2>&1
cat $O/r02_cb_plain.log
