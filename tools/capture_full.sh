# ncu --set full captures of the three hot GEMM kernels at one representative unet layer each (tools/bench_conv.py /
# bench_wgrad.py run 7 launches of the layer: skip 3, capture 1).  Each capture follows a plain run of the same command.
set -e
O=gpurun_out
ONLY="c3 512->512" python tools/bench_conv.py > $O/r02_t_plain1.log 2>&1 && ONLY="c3 512->512" ncu --set full --clock-control none --import-source on -k regex:igemm_fwd2 -s 3 -c 1 -o $O/r02_t_fwd2_256_pair python tools/bench_conv.py > $O/r02_t_ncu1.log 2>&1
ONLY="c3 64->64" python tools/bench_conv.py > $O/r02_t_plain2.log 2>&1 && ONLY="c3 64->64" ncu --set full --clock-control none --import-source on -k regex:igemm_fwd2 -s 3 -c 1 -o $O/r02_t_fwd2_64 python tools/bench_conv.py > $O/r02_t_ncu2.log 2>&1
ONLY=128,256,256 python tools/bench_wgrad.py > $O/r02_t_plain3.log 2>&1 && ONLY=128,256,256 ncu --set full --clock-control none --import-source on -k regex:igemm_wgrad2 -s 3 -c 1 -o $O/r02_t_wgrad2_wide_pair python tools/bench_wgrad.py > $O/r02_t_ncu3.log 2>&1
cat $O/r02_t_plain1.log $O/r02_t_plain2.log $O/r02_t_plain3.log
