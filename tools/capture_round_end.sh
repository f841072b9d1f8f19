# ncu launch lists (durations + DRAM bytes per launch) of one step of unet (BASELINE configs[1]) and u2net, final state of
# the round; run ONLY after the same commands exited 0 without ncu (they do, first, to write the tag files).
set -e
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
python tools/ncu_step.py --model unet --batch 32 --tags gpurun_out/r02_bh_tags_unet.json && ncu --metrics $M --clock-control none -c 3000 --csv --log-file gpurun_out/r02_bh_launches_unet.csv python tools/ncu_step.py --model unet --batch 32 --tags gpurun_out/r02_bh_tags_unet.json > gpurun_out/r02_bh_ncu1.log 2>&1
python tools/ncu_step.py --model u2net --batch 32 --tags gpurun_out/r02_bh_tags_u2net.json && ncu --metrics $M --clock-control none -c 9000 --csv --log-file gpurun_out/r02_bh_launches_u2net.csv python tools/ncu_step.py --model u2net --batch 32 --tags gpurun_out/r02_bh_tags_u2net.json > gpurun_out/r02_bh_ncu2.log 2>&1
echo done
