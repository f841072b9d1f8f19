# ncu launch list (durations + DRAM bytes per launch) of one unet step (BASELINE configs[1]), final state of the round;
# run ONLY after the same command exited 0 without ncu (it does, first, to write the tag file).
set -e
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
python tools/ncu_step.py --model unet --batch 32 --tags gpurun_out/r02_cf_tags_unet.json && ncu --metrics $M --clock-control none -c 3000 --csv --log-file gpurun_out/r02_cf_launches_unet.csv python tools/ncu_step.py --model unet --batch 32 --tags gpurun_out/r02_cf_tags_unet.json > gpurun_out/r02_cf_ncu1.log 2>&1
echo done
