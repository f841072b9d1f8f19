set -e
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
python tools/ncu_step.py --model resunet --batch 32 --eval --tags gpurun_out/r02_n_tags_resunet_eval.json && ncu --metrics $M --clock-control none -c 3000 --csv --log-file gpurun_out/r02_n_launches_resunet_eval.csv python tools/ncu_step.py --model resunet --batch 32 --eval --tags gpurun_out/r02_n_tags_resunet_eval.json > gpurun_out/r02_n_ncu1.log 2>&1
python tools/ncu_step.py --model nested_unet --batch 8 --tags gpurun_out/r02_n_tags_nested.json && ncu --metrics $M --clock-control none -c 4000 --csv --log-file gpurun_out/r02_n_launches_nested.csv python tools/ncu_step.py --model nested_unet --batch 8 --tags gpurun_out/r02_n_tags_nested.json > gpurun_out/r02_n_ncu2.log 2>&1
python tools/ncu_step.py --model attention_unet --batch 16 --tags gpurun_out/r02_n_tags_attention.json && ncu --metrics $M --clock-control none -c 4000 --csv --log-file gpurun_out/r02_n_launches_attention.csv python tools/ncu_step.py --model attention_unet --batch 16 --tags gpurun_out/r02_n_tags_attention.json > gpurun_out/r02_n_ncu3.log 2>&1
echo done
