# lego.yml render: frame time, launch list of one chunk, ncu --set full of the fused chain kernel (fine pass)
set -x
python scripts/time_lego_render.py 640000 gpurun_out/lego_now.json > gpurun_out/lego_now.log 2>&1
tail -3 gpurun_out/lego_now.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/lego_launches.csv python scripts/time_lego_render.py 8192 > gpurun_out/lego_ncu.log 2>&1
python scripts/launch_summary.py gpurun_out/lego_launches.csv 12
ncu --set full --clock-control none --import-source on -k regex:flex_fwd_tc -s 2 -c 2 -f -o gpurun_out/flex_fwd python scripts/time_lego_render.py 8192 > gpurun_out/lego_ncu_full.log 2>&1
python scripts/ncu_summary.py gpurun_out/flex_fwd.ncu-rep gpurun_out/flex_fwd_ncu.txt | tail -32
