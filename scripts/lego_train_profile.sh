python scripts/time_lego_train.py 3 gpurun_out/lego_train.json 2>&1 | tail -3
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/lego_train_launches.csv python scripts/time_lego_train.py 1 > gpurun_out/lego_train_ncu.log 2>&1
python scripts/launch_summary.py gpurun_out/lego_train_launches.csv 25
