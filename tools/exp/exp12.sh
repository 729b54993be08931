#!/bin/bash
L=gpurun_out/exp12.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest3.log 2>&1; tail -3 gpurun_out/pytest3.log > $L
python bench.py --steps 10 --warmup 3 > gpurun_out/bench3.log 2>&1 && tail -1 gpurun_out/bench3.log >> $L
python bench.py --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 250 -c 80 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log >> $L
