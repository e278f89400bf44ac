#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:lstm_cluster --csv --log-file gpurun_out/lstm_times.csv python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_lstm_times.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/lstm_times.csv')) if len(r)>8]
ix={h:i for i,h in enumerate(rows[0])}
print('lstm launch ms:', [round(float(r[ix['Metric Value']])/1e6,2) for r in rows[1:]])
PY
