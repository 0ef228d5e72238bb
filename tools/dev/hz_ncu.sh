#!/bin/bash
# ncu capture of the two row-operand conv1 kernels (plain run first); summaries are parsed on the box
python tools/dev/hz_probe.py 20 2>&1 | grep True
for k in stats:2 pool:7; do
  n=${k%%:*}; sk=${k##*:}
  ncu --set full --clock-control none --import-source on -k regex:conv1_hz -s $sk -c 1 -f -o /tmp/hz_$n python tools/dev/hz_probe.py 2 > /dev/null 2>&1
  ncu -i /tmp/hz_$n.ncu-rep --page raw --csv > gpurun_out/hz_${n}_raw.csv
  ncu -i /tmp/hz_$n.ncu-rep --page source --csv > gpurun_out/hz_${n}_src.csv 2>/dev/null
done
