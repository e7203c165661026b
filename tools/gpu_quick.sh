#!/bin/bash
mkdir -p gpurun_out
timeout 300 ncu --set full --import-source on --clock-control none -k regex:householder_qr --launch-skip 2 -c 1 -f -o gpurun_out/qr_full python tools/qr_probe.py 974 520 1 2>&1 | tail -2
