#!/bin/bash
for sh in 0/8 1/8 2/8 3/8 4/8 5/8 6/8 7/8; do SPLIT=1 GRAPH=1 SHARD=$sh timeout 300 python tools/gemm_probe.py 2>&1 | tail -2 | tr '\n' ' ' | cut -c1-150; echo; done
