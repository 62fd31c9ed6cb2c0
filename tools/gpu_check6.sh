#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 600 python tools/timeline_cta.py --layers 4 2>&1 | tail -40 | tee gpurun_out/timeline_cta.txt
