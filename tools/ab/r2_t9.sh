#!/bin/bash
cd /root/repo
python tools/overlap_probe.py 16384
PROBE_MAIN_STREAM=1 python tools/overlap_probe.py 16384
