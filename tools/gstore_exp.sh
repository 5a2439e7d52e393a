#!/bin/bash
echo "== wide"; timeout 300 python tools/check_gstore.py 1000 32768 | grep -v "mode 1"
echo "== BSC_BWD_WIDE=0"; BSC_BWD_WIDE=0 timeout 300 python tools/check_gstore.py 32768 | grep "mode 2"
