#!/bin/bash
for w in 1 2; do
  echo "== BSC_DBG_W=$w"; BSC_DBG_W=$w timeout 300 python tools/check_gstore.py 32768 | tail -1
done
