#!/bin/bash
# experiments on the stored-weight sweep (timings only)
for js in 1 2; do
  echo "== BSC_BWD_JSPLIT=$js"; BSC_BWD_JSPLIT=$js timeout 300 python tools/check_gstore.py 32768 | tail -1
done
echo "== nostore"; BSC_DBG_NOSTORE=1 timeout 300 python tools/check_gstore.py 32768 | tail -1
