#!/bin/bash
# CLE kernel time against resident blocks per SM: tools/cle_bps.sh
for b in 1 2 3 4; do
  DFQ_CLE_BLOCKS_PER_SM=$b timeout 120 python bench.py --no-eval --no-micro --no-cpu-baseline --steps 10 2>/dev/null > /tmp/cle_bps.json
  python -c "
import json;d=json.load(open('/tmp/cle_bps.json'));print('blocks_per_sm', $b, 'pipeline_ms', d['value'], 'cle_kernel_ms', d['roofline']['kernel_ms'])"
done
