#!/bin/bash
# e2e throughput of orbx_extract_batch vs lanes / chunk size (host-side tuning; results are identical for any setting)
for l in 2 3 4; do for c in 16 32 64; do
  ORBX_LANES=$l ORBX_CHUNK=$c python bench.py --steps 6 --warmup 3 --skip-cpu --skip-hamming --skip-workloads --skip-parity 2>/dev/null | python -c "
import json,sys; b=json.loads(sys.stdin.read()); print('lanes', $l, 'chunk', $c, 'e2e', round(b['e2e']['value']), 'device', round(b['value']))"
done; done
