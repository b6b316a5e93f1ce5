#!/bin/bash
# chunk size / taper sweep of the synchronous drop-in call (bench.py e2e leg only)
for mb in 2 4 8 16; do for tp in 0; do
  echo -n "chunk_mb=$mb taper=$tp  "
  TS_HOST_CALL_CPP_MB=$mb TS_HOST_CALL_TAPER=$tp timeout 200 python bench.py --steps 50 --warmup 3 --skip-per --skip-cpu --skip-updates 2>/dev/null | python -c "import json,sys; b=json.load(sys.stdin); print(b['e2e']['ms_per_step'], b['e2e']['variants_ms_per_step'])"
done; done
