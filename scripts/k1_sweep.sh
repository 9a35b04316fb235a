#!/bin/bash
# K1 time for several start skews of the second CTA on each SM
for ns in 0 2000 4000 6000 8000 12000; do
  echo "skew $ns ns"; MIXFE_K1_SKEW_NS=$ns timeout 100 python scripts/k1_time.py | head -2
done
