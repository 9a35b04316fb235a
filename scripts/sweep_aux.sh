#!/bin/bash
# Sweep the shared-memory budgets (frames per FFT group) of K1G and K6 on the bench batch.
for kb in 50 72 100 110; do
  echo "== MIXFE_K6_SMEM_KB=$kb MIXFE_K1G_SMEM_KB=$kb"
  MIXFE_K6_SMEM_KB=$kb MIXFE_K1G_SMEM_KB=$kb python scripts/time_aux.py
done
