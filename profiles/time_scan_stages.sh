#!/bin/bash
# A/B of the prefilter kernels' TMA ring depth (one process per setting: the depth is read once per process)
for st in 2 3 4; do
  echo "== TRX_FILTER_STAGES=$st"
  TRX_FILTER_STAGES=$st python profiles/time_scan.py "LMA(E|Q)GLYN" "C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H" "W.{2,4}C.{5,9}CW" 2>&1 | grep "filter=on"
done
