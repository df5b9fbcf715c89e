#!/bin/bash
# A/B of the prefilter kernels' thread mapping (threads per 128-byte row; one process per setting)
for sp in 1 2; do
  echo "== TRX_FILTER_SPLIT=$sp"
  TRX_FILTER_SPLIT=$sp python profiles/time_scan.py "LMA(E|Q)GLYN" "C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H" "W.{2,4}C.{5,9}CW" "D.{0,3}EW" 2>&1 | grep "filter=on"
done
