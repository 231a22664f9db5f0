#!/usr/bin/env bash
# no-alloc K4 entry point: its test alone in a fresh process, then the whole GPU suite with -x
set -u
mkdir -p gpurun_out
O=gpurun_out/r02_step7
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > ${O}_alone.log 2>&1; echo "alone rc=$?" | tee -a ${O}_alone.log
tail -n 15 ${O}_alone.log
python -m pytest tests -m gpu -x -q > ${O}_full.log 2>&1; echo "full rc=$?" | tee -a ${O}_full.log
tail -n 8 ${O}_full.log
