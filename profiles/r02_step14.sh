#!/usr/bin/env bash
# round 2, step 14: geometry sweep of the CTA variant at d=11 (threads, streams, batch, CTAs per SM)

export QEC_B200_LIN_THREADS=1024
python tools/tune_lin.py ':/@11,11!nb' 2>&1 | cut -c1-200
QEC_B200_LIN_STREAMS=32 python tools/tune_lin.py ':/@11,11!nb' 2>&1 | cut -c1-200
QEC_B200_LIN_STREAMS=128 python tools/tune_lin.py ':/@11,11!nb' 2>&1 | cut -c1-200
QEC_B200_LIN_CTAS=1 QEC_B200_LIN_LGB=9 python tools/tune_lin.py ':/@11,11!nb' 2>&1 | cut -c1-200
QEC_B200_LIN_THREADS=512 QEC_B200_LIN_CTAS=2 QEC_B200_LIN_LGB=9 python tools/tune_lin.py ':/@11,11!nb' 2>&1 | cut -c1-200
python tools/tune_lin.py ':/@11,11' 2>&1 | cut -c1-200
python tools/tune_lin.py 'cta:/@7,7' 'cta:/@5,5' 2>&1 | cut -c1-200
