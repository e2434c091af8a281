#!/bin/bash
# A/B: k_compact3 of batch i beside the front stages of batch i+1 (QUADE_OVERLAP=1) at several CTA splits
run() { echo "== $*"; env "$@" QUADE_SLOTS=4 timeout 200 python tools/probe_steps.py 2>&1 | tail -2 | head -1 | sed 's/submit ms.*kernels/kernels/'; }
run QUADE_OVERLAP=0
run QUADE_OVERLAP=1
run QUADE_OVERLAP=1 QUADE_F6_CTAS=3 QUADE_C3_CTAS=10
run QUADE_OVERLAP=1 QUADE_F6_CTAS=3 QUADE_C3_CTAS=12
run QUADE_OVERLAP=1 QUADE_F6_CTAS=4 QUADE_C3_CTAS=8
run QUADE_OVERLAP=1 QUADE_F6_CTAS=4 QUADE_C3_CTAS=6
run QUADE_OVERLAP=1 QUADE_F6_CTAS=2 QUADE_C3_CTAS=14
run QUADE_OVERLAP=1 QUADE_F6_CTAS=5 QUADE_C3_CTAS=4
