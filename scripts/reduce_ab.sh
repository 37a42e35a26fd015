#!/bin/bash
for acc in ${ACCS:-0 1 2}; do for mb in ${MBS:-1 2}; do echo "ACC=$acc MB=$mb"; DLAP_REDUCE_ACC=$acc DLAP_REDUCE_MB=$mb python scripts/bwd_timeline.py 2>&1 | grep "REDUCE CTA 0"; done; done
