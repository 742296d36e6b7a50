#!/bin/bash
ORBX_BOW_PROBE=1 timeout 300 python scripts/exp_call_once.py bow 8 2>&1 | tail -8
