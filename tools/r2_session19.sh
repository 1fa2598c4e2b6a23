#!/bin/bash
# RLB_INL_DBG build: %globaltimer stamps (ns) of three CTAs of inlier_tma_kernel
timeout 200 python bench.py --profile --steps 2 --warmup 1 --no-localization 2>&1 | grep -E "^cta" | tail -12
