#!/bin/bash
# e2e per-step parts (page-locked and pageable inputs), then the GPU suite
for v in 0 1; do
PTB200_E2E_PAGEABLE=$v PTB200_E2E_DEBUG=1 timeout 600 python bench.py --steps 4 --warmup 2 --no-cpu --no-secondary 2>&1 >/dev/null | grep "e2e step" | tr '\n' ';'; echo
done
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 2
