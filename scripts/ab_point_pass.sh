#!/bin/bash
# A/B of the operator-pass variants on the GPU box (stage times)
for v in 0 1 2; do echo "== SBA_CP_VARIANT=$v"; SBA_CP_VARIANT=$v python scripts/stage_times.py venice 20 | grep -E "matvec|iterations"; done
