#!/bin/bash
# usage: scripts/stage_ms.sh [label]  -- short bench run, prints per-stage ms (debug helper)
python bench.py --steps 50 --warmup 5 --no-cpu 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step'],4), d['stage_ms'], 'e2e', round(d['e2e']['value'],1))"
