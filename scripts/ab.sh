#!/bin/bash
# A/B harness: time bench.py with every library variant under scripts/ab/ on the same box
for rep in 1 2; do
for so in scripts/ab/*.so; do
  KOROLLUP_LIB=$PWD/$so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('$so', round(d['ms_per_step'],3), round(d['roofline']['frac'],4), d.get('series_needing_extra_pass'))"
done; done
