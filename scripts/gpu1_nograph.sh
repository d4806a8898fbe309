#!/bin/bash
for flags in "--no-graph" ""; do
timeout 300 python bench.py --steps 500 --warmup 10 --no-cpu-baseline --no-gpt $flags 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$flags', d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['launch'], d['roofline']['us_per_launch'], d['roofline_stage1']['us_per_launch'])"
done
