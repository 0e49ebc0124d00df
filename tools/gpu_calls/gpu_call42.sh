#!/bin/bash
mkdir -p gpurun_out
python bench.py --workload spheres_1k --no-ref-gpu --no-cpu-baseline > gpurun_out/sph_a.json 2>/dev/null
mkdir -p /tmp/nosmi; ln -sf $(which python) /tmp/nosmi/python
PATH=/tmp/nosmi python bench.py --workload spheres_1k --no-ref-gpu --no-cpu-baseline > gpurun_out/sph_b.json 2>/dev/null
python bench.py --workload spheres_1k --no-ref-gpu --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/sph_c.json 2>/dev/null
