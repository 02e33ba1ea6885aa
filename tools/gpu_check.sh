#!/bin/bash
# Development aid (run under gpurun): quick extractor parity + bench summary.
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "${1:-extractor or batch}" 2>&1 | tail -3
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > /tmp/b.json 2>/tmp/b.err; tail -2 /tmp/b.err
python tools/show_bench.py < /tmp/b.json
