#!/bin/bash
# single-problem latency through the drop-in planner API (configs[0]) + a quick parity run
timeout 600 python -m pytest tests -m gpu -q -x -k "${PYTEST_K:-onf or lockstep or api}" 2>&1 | tail -3
python -c "
import bench, json
print(json.dumps(bench.single_problem_bench()))
"
