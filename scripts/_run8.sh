python -m pytest tests/test_gpu_planning.py tests/test_gpu_planner_api.py -m gpu -q -x 2>&1 | tail -8
