python -m pytest tests/test_gpu_distance.py tests/test_gpu_async.py -m gpu -q -x 2>&1 | tail -30
python scripts/_lat.py
