python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -5
python -m pytest tests -m gpu -q -x 2>&1 | tail -5
