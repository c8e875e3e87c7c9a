timeout 600 python -m pytest tests/test_gpu_api.py -m gpu -q -x 2>&1 | tail -8
