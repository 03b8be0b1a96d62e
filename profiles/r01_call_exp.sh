python -m pytest tests/test_gpu_tracer.py -q -k "ale_producers" 2>&1 | tail -25 | cut -c1-220
