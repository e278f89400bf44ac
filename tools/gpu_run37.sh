#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_lstm_cluster.py tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -3
