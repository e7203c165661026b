#!/bin/bash
timeout 600 python -m pytest tests/test_live_reference.py tests/test_full_size.py -m gpu -x -q 2>&1 | tail -5
