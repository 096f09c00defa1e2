#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -2
