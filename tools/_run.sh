#!/bin/bash
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python tools/e2e_profile.py 2>&1 | tail -2
