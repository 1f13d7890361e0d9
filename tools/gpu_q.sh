#!/bin/bash
python -m pytest tests/test_hatx.py -m gpu -x -q 2>&1 | tail -15
