#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "divide" > gpurun_out/pytest_divide.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_divide.log
