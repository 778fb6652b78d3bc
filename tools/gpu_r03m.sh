#!/bin/bash
mkdir -p gpurun_out
timeout -k 10 1200 python -m pytest tests -x -q -m gpu > gpurun_out/r03m_pytest_gpu.txt 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r03m_pytest_gpu.txt | cut -c1-300
