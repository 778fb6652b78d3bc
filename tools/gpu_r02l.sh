#!/bin/bash
mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests -m gpu -x -q -k "incremental or scripted or golden or adapter or placing or smoke or waypoints" > gpurun_out/pytest_r02l.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_r02l.log | cut -c1-250
