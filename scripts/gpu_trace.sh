#!/bin/bash
mkdir -p gpurun_out
export B200DDP_NVCC_DEFINES=-DB200_CONV_TRACE
timeout 300 python scripts/conv_trace.py 2>&1 | tee gpurun_out/conv_trace.log
