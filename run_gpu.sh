#!/bin/bash
mkdir -p gpurun_out
timeout 300 ./build_variants/pcie_write_probe 2>&1 | tee gpurun_out/pcie_write_probe.txt
