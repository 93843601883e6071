#!/bin/bash
# Host<->device copy bandwidth with N GPUs copying at the same time (one process per GPU): names what limits the
# host-buffer entry (ctc_b200_compute_host) when every rank of a box runs it.  Usage: tools/pcie_concurrency.sh
cd tools/ubench && nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pcie_probe pcie_probe.cu 2>/dev/null; cd ../..
for N in 1 2 4 8; do
  echo "== $N GPUs copying concurrently"
  for ((g = 0; g < N; ++g)); do
    CUDA_VISIBLE_DEVICES=$g tools/ubench/pcie_probe 2>&1 | grep -E "^(H2D|D2H) contiguous|duplex DMA" | sed "s/^/gpu$g /" &
  done
  wait
done
