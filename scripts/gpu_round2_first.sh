#!/bin/bash
# First GPU call of round 2: build and run the two prototypes of scripts/proto/ (each under its own timeout: a protocol
# bug traps after a bounded spin instead of hanging), then the attention timings of the library for comparison.
#   gpurun --timeout 300 -- 'bash scripts/gpu_round2_first.sh'
mkdir -p gpurun_out
NV="nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a"
$NV -o /tmp/pv_proto scripts/proto/pv_tma_proto.cu -lcuda > gpurun_out/r02_proto_build.log 2>&1
$NV -o /tmp/qk_proto scripts/proto/qk_heads_softmax_proto.cu -lcuda >> gpurun_out/r02_proto_build.log 2>&1
{
  echo "== pv_proto small (raw fp32 as the hi plane)";  timeout 60 /tmp/pv_proto 4 256 256;  echo "rc=$?"
  echo "== pv_proto small (masked hi plane)";            timeout 60 /tmp/pv_proto 4 256 256 0; echo "rc=$?"
  echo "== pv_proto GPT-2 size";                          timeout 60 /tmp/pv_proto 96 1024 1024; echo "rc=$?"
  echo "== qk_proto small";                               timeout 60 /tmp/qk_proto 1 256;  echo "rc=$?"
  echo "== qk_proto GPT-2 size";                          timeout 90 /tmp/qk_proto 8 1024; echo "rc=$?"
  echo "== library today";                                timeout 60 python scripts/attn_profile.py; echo "rc=$?"
} > gpurun_out/r02_proto_run.log 2>&1
cat gpurun_out/r02_proto_run.log
