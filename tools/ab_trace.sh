#!/bin/bash
# A/B of the ray query kernels on one GPU box: tuner's pick (-1) vs phased (0) / one thread per ray (1) / queued (2)
# THR: thresholds of the queued kernel, flushWaiting | instLanes << 8 | refillLanes << 16
out=gpurun_out/${1:-ab_trace}.txt
: > $out
for K in ${KERNELS:--1 2}; do
  for F in ${THR:-264196}; do
    echo "== trace kernel $K thresholds $F" >> $out
    THEIA_WF_TRACE_KERNEL=$K THEIA_WF_SHADOW_TRACE_KERNEL=$K THEIA_WF_FLUSH_WAITING=$F python tools/quick.py ${WORKLOADS:-c2,c3,c5} >> $out 2>&1
  done
done
cat $out
