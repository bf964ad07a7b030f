#!/bin/bash
# single-frame kernel graphs: parity tests of the replay path, then latency with the graphs off / on
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x --timeout 300 -k "graph_replay or batch_equals_single or bundled_frames or launch_counter or frontend" > gpurun_out/pytest_graph.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_graph.log
tail -8 gpurun_out/pytest_graph.log
for g in 0 1; do
  for rep in 1 2; do
    MSLAM_B200_GRAPH=$g timeout 300 python tools/latency.py > gpurun_out/latency_graph${g}_$rep.json 2> gpurun_out/latency_graph${g}_$rep.err
    echo "graph=$g: $(cat gpurun_out/latency_graph${g}_$rep.json)"
  done
done
