set -x
mkdir -p gpurun_out
timeout 300 python tools/probe_vs_wheel_anomaly.py > gpurun_out/r3c_anomaly.txt 2>&1; cat gpurun_out/r3c_anomaly.txt | tail -12
