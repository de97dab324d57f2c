python tools/trace_x3.py 0 1200 > gpurun_out/trace_pts.txt 2>&1; head -1 gpurun_out/trace_pts.txt
