# Developer tool (GPU box): pcnn_knn_host call time against stage size / stream count / ramp
for ramp in 1 0; do for st in 6 8; do for ch in 98304 131072 163840; do echo "ramp $ramp streams $st chunk $ch"; PCNN_HOST_RAMP=$ramp PCNN_HOST_STREAMS=$st PCNN_HOST_CHUNK=$ch ITERS=8 python tools/pipe_trace.py 2>&1 | grep "^call" | tail -4 | tr '\n' ' '; echo; done; done; done
PCNN_PIPE_TRACE=1 python tools/pipe_trace.py 2>&1 | tail -32
