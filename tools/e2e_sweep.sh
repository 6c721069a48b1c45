cd $GRAFT_REPO_ROOT
S=$PWD/hands-on-point-cloud-processing_b200/libpcnn_study.so
for c in 0 65536 131072 200000; do
  echo "chunk $c"; PCNN_LIB=$S PCNN_HOST_CHUNK=$c ITERS=6 python tools/pipe_trace.py 2>&1 | grep "^call" | tail -3 | tr '\n' ' '; echo
done
PCNN_LIB=$S PCNN_PIPE_TRACE=1 ITERS=3 python tools/pipe_trace.py 2>&1 | tail -32 | grep "search done\|D2H done\|returned" | tr '\n' ';' | cut -c1-1600
