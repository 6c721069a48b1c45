cd $GRAFT_REPO_ROOT
S=$PWD/hands-on-point-cloud-processing_b200/libpcnn_study.so
for c in 0 100000; do
  echo "chunk $c: $(PCNN_LIB=$S PCNN_HOST_CHUNK=$c ITERS=8 python tools/pipe_trace.py 2>&1 | grep '^call' | tail -4 | tr '\n' ' ')"
done
PCNN_LIB=$S PCNN_PIPE_TRACE=1 ITERS=3 python tools/pipe_trace.py 2>&1 | tail -34 | grep "H2D done\|search done\|D2H done\|returned" | sed 's/pcnn_knn_host trace (ms since call; device | host enqueue), //; s/  stage /s/; s/ done *//' | tr '\n' ';' | cut -c1-2400; echo
