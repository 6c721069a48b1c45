set -x
cd $GRAFT_REPO_ROOT
python tools/sanitize_subset.py > gpurun_out/san_plain.log 2>&1; tail -1 gpurun_out/san_plain.log
for tool in memcheck synccheck racecheck; do
  SAN_N=4000 timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --log-file gpurun_out/r02_sanitizer_$tool.txt python tools/sanitize_subset.py > gpurun_out/san_$tool.log 2>&1
  echo "$tool rc=$?"; tail -3 gpurun_out/r02_sanitizer_$tool.txt; tail -2 gpurun_out/san_$tool.log
done
