# Developer tool (GPU box): time builds of the library that differ in a compile-time constant of the kNN kernels
cd $GRAFT_REPO_ROOT
for so in libpcnn.so libpcnn_q8.so; do
  echo "== $so"; PCNN_LIB=$PWD/hands-on-point-cloud-processing_b200/$so KS=32,20,24 REPS=4 python tools/knn_time.py 2>&1 | grep "^k="
done
