# Developer tool (GPU box): time builds of the library that differ in a compile-time constant of the kNN kernels
cd $GRAFT_REPO_ROOT
for so in ${SOS:-libpcnn.so}; do
  echo "== $so"; PCNN_LIB=$PWD/hands-on-point-cloud-processing_b200/$so KS=${KS:-8,10,16,32} REPS=4 python tools/knn_time.py 2>&1 | grep "^k="
done
