# Developer tool (GPU box): time builds of the library that differ in a compile-time constant of the ICP walk
cd $GRAFT_REPO_ROOT
for so in ${SOS:-libpcnn.so}; do
  echo "== $so"; PCNN_LIB=$PWD/hands-on-point-cloud-processing_b200/$so python tools/icp_iter_times.py 2>&1 | grep "icp_pdl=1\|nn1"
done
