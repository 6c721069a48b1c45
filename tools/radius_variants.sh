# Developer tool (GPU box): time builds of the library that differ in a compile-time constant of the radius kernels
cd $GRAFT_REPO_ROOT
for so in ${SOS:-libpcnn.so}; do
  echo "== $so"; PCNN_LIB=$PWD/hands-on-point-cloud-processing_b200/$so NQ=1000000 RADII=${RADII:-0.5} CFG=1,0,45 python tools/tune_radius.py 2>&1 | tail -3
done
