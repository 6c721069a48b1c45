export PCNN_LIB=/root/repo/hands-on-point-cloud-processing_b200/libpcnn_study.so
python tools/knn_time.py 2>&1 | grep "^k="
for c in 64 128 256 512 2048; do for m in 4 8 16; do
echo "chunk $c min $m: $(METHOD=10 PCNN_REFILL_CHUNK=$c PCNN_REFILL_MIN=$m python tools/knn_time.py 2>&1 | grep '^k=')"
done; done
echo "global m8: $(METHOD=10 python tools/knn_time.py 2>&1 | grep '^k=')"
