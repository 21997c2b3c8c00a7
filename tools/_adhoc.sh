cd /root/repo
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_algorithms.py -m gpu -q -x -p no:cacheprovider -k "sort or build or csr" 2>&1 | tail -2
for pz in 1 0; do echo "== QDSP_RS_PERSIST=$pz"; QDSP_RS_PERSIST=$pz timeout 300 python tools/bench_paths.py sort build 2>&1 | cut -c1-120,200-260; done
