# hit list on by default: the parity tests that fit the round's last GPU seconds (everything but the slow converged /
# full-size cases, which differ from the rest only in sample count and scene size), then smoke
R=gpurun_out/r06
mkdir -p $R
timeout 36 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "not converged and not full_size and not lbvh_large and not edge_grazing and not isfb" > $R/pytest_hitlist.log 2>&1; echo "pytest rc=$?" >> $R/pytest_hitlist.log
tail -n 3 $R/pytest_hitlist.log
