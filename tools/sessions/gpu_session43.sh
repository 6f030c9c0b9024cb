# final-state validation of the round's last tree (after the frame-buffer cache and the CLI's parallel replication)
R=gpurun_out/r06
mkdir -p $R
( time timeout 150 python -m pytest tests -m gpu -x -q --durations=5 ) > $R/pytest_final.log 2>&1; echo "pytest rc=$?" >> $R/pytest_final.log
timeout 40 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $R/smoke_final.log 2>&1
tail -n 4 $R/pytest_final.log; tail -n 2 $R/smoke_final.log
