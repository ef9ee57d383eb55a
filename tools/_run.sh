timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_t7.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_t7.log
timeout 900 python tools/tree_sweep.py --caps 16 18 20 22 24 26 --cpu > gpurun_out/r02_tree_sweep.txt 2>&1; echo "sweep rc=$?"; cat gpurun_out/r02_tree_sweep.txt
