timeout 600 python tools/gather_sweep.py --quick > gpurun_out/r02_gather_sweep.txt 2>&1; echo "sweep rc=$?"; tail -3 gpurun_out/r02_gather_sweep.txt
timeout 900 python tools/tree_sweep.py --caps 16 18 20 22 24 26 --cpu > gpurun_out/r02_tree_sweep.txt 2>&1; echo "tree rc=$?"; tail -3 gpurun_out/r02_tree_sweep.txt
timeout 300 python tools/update_phases.py --batch 512 1024 > gpurun_out/r02_update_phases.txt 2>&1; tail -2 gpurun_out/r02_update_phases.txt
