cd $GRAFT_REPO_ROOT
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"fill_blocks" -c 2 -o gpurun_out/s33_fill python scripts/prof_matvec.py 1024 4 1 > gpurun_out/s33_fill.log 2>&1
