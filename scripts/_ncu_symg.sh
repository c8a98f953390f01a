cd $GRAFT_REPO_ROOT
for np in 5 6; do
timeout 500 ncu --set full --clock-control none --import-source on -k regex:symg_ring -c 2 -o gpurun_out/s33_symgring$np python scripts/prof_matvec.py 512 $np 1 > gpurun_out/s33_symgring$np.log 2>&1
done
