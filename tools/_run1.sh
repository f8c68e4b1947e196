mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python tools/prof_all.py > gpurun_out/prof_all_plain.log 2>&1; echo "prof_all rc $?"; tail -2 gpurun_out/prof_all_plain.log
timeout 1500 ncu --set full --clock-control none -o gpurun_out/all_r2b -f python tools/prof_all.py > gpurun_out/prof_all_ncu.log 2>&1; echo "ncu rc $?"
ncu -i gpurun_out/all_r2b.ncu-rep --page raw --csv > gpurun_out/all_r2b_raw.csv 2>/dev/null; ls -la gpurun_out/all_r2b*; 
sz=$(stat -c %s gpurun_out/all_r2b.ncu-rep); if [ "$sz" -gt 40000000 ]; then rm gpurun_out/all_r2b.ncu-rep; fi
