python bench.py --steps 1 --warmup 1 --batch 2 --workers 1 --no-cpu --no-clocks > gpurun_out/p7_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_s19.csv python bench.py --steps 1 --warmup 1 --batch 2 --workers 1 --no-cpu --no-clocks > gpurun_out/p7_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_qa_pass -s 10 -c 10 -f -o gpurun_out/prof_qa_s19 python bench.py --steps 1 --warmup 1 --batch 2 --workers 1 --no-cpu --no-clocks > gpurun_out/p7_ncu2.log 2>&1
ls -la gpurun_out/ | tail -5
