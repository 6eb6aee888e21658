mkdir -p gpurun_out
timeout 900 bash scripts/ncu_traffic.sh c3 2 24 > gpurun_out/ncu_traffic.out 2>&1; tail -5 gpurun_out/ncu_traffic.out
cp gpurun_out/launches_c3_24h_auto.csv gpurun_out/traffic_c3_24h_auto.csv gpurun_out/traffic_c3_24h_auto_filter.csv gpurun_out/ 2>/dev/null
timeout 400 bash scripts/ncu_kernel.sh xf_pair_kernel r2_xf_pair 6 --hours 2; cp profiles/r2_xf_pair.ncu_summary.txt gpurun_out/
timeout 400 bash scripts/ncu_kernel.sh xf_spectra_kernel r2_xf_spectra 6 --hours 2; cp profiles/r2_xf_spectra.ncu_summary.txt gpurun_out/
timeout 400 bash scripts/ncu_kernel.sh xf_verify_items_kernel r2_xf_verify_items 3 --hours 2; cp profiles/r2_xf_verify_items.ncu_summary.txt gpurun_out/
rm -f gpurun_out/*.ncu-rep
head -12 profiles/r2_xf_pair.ncu_summary.txt; cat gpurun_out/ncu_traffic.json | head -c 1500
