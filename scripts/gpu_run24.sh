mkdir -p gpurun_out
timeout 600 bash scripts/ncu_kernel.sh xf_spectra_kernel r2_xf_spectra 6 --hours 2
cp profiles/r2_xf_spectra.ncu_summary.txt gpurun_out/
ncu -i gpurun_out/r2_xf_spectra.ncu-rep --page source --csv > gpurun_out/r2_xf_spectra.source.csv 2>/dev/null
rm -f gpurun_out/r2_xf_spectra.ncu-rep
cat profiles/r2_xf_spectra.ncu_summary.txt
