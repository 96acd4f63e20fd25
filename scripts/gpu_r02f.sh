mkdir -p gpurun_out
bash scripts/run_variants.sh base pfl1 early3 early2
