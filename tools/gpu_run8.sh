# A/B of compile-time variants on the box: $1 = extra nvcc defines for the variant
run() { python bench.py --steps 30 --no-cpu-baseline > gpurun_out/$1.json 2>gpurun_out/$1.err; python - <<PY
import json
d=json.loads(open("gpurun_out/$1.json").read().strip().splitlines()[-1])
r=d["roofline"]
print("$1", "%.4g"%d["value"], "ms/step %.4f"%d["ms_per_step"], "kernel us %.1f"%(1e3*r["kernel_ms_total"]/r["kernel_launches"]), "frac %.3f"%r["frac"], "mix %.3f"%r["mix_ceiling"]["frac_of_peak"], "prop %.4g"%d["proposals"]["value"], d["total_loglik"])
PY
}
run base
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -shared -Xcompiler -fPIC -I include"
for v in "$@"; do
  nvcc $FLAGS $v desi_mcmc_b200/csrc/celeste_b200.cu -o desi_mcmc_b200/libceleste_b200.so || exit 1
  run "var$(echo $v | tr -c 'A-Za-z0-9\n' '_')"
done
