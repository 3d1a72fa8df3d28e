# A/B runs of the main E-step launch on ONE box: bench.py with the default build and with variant builds
# (ONI_LDA_LIB) or runtime knobs.  usage: bash profiles/ab_estep.sh name=ENV=VALUE[,ENV=VALUE] ...
run() { # name, env...
  name=$1; shift
  env "$@" python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-cli --no-scoring > gpurun_out/ab_$name.json 2>/dev/null
  python - <<PY
import json; d=json.load(open("gpurun_out/ab_$name.json"))
print("%-18s step %.2f ms  E-step %.2f ms  clocks %s" % ("$name", d["ms_per_step"], d["breakdown_ms"]["estep_kernel"], d["clocks"]["sm_mhz"]))
PY
}
run base A=1
for spec in "$@"; do
  name=${spec%%=*}; envs=${spec#*=}
  run $name $(echo $envs | tr ',' ' ')
done
run base2 A=1
