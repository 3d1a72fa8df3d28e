run() { # name, env...
  name=$1; shift
  env "$@" python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-cli --no-scoring > gpurun_out/ab_$name.json 2>/dev/null
  python - <<PY
import json; d=json.load(open("gpurun_out/ab_$name.json"))
print("%-18s step %.2f ms  E-step %.2f ms  clocks %s" % ("$name", d["ms_per_step"], d["breakdown_ms"]["estep_kernel"], d["clocks"]["sm_mhz"]))
PY
}
run base A=1
run l2persist ONI_LDA_L2_PERSIST=1
run repl ONI_LDA_LIB=$PWD/oni_ml_b200/variants/liboni_lda_repl.so
run repl_l2 ONI_LDA_LIB=$PWD/oni_ml_b200/variants/liboni_lda_repl.so ONI_LDA_L2_PERSIST=1
run base2 A=1
