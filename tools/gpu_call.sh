#!/bin/bash
# The body of ONE gpurun call, composed from steps given on the command line (runs ON the GPU box, from the repo root):
#   tools/grun.sh <tag> <timeout_s> 'tools/gpu_call.sh <tag> step [step ...]'
# steps:
#   tests[:<pytest -k expr>]     the -m gpu suite (tail of the report -> gpurun_out/<tag>_tests.log)
#   smoke                        __graft_entry__.smoke()
#   bench                        the default bench line  -> gpurun_out/<tag>_bench_default.json
#   w:<workload>[:ENV=V,...]     one workload alone      -> gpurun_out/<tag>_<workload>[_ENV=V...].json, one summary line
#   env:VAR=value / unset:VAR    set / clear a variable for the following steps (values may hold ':' and ',')
#   lib:<path.so>                the following steps load this build of the core (RT_CORE_LIB; experiment builds)
#   ncu:<workload>:<spp>         ncu --set full of k_render on one launch -> gpurun_out/<tag>_<workload>.ncu-rep
#   launches                     ncu launch list of the default bench     -> gpurun_out/<tag>_launches.csv
#   py:<script.py>               python <script.py> <tag>                 (a measurement script under tools/)
tag=$1; shift
mkdir -p gpurun_out
line() { python - "$1" "$2" <<'P'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d["roofline"]; a = r.get("algorithmic", {})
    print(f"{sys.argv[2]:46s} {d['value']:9.1f} Msamples/s  e2e {d['e2e']['value']:9.1f}  frame {d['ms_per_step']:8.3f} ms  build {d['build']['ms']:.3f} ms"
          f"  frac {r['frac']:.3f}  box/ray {a.get('box_tests_per_ray', 0):.1f}  rec/ray {a.get('record_tests_per_ray', 0):.2f}")
except Exception as e:
    print(sys.argv[2], "FAILED", repr(e)[:120])
P
}
for step in "$@"; do
  case $step in
    tests*) k=${step#tests}; k=${k#:}
      timeout 1500 python -m pytest tests -m gpu -x -q ${k:+-k "$k"} > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$? $(tail -1 gpurun_out/${tag}_tests.log)";;
    smoke) timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1;;
    bench) ( time timeout 900 python bench.py > gpurun_out/${tag}_bench_default.json 2> gpurun_out/${tag}_bench_default.err ) 2>&1 | grep real
      line gpurun_out/${tag}_bench_default.json default;;
    env:*) export "${step#env:}"; ENVTAG="${ENVTAG} ${step#env:}"; echo "export ${step#env:}";;
    unset:*) unset "${step#unset:}"; ENVTAG=""; echo "unset ${step#unset:}";;
    lib:*) export RT_CORE_LIB=$PWD/${step#lib:}; echo "RT_CORE_LIB=$RT_CORE_LIB";;
    w:*) IFS=: read -r _ w envs <<< "$step"
      n=$((n + 1)); out=gpurun_out/${tag}_${n}_${w}${envs:+_${envs//,/_}}${RT_CORE_LIB:+_$(basename $RT_CORE_LIB .so)}.json
      ( [ -n "$envs" ] && export ${envs//,/ }; timeout 600 python bench.py --workload $w --no-cpu-baseline --steps 3 --warmup 3 > $out 2> ${out%.json}.err )
      line $out "$w ${envs}${ENVTAG} ${RT_CORE_LIB:+$(basename $RT_CORE_LIB .so)}";;
    ncu:*) IFS=: read -r _ w spp <<< "$step"
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -f -o gpurun_out/${tag}_$w \
        python bench.py --workload $w --spp $spp --no-cpu-baseline --steps 1 --warmup 0 > gpurun_out/${tag}_ncu_$w.log 2>&1
      echo "ncu $w rc=$?";;
    launches) timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${tag}_launches.csv \
        python bench.py --only-headline --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/${tag}_launches.log 2>&1; echo "launches rc=$?";;
    py:*) timeout 900 python ${step#py:} $tag 2>&1 | tail -40;;
    *) echo "unknown step $step";;
  esac
done
