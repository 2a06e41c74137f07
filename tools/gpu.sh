#!/bin/bash
# The one GPU-box launcher (run through gpurun): tools/gpu.sh <tag> <step> [<step> ...]; every step under its own timeout, logs
# into gpurun_out/<tag>_*.  Steps:
#   tests              pytest -m gpu
#   bench              bench.py (default flags) -> <tag>_bench.json
#   quick              bench.py --no-cpu-baseline --steps 3 (fast A/B line)
#   ref                bench.py --impl reference
#   ab:<sfx>[,<sfx>..] quick bench for each build variant mitsuba3_plt_b200/libpltb200_<sfx>.so (tools/build_variant.sh); "" = default
#   launches[:scene:res:spp]  ncu launch list (gpu__time_duration.sum) of two renders
#   full[:scene:res:spp]      ncu --set full of one render; details / raw / source pages exported as text
#   configs[:C1,C4]    tools/config_bench.py
#   anim               tools/anim_bench.py
#   sanitize           compute-sanitizer memcheck + racecheck + initcheck of the smoke render (closed on this pool: see debugchecks)
#   debugchecks        pytest -m gpu against a -DPLTB_DEBUG_CHECKS build (device-side assertions)
T=${1:-r}; shift
mkdir -p gpurun_out
O=gpurun_out/$T
target() { IFS=: read -r _ sc res spp <<< "$1"; echo "${sc:-cbox-grating} ${res:-1024} ${spp:-8}"; }
for step in "$@"; do
  echo "== $step"
  case $step in
    tests) timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee ${O}_tests.log ;;
    bench) timeout 900 python bench.py > ${O}_bench.json 2> ${O}_bench.err; cat ${O}_bench.json ;;
    quick) timeout 300 python bench.py --no-cpu-baseline --steps 3 > ${O}_quick.json 2> ${O}_quick.err; cat ${O}_quick.json ;;
    ref) timeout 600 python bench.py --impl reference > ${O}_bench_ref.json 2> ${O}_bench_ref.err; cat ${O}_bench_ref.json ;;
    ab:*)
      IFS=, read -ra VS <<< "${step#ab:},"
      for v in "" "${VS[@]}"; do
        lib=mitsuba3_plt_b200/libpltb200${v:+_$v}.so; [ -f $lib ] || continue
        echo "-- variant '${v:-default}'"
        PLTB200_LIB=$lib timeout 200 python bench.py --no-cpu-baseline --steps 3 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['kernel_ms_per_step'], {k: round(v['ms'], 1) for k, v in d['kernels'].items()})"
      done 2>&1 | tee ${O}_ab.log ;;
    launches*)
      a=$(target $step)
      timeout 300 python tools/ncu_target.py $a 2 > ${O}_target.log 2>&1 && \
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file ${O}_launches.csv \
          python tools/ncu_target.py $a 2 > ${O}_ncu_l.log 2>&1 ;;
    full*)
      a=$(target $step)
      timeout 300 python tools/ncu_target.py $a 1 > ${O}_target.log 2>&1 && \
      timeout 2400 ncu --set full --clock-control none --import-source on -o /tmp/${T}_full -f \
          python tools/ncu_target.py $a 1 > ${O}_ncu_f.log 2>&1
      ncu -i /tmp/${T}_full.ncu-rep --page details > ${O}_details.txt 2>/dev/null
      ncu -i /tmp/${T}_full.ncu-rep --page raw --csv > ${O}_raw.csv 2>/dev/null
      ncu -i /tmp/${T}_full.ncu-rep --page source --csv --kernel-name regex:"k_shade|k_nee_post|k_trace|k_raygen|k_classify|k_nee_light" > ${O}_source.csv 2>/dev/null
      gzip -f ${O}_source.csv ${O}_raw.csv ${O}_details.txt
      tail -n 3 ${O}_ncu_f.log ;;
    configs*)
      only=${step#configs}; only=${only#:}
      timeout 1500 python tools/config_bench.py ${only:+--only $only} 2> ${O}_configs.err | tee ${O}_configs.jsonl ;;
    anim) timeout 600 python tools/anim_bench.py 2> ${O}_anim.err | tee ${O}_anim.jsonl ;;
    debugchecks)   # the GPU test-suite against a build with device-side assertions (-DPLTB_DEBUG_CHECKS); stands in for compute-sanitizer
      [ -f mitsuba3_plt_b200/libpltb200_dbg.so ] || bash tools/build_variant.sh dbg -DPLTB_DEBUG_CHECKS > ${O}_debugchecks_build.log 2>&1   # (build it before the call: box time is GPU budget)
      PLTB200_LIB=mitsuba3_plt_b200/libpltb200_dbg.so timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -5 | tee ${O}_debugchecks.log ;;
    sanitize)
      for tool in memcheck racecheck initcheck; do
        timeout 1200 compute-sanitizer --tool $tool python -c "import __graft_entry__ as g; g.smoke()" > ${O}_sanitize_$tool.log 2>&1
        tail -n 4 ${O}_sanitize_$tool.log
      done ;;
    *) echo "unknown step $step" ;;
  esac
done
du -sh gpurun_out
