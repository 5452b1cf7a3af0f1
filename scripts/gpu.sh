#!/bin/bash
# One parameterised runner for everything that needs the GPU box (replaces round 1's one-off gpu_*.sh scripts).
#   gpurun --timeout T -- 'bash scripts/gpu.sh TASK [ARGS...] ; bash scripts/gpu.sh TASK2 ...'
# Every task writes under gpurun_out/<tag>_*; summaries worth keeping are copied into profiles/ by hand afterwards.
#
#   tests [pytest args]          python -m pytest tests -m gpu -x -q
#   bench TAG [bench args]       one bench.py line                        -> gpurun_out/TAG_bench.json
#   variants TAG "BENCH ARGS" V1 V2 ...   the same bench line for build/libminipath_b200_V.so (scripts/build_variant.sh), table at the end
#   launches TAG [bench args]    ncu launch list (gpu__time_duration) of one step     -> gpurun_out/TAG_launches.csv
#   ncu TAG REGEX SKIP COUNT [bench args]   ncu --set full of COUNT launches matching REGEX after SKIP -> gpurun_out/TAG.ncu-rep + TAG_summary.txt
#   traffic TAG KEY SPP [bench args]        DRAM bytes of the traversal launches of one step -> profiles/traffic.json (scripts/ncu_traffic.py)
#   gates [args]                 python scripts/gate_full.py (full-size parity gates)  -> gpurun_out/gates.txt
#   scale TAG N [bench args]     torchrun bench.py on N GPUs (needs gpurun --gpus N)   -> gpurun_out/TAG_scale_nN.json
set -u
mkdir -p gpurun_out
task=$1; shift
case $task in
  tests)
    # paths / options given: only those (e.g. tests/test_gpu_builder.py -s); none: the whole GPU suite
    if [ $# -gt 0 ]; then timeout 1500 python -m pytest -m gpu -x -q "$@" 2>&1 | tail -40 | tee gpurun_out/tests.txt
    else timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/tests.txt; fi ;;
  bench)
    tag=$1; shift
    timeout 900 python bench.py "$@" > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || tail -5 gpurun_out/${tag}_bench.err
    python - "$tag" <<'PY'
import json,sys
try:
    d=json.loads(open(f"gpurun_out/{sys.argv[1]}_bench.json").read().strip().splitlines()[-1])
    r=d.get("roofline") or {}
    print(sys.argv[1], d["config"]["workload"], "value", round(d["value"]), "ms", round(d["ms_per_step"],2), "e2e", d.get("e2e") and round(d["e2e"]["value"]), "bound", r.get("bound"), "frac", r.get("frac") and round(r["frac"],3),
          "l2", round(r["l2"]["frac"],3), "fp32", round(r["fp32"]["frac"],3), "l1g", round(r["l1_gather"]["frac"],3))
except Exception as e: print(sys.argv[1], "ERR", e)
PY
    ;;
  variants)
    tag=$1; bargs=$2; shift 2
    for v in "$@"; do
      if [ "$v" = default ]; then unset MINIPATH_B200_LIB; else export MINIPATH_B200_LIB=$PWD/build/libminipath_b200_$v.so; fi
      timeout 400 python bench.py --no-cpu-baseline --no-e2e --no-extra $bargs > gpurun_out/${tag}_${v}.json 2> gpurun_out/${tag}_${v}.err
    done
    unset MINIPATH_B200_LIB
    python - "$tag" "$@" <<'PY' | tee gpurun_out/${tag}_table.txt
import json,sys
tag=sys.argv[1]
for v in sys.argv[2:]:
    f=f"gpurun_out/{tag}_{v}.json"
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f"{v:28s} {d['value']:9.1f} Mrays/s {d['ms_per_step']:9.3f} ms/step  frac {d['roofline']['frac']:.3f}")
    except Exception as e: print(v,"ERR",open(f.replace('.json','.err')).read()[-400:])
PY
    ;;
  launches)
    tag=$1; shift
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv \
        python bench.py --no-cpu-baseline --no-e2e --no-extra --steps 1 --warmup 0 "$@" > gpurun_out/${tag}_launches.log 2>&1
    python scripts/ncu_shares.py gpurun_out/${tag}_launches.csv | tee gpurun_out/${tag}_shares.txt ;;
  ncu)
    tag=$1; regex=$2; skip=$3; count=$4; shift 4
    timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base ${NCU_BASE:-function} -k regex:"$regex" -s $skip -c $count -f -o gpurun_out/$tag \
        python bench.py --no-cpu-baseline --no-e2e --no-extra --steps 1 --warmup 0 "$@" > gpurun_out/${tag}_ncu.log 2>&1
    python scripts/ncu_summary.py gpurun_out/$tag.ncu-rep | tee gpurun_out/${tag}_summary.txt ;;
  traffic)
    tag=$1; key=$2; spp=$3; shift 3
    timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_traffic.csv \
        python bench.py --no-cpu-baseline --no-e2e --no-extra --steps 1 --warmup 0 --spp $spp "$@" > gpurun_out/${tag}_traffic.log 2>&1
    python scripts/ncu_traffic.py gpurun_out/${tag}_traffic.csv "$key" $spp
    cp profiles/traffic.json gpurun_out/traffic.json ;;
  gates)
    timeout 3000 python scripts/gate_full.py "$@" 2>&1 | tee gpurun_out/gates.txt ;;
  scale)
    tag=$1; n=$2; shift 2
    if [ "$n" = 1 ]; then
      timeout 900 python bench.py --gpus 1 "$@" > gpurun_out/${tag}_scale_n1.json 2> gpurun_out/${tag}_scale_n1.err
    else
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n "$@" \
        > gpurun_out/${tag}_scale_n$n.json 2> gpurun_out/${tag}_scale_n$n.err
    fi
    tail -c 600 gpurun_out/${tag}_scale_n$n.err; python - gpurun_out/${tag}_scale_n$n.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][-1])
    print(sys.argv[1], "n", d["n_gpus"], d["scaling"], "value", round(d["value"]), "ms", round(d["ms_per_step"],2), "e2e", d.get("e2e") and round(d["e2e"]["value"]))
except Exception as e: print(sys.argv[1], "ERR", e)
PY
    ;;
  *) echo "unknown task $task"; exit 2 ;;
esac
